"""The ``Trajectory`` record of ``kbmod.search`` (common.h:55-115,164-198) as a plain Python class.

Kept apart from ``search.py`` so host-only code (input generators, the benchmark's reference arm) can use it
without loading the CUDA library.
"""
import math

import numpy as np


def _narrow(v):
    return float(np.float32(v))


class Trajectory:
    """common.h:55-115,164-198.  vx, vy, lh, flux are stored as float32 like the C++ struct."""

    __slots__ = ("_vx", "_vy", "_lh", "_flux", "x", "y", "obs_count")

    def __init__(self, x=0, y=0, vx=0.0, vy=0.0, flux=0.0, lh=0.0, obs_count=0):
        self.x, self.y, self.obs_count = int(x), int(y), int(obs_count)
        self._vx, self._vy, self._flux, self._lh = _narrow(vx), _narrow(vy), _narrow(flux), _narrow(lh)

    vx = property(lambda s: s._vx, lambda s, v: setattr(s, "_vx", _narrow(v)))
    vy = property(lambda s: s._vy, lambda s, v: setattr(s, "_vy", _narrow(v)))
    lh = property(lambda s: s._lh, lambda s, v: setattr(s, "_lh", _narrow(v)))
    flux = property(lambda s: s._flux, lambda s, v: setattr(s, "_flux", _narrow(v)))

    def get_x_pos(self, time, centered=True):
        return _narrow(self.x + time * self._vx + 0.5) if centered else _narrow(self.x + time * self._vx)

    def get_y_pos(self, time, centered=True):
        return _narrow(self.y + time * self._vy + 0.5) if centered else _narrow(self.y + time * self._vy)

    def get_x_index(self, time):
        return int(math.floor(self.get_x_pos(time, True)))

    def get_y_index(self, time):
        return int(math.floor(self.get_y_pos(time, True)))

    def is_valid(self):
        return bool(np.isfinite([self._vx, self._vy, self._lh, self._flux]).all() and self.obs_count >= 0)

    def clear(self):
        self.x = self.y = self.obs_count = 0
        self._vx = self._vy = self._lh = self._flux = 0.0

    def __str__(self):
        return (f"lh: {self._lh:.6f} flux: {self._flux:.6f} x: {self.x} y: {self.y} vx: {self._vx:.6f} "
                f"vy: {self._vy:.6f} obs_count: {self.obs_count}")

    def __repr__(self):
        return f"Trajectory({self})"

    def __eq__(self, o):
        return isinstance(o, Trajectory) and self._astuple() == o._astuple()

    def _astuple(self):
        return (self._vx, self._vy, self._lh, self._flux, self.x, self.y, self.obs_count)

    def __getstate__(self):
        return self._astuple()

    def __setstate__(self, t):
        self._vx, self._vy, self._lh, self._flux, self.x, self.y, self.obs_count = t
