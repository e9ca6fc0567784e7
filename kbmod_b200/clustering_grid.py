"""Mirror of the reference's ``kbmod.filters.clustering_grid`` (filters/clustering_grid.py) -- the near-duplicate
filter the pipeline applies to the search results (run_search.py:293-301) -- evaluated on the device through
``kb_grid_filter`` (csrc/results.cu: stable radix sorts by lh and by the four bin coordinates).  No CPU path."""
import ctypes as C

import numpy as np

from ._lib import check, lib, ptr, require_gpu
from .search import array_to_trajectories, trajectories_to_array


def _grid_indices(arr, bin_width, max_time):
    require_gpu()   # after the callers' argument checks
    idx = np.zeros(max(arr.size, 1), dtype=np.uint32)
    n = C.c_uint64(0)
    check(lib.kb_grid_filter(ptr(arr), C.c_uint64(arr.size), C.c_double(bin_width), C.c_double(max_time),
                             ptr(idx), C.byref(n)))
    return idx[: n.value].astype(np.int64)


class TrajectoryClusterGrid:
    """clustering_grid.py:13-150.  Trajectories are collected on the host; the table (best trajectory per spatial bin,
    bins in first-seen order) is computed on the device when it is read."""

    def __init__(self, bin_width=10, max_time=1.0):
        if bin_width < 1 or not np.isfinite(bin_width):
            raise ValueError(f"Bin width must be at least 1. Got {bin_width}.")
        self.bin_width = bin_width
        if max_time < 0 or not np.isfinite(max_time):
            raise ValueError(f"Max time must be >= 0. Got {max_time}.")
        self.max_time = max_time
        self.total_count = 0
        self._trjs = []
        self._idx = []
        self._cache = None

    def _solve(self):
        if self._cache is None:
            self._cache = _grid_indices(trajectories_to_array(self._trjs), self.bin_width, self.max_time)
        return self._cache

    def __len__(self):
        return len(self._solve())

    def _keys(self):
        """Bin key of every collected trajectory (host arithmetic identical to the device's: double products in
        add_trajectory, clustering_grid.py:72-75, and -- the velocity columns being Python floats -- in
        add_trajectory_list, :109-113)."""
        keys = []
        for trj in self._trjs:
            ex, ey = self.max_time * trj.vx, self.max_time * trj.vy
            keys.append((int(trj.x / self.bin_width), int(trj.y / self.bin_width),
                         int((trj.x + ex) / self.bin_width), int((trj.y + ey) / self.bin_width)))
        return keys

    @property
    def table(self):
        """{bin key: best trajectory}, bins in first-seen order (the reference's ``table`` dict)."""
        keys = self._keys()
        return {keys[i]: self._trjs[i] for i in self._solve()}

    @property
    def idx_table(self):
        keys = self._keys()
        return {keys[i]: self._idx[i] for i in self._solve()}

    @property
    def count(self):
        """{bin key: number of trajectories that fell into the bin}."""
        out = {}
        for k in self._keys():
            out[k] = out.get(k, 0) + 1
        return out

    def add_trajectory(self, trj, idx=None):
        if idx is None:
            idx = self.total_count
        self._trjs.append(trj)
        self._idx.append(idx)
        self.total_count += 1
        self._cache = None

    def add_trajectory_list(self, trj_list):
        for idx, trj in enumerate(trj_list):
            self._trjs.append(trj)
            self._idx.append(idx)
        self.total_count += len(trj_list)
        self._cache = None

    def get_trajectories(self):
        return [self._trjs[i] for i in self._solve()]

    def get_indices(self):
        return [self._idx[i] for i in self._solve()]


def apply_trajectory_grid_filter(trajectories, bin_width, max_dt):
    """clustering_grid.py:153-176: (best trajectory of every bin, their indices in ``trajectories``)."""
    if bin_width < 1 or not np.isfinite(bin_width):
        raise ValueError(f"Bin width must be at least 1. Got {bin_width}.")
    if max_dt < 0 or not np.isfinite(max_dt):
        raise ValueError(f"Max time must be >= 0. Got {max_dt}.")
    if isinstance(trajectories, np.ndarray):
        keep = _grid_indices(np.ascontiguousarray(trajectories), bin_width, max_dt)
        return trajectories[keep], [int(i) for i in keep]
    trajectories = list(trajectories)
    keep = _grid_indices(trajectories_to_array(trajectories), bin_width, max_dt)
    return [trajectories[i] for i in keep], [int(i) for i in keep]
