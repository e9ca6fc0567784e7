"""Synthetic inputs for the shift-and-stack path (host-side, numpy only).

The reference's generators live in ``kbmod.fake_data`` / ``kbmod.trajectory_generator`` /
``kbmod.core.psf`` and need astropy to import, so the few that the hot path's tests and benchmark use
are restated here with the same argument meaning and the same RNG call order (so a given seed gives the
same arrays as the reference):

* ``create_fake_times``            fake_data/fake_data_creator.py:23-59
* ``make_fake_image_stack``        fake_data/fake_data_creator.py:62-107 (sci drawn in t order, var = noise^2)
* ``add_random_masks``             fake_data/fake_data_creator.py:110-132
* ``add_fake_object``              fake_data/fake_data_creator.py:135-172
* ``make_gaussian_kernel``         core/psf.py:48-74
* ``FakeDataSet``                  fake_data/fake_data_creator.py:175-450 (stack + object insertion only)
* ``VelocityGridSearch``           trajectory_generator.py:205-268
* ``KBMODV1Search``                trajectory_generator.py:380-456
* ``EclipticCenteredSearch``       trajectory_generator.py:523-625 (pixel/day + radian form only)
"""
import math

import numpy as np

from .trajectory import Trajectory


def create_fake_times(num_times, t0=0.0, obs_per_day=1, intra_night_gap=0.01, inter_night_gap=1):
    if num_times <= 0:
        raise ValueError(f"Invalid number of times {num_times}")
    out, seen, day = [], 0, 0
    for _ in range(num_times):
        out.append(t0 + day + seen * intra_night_gap)
        seen += 1
        if seen == obs_per_day:
            seen = 0
            day += inter_night_gap
    return out


def make_gaussian_kernel(stddev, normalize=True):
    if stddev < 0:
        raise ValueError("Standard deviation must be non-negative.")
    radius = int(3 * stddev)
    x = np.arange(-radius, radius + 1)
    xx, yy = np.meshgrid(x, x)
    kernel = np.exp(-0.5 * (xx**2 + yy**2) / stddev**2)
    if normalize:
        kernel /= np.sum(kernel)
    return kernel.astype(np.float32)


class ImageStack:
    """Minimal stand-in for ``kbmod.core.image_stack_py.ImageStackPy`` (times, sci, var, psfs)."""

    def __init__(self, times, sci, var, psfs):
        self.times = np.asarray(times, dtype=float)
        self.zeroed_times = self.times - self.times[0]
        self.sci, self.var, self.psfs = sci, var, psfs
        self.num_times = len(self.times)
        self.height, self.width = sci[0].shape


def make_fake_image_stack(height, width, times, noise_level=2.0, psf_val=0.5, psfs=None, rng=None):
    if rng is None:
        rng = np.random.default_rng()
    times = np.asarray(times)
    sci = [rng.normal(0.0, noise_level, (height, width)).astype(np.float32) for _ in range(len(times))]
    var = [np.full((height, width), noise_level**2).astype(np.float32) for _ in range(len(times))]
    if psfs is None:
        k = make_gaussian_kernel(psf_val)
        psfs = [k for _ in range(len(times))]
    elif len(psfs) != len(times):
        raise ValueError(f"The number of PSFs ({len(psfs)}) must be the same as times ({len(times)}).")
    return ImageStack(times, sci, var, psfs)


def add_random_masks(stack, mask_fraction, rng=None):
    if not (0.0 <= mask_fraction <= 1.0):
        raise ValueError(f"Invalid mask fraction {mask_fraction}. Must be between 0 and 1.")
    if rng is None:
        rng = np.random.default_rng()
    for idx in range(len(stack.sci)):
        mask = rng.random(stack.sci[idx].shape) < mask_fraction
        stack.sci[idx][mask] = np.nan
        stack.var[idx][mask] = np.nan


def add_fake_object(stack, x, y, vx, vy, *, flux=100.0):
    for idx, t in enumerate(stack.zeroed_times):
        k = stack.psfs[idx]
        dim = k.shape[0]
        rad = dim // 2
        px = int(x + vx * t + 0.5)
        py = int(y + vy * t + 0.5)
        x0, x1 = max(px - rad, 0), min(px + rad + 1, stack.width)
        y0, y1 = max(py - rad, 0), min(py + rad + 1, stack.height)
        if x0 >= x1 or y0 >= y1:
            continue
        sub = stack.sci[idx][y0:y1, x0:x1]
        ker = k[y0 - (py - rad) : y1 - (py - rad), x0 - (px - rad) : x1 - (px - rad)]
        ok = np.isfinite(sub)
        sub[ok] += (np.float32(flux) * ker)[ok]  # float32 product, as numpy>=2 computes `flux * psf[y, x]`


class FakeDataSet:
    """Seeded synthetic stack with inserted linear movers (reference: fake_data_creator.py:175-450)."""

    def __init__(self, width, height, times, *, mask_fraction=0.0, noise_level=2.0, psf_val=0.5, psfs=None,
                 use_seed=-1):
        if len(times) == 0:
            raise ValueError("The list of times must not be empty.")
        if width <= 0 or height <= 0:
            raise ValueError(f"Invalid image dimensions: width={width}, height={height}")
        self.width, self.height, self.times = width, height, times
        self.rng = np.random.default_rng() if use_seed < 0 else np.random.default_rng(use_seed)
        self.stack_py = make_fake_image_stack(height, width, times, noise_level=noise_level, psf_val=psf_val,
                                              psfs=psfs, rng=self.rng)
        if mask_fraction > 0.0:
            add_random_masks(self.stack_py, mask_fraction, rng=self.rng)
        self.trajectories = []

    def insert_object(self, trj):
        add_fake_object(self.stack_py, trj.x, trj.y, trj.vx, trj.vy, flux=trj.flux)
        self.trajectories.append(trj)

    def trajectory_is_within_bounds(self, trj):
        dt = self.times[-1] - self.times[0]
        xe, ye = trj.x + trj.vx * dt, trj.y + trj.vy * dt
        return 0 <= trj.x < self.width and 0 <= trj.y < self.height and 0 <= xe < self.width and 0 <= ye < self.height

    def insert_random_objects_from_generator(self, num_trj, generator, flux):
        cands = [t for t in generator]
        if not cands:
            raise ValueError("The generator did not produce any trajectories.")
        out, itr = [], 0
        while len(out) < num_trj and itr < 10 * num_trj:
            v = cands[int(self.rng.choice(len(cands)))]
            trj = Trajectory(x=int(self.rng.random() * self.width), y=int(self.rng.random() * self.height),
                             vx=v.vx, vy=v.vy, flux=flux)
            if self.trajectory_is_within_bounds(trj):
                self.insert_object(trj)
                out.append(trj)
            itr += 1
        return out


class _Generator:
    def __iter__(self):
        return self.generate()

    def to_arrays(self):
        """Candidate velocities as two float32 arrays (what StackSearch.search_all consumes)."""
        v = np.array([(t.vx, t.vy) for t in self], dtype=np.float32).reshape(-1, 2)
        return np.ascontiguousarray(v[:, 0]), np.ascontiguousarray(v[:, 1])


class VelocityGridSearch(_Generator):
    def __init__(self, vx_steps, min_vx, max_vx, vy_steps, min_vy, max_vy):
        if vx_steps < 2 or vy_steps < 2:
            raise ValueError("VelocityGridSearch requires at least 2 steps in each dimension")
        if max_vx < min_vx or max_vy < min_vy:
            raise ValueError("Invalid VelocityGridSearch bounds.")
        self.vx_steps, self.min_vx, self.max_vx = vx_steps, min_vx, max_vx
        self.vy_steps, self.min_vy, self.max_vy = vy_steps, min_vy, max_vy
        self.vx_stepsize = (max_vx - min_vx) / float(vx_steps - 1)
        self.vy_stepsize = (max_vy - min_vy) / float(vy_steps - 1)

    def __len__(self):
        return self.vy_steps * self.vx_steps

    def generate(self):
        for vy_i in range(self.vy_steps):
            for vx_i in range(self.vx_steps):
                yield Trajectory(vx=self.min_vx + vx_i * self.vx_stepsize, vy=self.min_vy + vy_i * self.vy_stepsize)


class KBMODV1Search(_Generator):
    def __init__(self, vel_steps, min_vel, max_vel, ang_steps, min_ang, max_ang):
        if vel_steps < 1 or ang_steps < 1:
            raise ValueError("KBMODV1Search requires at least 1 step in each dimension")
        if max_vel < min_vel or max_ang < min_ang:
            raise ValueError("Invalid KBMODV1Search bounds.")
        self.vel_steps, self.min_vel, self.max_vel = vel_steps, min_vel, max_vel
        self.ang_steps, self.min_ang, self.max_ang = ang_steps, min_ang, max_ang
        self.vel_stepsize = (max_vel - min_vel) / float(vel_steps)
        self.ang_stepsize = (max_ang - min_ang) / float(ang_steps)

    def __len__(self):
        return self.ang_steps * self.vel_steps

    def generate(self):
        for ang_i in range(self.ang_steps):
            for vel_i in range(self.vel_steps):
                a = self.min_ang + ang_i * self.ang_stepsize
                v = self.min_vel + vel_i * self.vel_stepsize
                yield Trajectory(vx=math.cos(a) * v, vy=math.sin(a) * v)


class EclipticCenteredSearch(_Generator):
    """velocities=[min,max,steps] in pix/day, angles=[min,max,steps] offsets in radians (inclusive ends)."""

    def __init__(self, velocities, angles, given_ecliptic=0.0):
        if len(angles) != 3 or len(velocities) != 3 or angles[2] < 1 or velocities[2] < 1:
            raise ValueError("Invalid EclipticCenteredSearch parameters")
        self.velocities, self.angles = list(velocities), list(angles)
        self.vel_stepsize = (velocities[1] - velocities[0]) / float(max(velocities[2] - 1, 1))
        self.min_ang = given_ecliptic + angles[0]
        self.max_ang = given_ecliptic + angles[1]
        self.ang_stepsize = (self.max_ang - self.min_ang) / float(max(angles[2] - 1, 1))

    def __len__(self):
        return self.angles[2] * self.velocities[2]

    def generate(self):
        for ang_i in range(self.angles[2]):
            for vel_i in range(self.velocities[2]):
                a = self.min_ang + ang_i * self.ang_stepsize
                v = self.velocities[0] + vel_i * self.vel_stepsize
                yield Trajectory(vx=math.cos(a) * v, vy=math.sin(a) * v)
