"""Mirror of the reference's ``kbmod.filters.sigma_g_filter`` (filters/sigma_g_filter.py) -- the step that follows
``StackSearch.search_all`` in its pipeline (run_search.py:311-321) -- on the device through the C-ABI
(``kb_sigmag_clip_matrix`` / ``kb_update_likelihood`` / ``kb_sigmag_filter_curves``, csrc/curve_filters.cu).

Same class, constructor arguments, validation and return types as the reference; there is no CPU path.
"""
import ctypes as C
import math

import numpy as np

from ._lib import check, f32, lib, ptr, require_gpu


def _inv_gauss_cdf(z):
    """sigma_g_filter.py:76-83: sqrt(2) * erfinv(2z - 1); without scipy, bisection on math.erf (equal to ~1 ulp)."""
    try:  # the reference's own expression, when scipy is installed (identical coefficient to the last bit)
        from scipy.special import erfinv
        sign = -1 if z < 0.5 else 1
        return sign * np.sqrt(2) * erfinv(sign * (2 * z - 1))
    except ImportError:
        pass
    if z <= 0.0 or z >= 1.0:
        return math.copysign(math.inf, z - 0.5)
    target = 2.0 * z - 1.0
    lo, hi = -6.0, 6.0
    for _ in range(200):
        mid = 0.5 * (lo + hi)
        if math.erf(mid) < target:
            lo = mid
        else:
            hi = mid
    return math.sqrt(2.0) * 0.5 * (lo + hi)


class SigmaGClipping:
    """filters/sigma_g_filter.py:19-168."""

    def __init__(self, low_bnd=25, high_bnd=75, n_sigma=2, clip_negative=False):
        if low_bnd > high_bnd or low_bnd <= 0.0 or high_bnd >= 100.0:
            raise ValueError(f"Invalid bounds [{low_bnd}, {high_bnd}]")
        if n_sigma <= 0.0:
            raise ValueError(f"Invalid n_sigma {n_sigma}")
        self.low_bnd = low_bnd
        self.high_bnd = high_bnd
        self.n_sigma = n_sigma
        self.coeff = SigmaGClipping.find_sigma_g_coeff(low_bnd, high_bnd)
        self.clip_negative = clip_negative

    @staticmethod
    def find_sigma_g_coeff(low_bnd, high_bnd):
        if (high_bnd <= low_bnd) or (low_bnd < 0) or (high_bnd > 100):
            raise ValueError(f"Invalid percentiles for sigma G coefficient [{low_bnd}, {high_bnd}]")
        x1 = SigmaGClipping.invert_gauss_cdf(low_bnd / 100.0)
        x2 = SigmaGClipping.invert_gauss_cdf(high_bnd / 100.0)
        return 1 / (x2 - x1)

    @staticmethod
    def invert_gauss_cdf(z):
        return float(_inv_gauss_cdf(z))

    def _args(self):
        # torch narrows the python double n_sigma * coeff when it meets the float32 tensor (sigma_g_filter.py:157)
        return (C.c_float(np.float32(self.low_bnd / 100.0)), C.c_float(np.float32(self.high_bnd / 100.0)),
                C.c_float(np.float32(self.n_sigma * self.coeff)), C.c_int(1 if self.clip_negative else 0))

    def compute_clipped_sigma_g_matrix(self, lh):
        """N x T likelihood curves -> N x T booleans (sigma_g_filter.py:114-168), computed on the device."""
        require_gpu()
        lh = f32(lh)
        if lh.ndim != 2:
            raise ValueError("lh must be an N x T matrix")
        out = np.zeros(lh.shape, dtype=np.uint8)
        if lh.size:
            check(lib.kb_sigmag_clip_matrix(ptr(lh), C.c_uint64(lh.shape[0]), C.c_uint32(lh.shape[1]), *self._args(),
                                            ptr(out)))
        return out.astype(bool)

    def compute_clipped_sigma_g(self, lh):
        """Single likelihood curve -> indices that pass (sigma_g_filter.py:86-112): numpy's float64 percentiles and a
        1e-8 floor -- different arithmetic from the matrix form -- evaluated on the device (kb_sigmag_clip_curve)."""
        require_gpu()
        lh = np.ascontiguousarray(lh, dtype=np.float64)
        if lh.ndim != 1:
            raise ValueError("lh must be a single likelihood curve")
        if lh.size == 0 or (self.clip_negative and np.count_nonzero(lh > 0) == 0):
            return np.array([]) if lh.size and self.clip_negative else np.array([], dtype=np.int64)
        idx = np.zeros(lh.size, dtype=np.int64)
        n = C.c_uint32(0)
        check(lib.kb_sigmag_clip_curve(ptr(lh), C.c_uint32(lh.size), C.c_double(self.low_bnd), C.c_double(self.high_bnd),
                                       C.c_double(self.n_sigma), C.c_double(self.coeff),
                                       C.c_int(1 if self.clip_negative else 0), ptr(idx), C.byref(n)))
        return idx[: n.value]


def update_likelihood(psi, phi, obs_valid=None):
    """Results._update_likelihood (results.py:608-645) on the device: (likelihood, flux, obs_count) per row."""
    require_gpu()
    psi, phi = f32(psi), f32(phi)
    if psi.ndim != 2 or psi.shape != phi.shape:
        raise ValueError("psi and phi must be N x T matrices of the same shape")
    R, T = psi.shape
    lh, flux, obs = np.zeros(R), np.zeros(R), np.zeros(R, dtype=np.int64)
    ov = None if obs_valid is None else np.ascontiguousarray(obs_valid, dtype=np.uint8)
    if ov is not None and ov.shape != psi.shape:
        raise ValueError("Wrong shape for obs_valid")
    if R:
        check(lib.kb_update_likelihood(ptr(psi), ptr(phi), ptr(ov) if ov is not None else None, C.c_uint64(R),
                                       C.c_uint32(T), ptr(lh), ptr(flux), ptr(obs)))
    return lh, flux, obs


def apply_clipped_sigma_g_arrays(clipper, psi, phi, obs_valid=None):
    """apply_clipped_sigma_g (sigma_g_filter.py:171-190) on plain arrays, fused on the device: likelihood curves
    (results.py:568-606) -> clip -> new obs_valid -> likelihood update.  Returns (obs_valid, likelihood, flux,
    obs_count)."""
    require_gpu()
    psi, phi = f32(psi), f32(phi)
    if psi.ndim != 2 or psi.shape != phi.shape:
        raise ValueError("psi and phi must be N x T matrices of the same shape")
    R, T = psi.shape
    has = obs_valid is not None
    ov = np.ascontiguousarray(obs_valid, dtype=np.uint8).copy() if has else np.zeros((R, T), dtype=np.uint8)
    lh, flux, obs = np.zeros(R), np.zeros(R), np.zeros(R, dtype=np.int64)
    if R:
        check(lib.kb_sigmag_filter_curves(ptr(psi), ptr(phi), C.c_uint64(R), C.c_uint32(T), *clipper._args(), ptr(ov),
                                          C.c_int(1 if has else 0), ptr(lh), ptr(flux), ptr(obs)))
    return ov.astype(bool), lh, flux, obs


def apply_clipped_sigma_g(clipper, result_data):
    """sigma_g_filter.py:171-190 for a reference ``Results``-like object (needs ``compute_likelihood_curves`` and
    ``update_obs_valid``)."""
    if len(result_data) == 0:
        return
    lh = result_data.compute_likelihood_curves(filter_obs=True, mask_value=np.nan)
    obs_valid = clipper.compute_clipped_sigma_g_matrix(lh)
    result_data.update_obs_valid(obs_valid)
