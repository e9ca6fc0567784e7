"""Drop-in mirror of the ``kbmod.search`` module for the shift-and-stack path, running on a B200.

Same names, argument meaning and error behaviour as the reference's pybind11 module
(``src/kbmod/search/bindings.cpp:20-41``); every compute call goes through the C-ABI of
``libkbmod_b200.so`` (``include/kbmod_b200.h``) into hand-written sm_100a CUDA.  There is no CPU fallback:
the reference's CPU-only entry points (``search_all(..., False)``, ``search_cpu_only``,
``evaluate_trajectory_cpu``, ``convolve_image_cpu``) raise ``RuntimeError`` here.

Deviation from the reference, on purpose: no 200-image limit (``common.h:31`` ``MAX_NUM_IMAGES``) -- the
search accepts any number of images (sigma-G on the device up to 1024).
"""
import ctypes as C
import enum
import logging as _pylogging
import math
import time as _time

import numpy as np

from . import _lib
from .trajectory import Trajectory, _narrow
from ._lib import KbMeta, KbParams, KbTiming, KbTrajectory, TRAJECTORY_DTYPE, check, f32, lib, ptr, ptr_array

KB_NO_DATA = float("nan")  # common.h:34
HAS_CUDA = True
HAS_OMP = False
MAX_KERNEL_RADIUS = 15


def pixel_value_valid(value):
    """common.h:41"""
    return bool(np.isfinite(np.float32(value)))


def kb_has_gpu():
    """kernel_helpers.cpp:23-29"""
    return _lib.device_count() > 0


def get_gpu_total_memory():
    free, total = C.c_size_t(0), C.c_size_t(0)
    check(lib.kb_gpu_memory(C.byref(free), C.byref(total)))
    return total.value


def get_gpu_free_memory():
    free, total = C.c_size_t(0), C.c_size_t(0)
    check(lib.kb_gpu_memory(C.byref(free), C.byref(total)))
    return free.value


def stat_gpu_memory_mb():
    return "GPU: %f MB free of %f MB total." % (get_gpu_free_memory() / 1048576.0, get_gpu_total_memory() / 1048576.0)


def print_cuda_stats():
    print("\n----- CUDA Debugging Log -----")
    print(f"Devices: {_lib.device_count()}  {stat_gpu_memory_mb() if kb_has_gpu() else 'no device'}")


def validate_gpu(req_memory=0):
    return kb_has_gpu() and get_gpu_free_memory() >= req_memory


class StampType(enum.IntEnum):
    STAMP_SUM = 0
    STAMP_MEAN = 1
    STAMP_MEDIAN = 2
    STAMP_VAR_WEIGHTED = 3


def trajectories_to_array(trjs):
    """list[Trajectory] (or structured array) -> numpy array of kb_trajectory records."""
    if isinstance(trjs, np.ndarray) and trjs.dtype == TRAJECTORY_DTYPE:
        return np.ascontiguousarray(trjs)
    out = np.zeros(len(trjs), dtype=TRAJECTORY_DTYPE)
    for i, t in enumerate(trjs):
        out[i] = (t.vx, t.vy, t.lh, t.flux, t.x, t.y, t.obs_count)
    return out


def array_to_trajectories(arr):
    return [Trajectory(x=int(r["x"]), y=int(r["y"]), vx=float(r["vx"]), vy=float(r["vy"]), flux=float(r["flux"]),
                       lh=float(r["lh"]), obs_count=int(r["obs_count"])) for r in arr]


class SearchParameters:
    """common.h:119-161,200-217"""

    def __init__(self):
        self.min_observations = 0
        self.min_lh = 0.0
        self.do_sigmag_filter = False
        self.sgl_L = 0.25
        self.sgl_H = 0.75
        self.sigmag_coeff = -1.0
        self.encode_num_bytes = -1
        self.x_start_min = 0
        self.x_start_max = 0
        self.y_start_min = 0
        self.y_start_max = 0
        self.results_per_pixel = 8
        self.total_results = 0

    def _c(self):
        return KbParams(int(self.min_observations), float(self.min_lh), int(bool(self.do_sigmag_filter)),
                        float(self.sgl_L), float(self.sgl_H), float(self.sigmag_coeff), int(self.encode_num_bytes),
                        int(self.x_start_min), int(self.x_start_max), int(self.y_start_min), int(self.y_start_max),
                        int(self.results_per_pixel), int(self.total_results))

    def __str__(self):
        out = f"Filtering Settings:\n  min_observations: {self.min_observations}\n  min_lh: {self.min_lh:.6f}"
        if self.do_sigmag_filter:
            out += f"\n  SigmaG: [{self.sgl_L:.6f}, {self.sgl_H:.6f}] coeff={self.sigmag_coeff:.6f}"
        else:
            out += "\n  SigmaG: OFF"
        out += f"\nResults per pixel: {self.results_per_pixel}\nencode_num_bytes: {self.encode_num_bytes}"
        out += f"\nBounds X=[{self.x_start_min}, {self.x_start_max}] Y=[{self.y_start_min}, {self.y_start_max}]"
        return out


class PsiPhi:
    """psi_phi_array_ds.h:34-37"""

    def __init__(self, psi=0.0, phi=0.0):
        self.psi, self.phi = _narrow(psi), _narrow(phi)

    def __repr__(self):
        return f"PsiPhi(psi={self.psi}, phi={self.phi})"


def encode_uint_scalar(value, min_val, max_val, scale):
    """psi_phi_array_ds.h:40-43 (host scalar helper; float32 arithmetic, '+ 1.0' in double)."""
    v, mn, mx, sc = np.float32(value), np.float32(min_val), np.float32(max_val), np.float32(scale)
    if not np.isfinite(v):
        return 0.0
    c = v if v < mx else mx
    c = mn if c < mn else c
    return float(np.float32(np.float64(np.float32(np.float32(c - mn) / sc)) + 1.0))


def decode_uint_scalar(value, min_val, scale):
    """psi_phi_array_ds.h:45-47"""
    v = np.float32(value)
    if v == 0.0:
        return KB_NO_DATA
    return float(np.float32((np.float64(v) - 1.0) * np.float64(np.float32(scale)) + np.float64(np.float32(min_val))))


def _as_image_list(imgs, name):
    out = []
    for im in imgs:
        a = np.ascontiguousarray(im, dtype=np.float32)
        if a.ndim != 2:
            raise RuntimeError(f"{name} images must be 2-dimensional float32 arrays.")
        out.append(a)
    return out


class PsiPhiArray:
    """psi_phi_array_ds.h:71-143 / psi_phi_array.cpp.  Owns a device-side handle once filled."""

    def __init__(self):
        self._h = None
        self._meta = KbMeta()
        self._meta.num_bytes = 4
        self._meta.psi_min_val = self._meta.phi_min_val = float(np.finfo(np.float32).max)
        self._meta.psi_max_val = self._meta.phi_max_val = -float(np.finfo(np.float32).max)
        self._meta.psi_scale = self._meta.phi_scale = 1.0
        self._times = []

    def __del__(self):
        self.clear()

    # -- filling ----------------------------------------------------------------------------------
    def _adopt(self, handle):
        self._h = handle
        check(lib.kb_get_meta(self._h, C.byref(self._meta)))

    def _require(self):
        if self._h is None:
            raise RuntimeError("PsiPhiArray is not allocated.")
        return self._h

    def set_meta_data(self, num_bytes, num_times, height, width):
        if num_bytes not in (-1, 1, 2, 4):
            raise RuntimeError(f"Invalid setting of num_bytes. Must be (-1 [use default], 1, 2, or 4). Got {num_bytes}")
        if num_times == 0:
            raise RuntimeError(f"Invalid num_times passed to set_meta_data: {num_times}")
        if width == 0:
            raise RuntimeError(f"Invalid width passed to set_meta_data: {width}")
        if height == 0:
            raise RuntimeError(f"Invalid height passed to set_meta_data: {height}")
        if self._h is not None:
            raise RuntimeError("Cannot change meta data with allocated arrays. Call clear() first.")
        m = self._meta
        m.num_bytes = num_bytes if num_bytes in (1, 2) else 4
        m.block_size = m.num_bytes
        m.num_times, m.width, m.height = num_times, width, height
        m.pixels_per_image = width * height
        m.num_entries = 2 * m.pixels_per_image * num_times
        m.total_array_size = m.block_size * m.num_entries

    def set_time_array(self, times):
        self._times = [float(t) for t in times]

    def clear(self):
        if getattr(self, "_h", None) is not None:
            lib.kb_destroy(self._h)
            self._h = None
        m = self._meta
        m.num_times = m.width = m.height = m.pixels_per_image = m.num_entries = m.total_array_size = 0
        m.psi_min_val = m.phi_min_val = float(np.finfo(np.float32).max)
        m.psi_max_val = m.phi_max_val = -float(np.finfo(np.float32).max)
        m.psi_scale = m.phi_scale = 1.0
        self._times = []

    def move_to_gpu(self):
        check(lib.kb_preload(self._require()))

    def clear_from_gpu(self):
        if self._h is not None:
            check(lib.kb_unload(self._h))

    # -- properties -------------------------------------------------------------------------------
    @property
    def on_gpu(self):
        if self._h is None:
            return False
        f = C.c_int(0)
        check(lib.kb_on_gpu(self._h, C.byref(f)))
        return bool(f.value)

    num_bytes = property(lambda s: s._meta.num_bytes)
    num_times = property(lambda s: s._meta.num_times)
    width = property(lambda s: s._meta.width)
    height = property(lambda s: s._meta.height)
    pixels_per_image = property(lambda s: s._meta.pixels_per_image)
    num_entries = property(lambda s: s._meta.num_entries)
    total_array_size = property(lambda s: s._meta.total_array_size)
    block_size = property(lambda s: s._meta.block_size)
    psi_min_val = property(lambda s: s._meta.psi_min_val)
    psi_max_val = property(lambda s: s._meta.psi_max_val)
    psi_scale = property(lambda s: s._meta.psi_scale)
    phi_min_val = property(lambda s: s._meta.phi_min_val)
    phi_max_val = property(lambda s: s._meta.phi_max_val)
    phi_scale = property(lambda s: s._meta.phi_scale)
    cpu_array_allocated = property(lambda s: s._h is not None)
    gpu_array_allocated = property(lambda s: s.on_gpu)

    def read_psi_phi(self, time, row, col):
        if self._h is None:
            return PsiPhi(KB_NO_DATA, KB_NO_DATA)
        out = (C.c_float * 2)()
        check(lib.kb_read_psi_phi(self._h, C.c_uint64(time), C.c_int(row), C.c_int(col), out))
        return PsiPhi(out[0], out[1])

    def read_time(self, time_index):
        if self._h is None:
            if time_index >= len(self._times):
                raise RuntimeError(f"Out of bounds read for time step. [{time_index}]")
            return self._times[time_index]
        out = C.c_double(0)
        check(lib.kb_read_time(self._h, C.c_uint64(time_index), C.byref(out)))
        return out.value

    def host_array(self):
        """The host copy in the reference layout [t][row][col][psi,phi] as a numpy array (a copy)."""
        p, n = C.c_void_p(), C.c_uint64(0)
        check(lib.kb_host_array(self._require(), C.byref(p), C.byref(n)))
        dt = {4: np.float32, 2: np.uint16, 1: np.uint8}[self.num_bytes]
        buf = (C.c_uint8 * n.value).from_address(p.value)
        return np.frombuffer(buf, dtype=dt).reshape(self.num_times, self.height, self.width, 2).copy()


def compute_scale_params_from_image_vect(imgs, num_bytes):
    """psi_phi_array.cpp:219-245 -- evaluated by the device reduction that the encoder uses."""
    imgs = _as_image_list(imgs, "input")
    zeros = [np.zeros_like(imgs[0]) for _ in imgs]
    tmp = PsiPhiArray()
    fill_psi_phi_array(tmp, num_bytes if num_bytes in (1, 2) else 2, imgs, zeros, [0.0] * len(imgs))
    mn, mx = tmp.psi_min_val, tmp.psi_max_val
    scale = tmp.psi_scale if num_bytes in (1, 2) else 1.0
    tmp.clear()
    return [mn, mx, scale]


def fill_psi_phi_array(result_data, num_bytes, psi_imgs, phi_imgs, zeroed_times):
    """psi_phi_array.cpp:321-372"""
    if result_data.cpu_array_allocated:
        return
    _lib.require_gpu()
    psi, phi = _as_image_list(psi_imgs, "psi"), _as_image_list(phi_imgs, "phi")
    if len(psi) == 0:
        raise RuntimeError("Trying to fill PsiPhi from empty vectors.")
    if len(phi) != len(psi):
        raise RuntimeError(f"Size mismatch error [psi and phi arrays]. Expected {len(psi)}. Found {len(phi)}")
    H, W = phi[0].shape
    for a in psi + phi:
        if a.shape != (H, W):
            raise RuntimeError("All psi and phi images must have the same dimensions.")
    times = np.ascontiguousarray(zeroed_times, dtype=np.float64)
    if times.size != len(psi):
        raise RuntimeError(f"Size mismatch error [psi array and zeroed times]. Expected {len(psi)}. Found {times.size}")
    h = C.c_void_p()
    check(lib.kb_create_from_psi_phi(ptr_array(psi), ptr_array(phi), C.c_int(len(psi)), C.c_int(H), C.c_int(W),
                                     ptr(times), C.c_int(num_bytes), C.byref(h)))
    result_data._adopt(h)
    result_data._times = [float(t) for t in times]


def fill_psi_phi_array_from_image_arrays(result_data, num_bytes, sci_imgs, var_imgs, psf_kernels, zeroed_times,
                                         rows=None):
    """psi_phi_array.cpp:374-410.  ``rows=(lo, hi)`` (extension, float32 stacks): build psi/phi for those image rows
    only -- what one rank of the multi-GPU search needs (its band of start rows plus the candidates' reach)."""
    _lib.require_gpu()
    sci, var = _as_image_list(sci_imgs, "science"), _as_image_list(var_imgs, "variance")
    psfs = _as_image_list(psf_kernels, "PSF")
    if len(sci) == 0:
        raise RuntimeError("Trying to fill PsiPhi from empty vectors.")
    if len(sci) != len(var):
        raise RuntimeError(f"Number of images in sci and var do not match. Sci={len(sci)}, Var={len(var)}")
    H, W = sci[0].shape
    for s, v in zip(sci, var):
        if s.shape != (H, W) or v.shape != (H, W):
            raise RuntimeError("Science and Variance images must be the same dimensions.")
    dims = (C.c_int * len(psfs))(*[k.shape[0] for k in psfs])
    for k in psfs:
        if k.shape[0] != k.shape[1]:
            raise RuntimeError("PSF kernels must be square.")
    times = np.ascontiguousarray(zeroed_times, dtype=np.float64)
    h = C.c_void_p()
    if rows is None:
        check(lib.kb_create_from_images(ptr_array(sci), ptr_array(var), ptr_array(psfs), dims, C.c_int(len(sci)),
                                        C.c_int(H), C.c_int(W), ptr(times), C.c_int(num_bytes), C.byref(h)))
    else:
        check(lib.kb_create_from_images_rows(ptr_array(sci), ptr_array(var), ptr_array(psfs), dims, C.c_int(len(sci)),
                                             C.c_int(H), C.c_int(W), ptr(times), C.c_int(num_bytes),
                                             C.c_int(int(rows[0])), C.c_int(int(rows[1])), C.byref(h)))
    result_data._adopt(h)
    result_data._times = [float(t) for t in times]


# ---- image utilities (image_utils_cpp.cpp:180-194) ------------------------------------------------------
def _need_f32_2d(a, name):
    if not isinstance(a, np.ndarray) or a.dtype != np.float32 or a.ndim != 2:
        raise TypeError(f"{name} must be a 2-d float32 numpy array (noconvert).")
    return np.ascontiguousarray(a)


def convolve_image_gpu(image, psf):
    _lib.require_gpu()
    image, psf = _need_f32_2d(image, "image"), _need_f32_2d(psf, "psf")
    out = np.empty_like(image)
    check(lib.kb_convolve(ptr(image), C.c_int(image.shape[0]), C.c_int(image.shape[1]), ptr(psf),
                          C.c_int(psf.shape[0]), ptr(out)))
    return out


def convolve_image(image, psf):
    """image_utils_cpp.cpp:103-108: GPU when one is present -- which this package requires."""
    return convolve_image_gpu(image, psf)


def convolve_image_cpu(image, psf):
    raise RuntimeError("kbmod_b200 provides no CPU implementation; use convolve_image / convolve_image_gpu.")


def square_psf_values(given_psf):
    k = _need_f32_2d(given_psf, "given_psf")
    return k * k


def generate_psi(sci, var, psf):
    _lib.require_gpu()
    sci, var, psf = _need_f32_2d(sci, "sci"), _need_f32_2d(var, "var"), _need_f32_2d(psf, "psf")
    if sci.shape != var.shape:
        raise RuntimeError(f"Science and Variance images must be the same dimensions.  Sci = ({sci.shape[0]},"
                           f"{sci.shape[1]}), Var = ({var.shape[0]},{var.shape[1]}).")
    out = np.empty_like(sci)
    check(lib.kb_generate_psi_phi(ptr(sci), ptr(var), C.c_int(sci.shape[0]), C.c_int(sci.shape[1]), ptr(psf),
                                  C.c_int(psf.shape[0]), ptr(out), None))
    return out


def generate_phi(var, psf):
    _lib.require_gpu()
    var, psf = _need_f32_2d(var, "var"), _need_f32_2d(psf, "psf")
    out = np.empty_like(var)
    check(lib.kb_generate_psi_phi(None, ptr(var), C.c_int(var.shape[0]), C.c_int(var.shape[1]), ptr(psf),
                                  C.c_int(psf.shape[0]), None, ptr(out)))
    return out


def sigmag_filtered_indices(values, sgl0, sgl1, sigma_g_coeff, width):
    """kernel_helpers.cpp:86-106, evaluated on the device by the routine the search kernel uses."""
    _lib.require_gpu()
    v = f32(values)
    if v.size == 0:
        return []
    idx = np.zeros(v.size, dtype=np.int32)
    n = C.c_int(0)
    check(lib.kb_sigmag_filtered_indices(ptr(v), C.c_int(v.size), C.c_float(sgl0), C.c_float(sgl1),
                                         C.c_float(sigma_g_coeff), C.c_float(width), ptr(idx), C.byref(n)))
    return [int(i) for i in idx[: n.value]]


def evaluate_trajectory_cpu(psi_phi, candidate):
    raise RuntimeError("kbmod_b200 provides no CPU search; use StackSearch.evaluate_single_trajectory(trj, True).")


def search_cpu_only(psi_phi_array, params, trj_to_search, results):
    raise RuntimeError("kbmod_b200 provides no CPU search; use StackSearch.search_all(candidates, True).")


# ---- TrajectoryList (trajectory_list.{h,cpp}) -----------------------------------------------------------
class TrajectoryList:
    def __init__(self, arg=0):
        if isinstance(arg, (int, np.integer)):
            if arg < 0:
                raise RuntimeError("Invalid TrajectoryList size.")
            self._list = [Trajectory() for _ in range(int(arg))]
        else:
            self._list = list(arg)
            self.assert_valid()
        self._on_gpu = False

    on_gpu = property(lambda s: s._on_gpu)

    def __len__(self):
        return len(self._list)

    def get_size(self):
        return len(self._list)

    def estimate_memory(self):
        return len(self._list) * 28

    def resize(self, new_size):
        self._guard()
        cur = len(self._list)
        if new_size < cur:
            del self._list[new_size:]
        else:
            self._list.extend(Trajectory() for _ in range(new_size - cur))

    def reset_all(self):
        self._guard()
        for t in self._list:
            t.clear()

    def _guard(self):
        if self._on_gpu:
            raise RuntimeError("Data on GPU")

    def get_trajectory(self, index):
        self._guard()
        if index < 0 or index >= len(self._list):
            raise RuntimeError("Index out of bounds.")
        return self._list[index]

    def set_trajectory(self, index, new_value):
        self._guard()
        if index < 0 or index >= len(self._list):
            raise RuntimeError("Index out of bounds.")
        if not new_value.is_valid():
            raise RuntimeError("Invalid trajectory")
        self._list[index] = new_value

    def set_trajectories(self, new_values):
        self._guard()
        self._list = list(new_values)
        self.assert_valid()

    def get_list(self):
        self._guard()
        return self._list

    def get_batch(self, start, count):
        self._guard()
        if count == 0:
            raise RuntimeError("count must be greater than 0")
        return self._list[start : start + count]

    def sort_by_likelihood(self):
        self._guard()
        self._list.sort(key=lambda t: -t.lh)

    def filter_by_likelihood(self, min_lh):
        self._guard()
        m = _narrow(min_lh)
        self._list = [t for t in self._list if not t.lh < m]

    def filter_by_obs_count(self, min_obs_count):
        self._guard()
        self._list = [t for t in self._list if not t.obs_count < min_obs_count]

    def move_to_gpu(self):
        if not kb_has_gpu():
            raise RuntimeError("GPU not available for TrajectoryList")
        self._on_gpu = True

    def move_to_cpu(self):
        self._on_gpu = False
        self.assert_valid()

    def assert_valid(self):
        for i, t in enumerate(self._list):
            if not t.is_valid():
                raise RuntimeError(f"Invalid trajectory detected at index {i}: {t}")


def _extract(trjs, field, dtype):
    if isinstance(trjs, np.ndarray):
        return trjs[field].astype(dtype)
    return np.array([getattr(t, field) for t in trjs], dtype=dtype)


def extract_all_trajectory_x(trjs): return _extract(trjs, "x", np.int32)
def extract_all_trajectory_y(trjs): return _extract(trjs, "y", np.int32)
def extract_all_trajectory_vx(trjs): return _extract(trjs, "vx", np.float32)
def extract_all_trajectory_vy(trjs): return _extract(trjs, "vy", np.float32)
def extract_all_trajectory_lh(trjs): return _extract(trjs, "lh", np.float32)
def extract_all_trajectory_flux(trjs): return _extract(trjs, "flux", np.float32)
def extract_all_trajectory_obs_count(trjs): return _extract(trjs, "obs_count", np.int32)


# ---- StackSearch (stack_search.{h,cpp}) -----------------------------------------------------------------
class StackSearch:
    """``StackSearch(sci_imgs, var_imgs, psf_kernels, zeroed_times, num_bytes=-1)`` -- stack_search.cpp:37-75.
    ``rows=(lo, hi)`` is an extension for the multi-GPU shard (see ``dist.band_rows``)."""

    def __init__(self, sci_imgs, var_imgs, psf_kernels, zeroed_times, num_bytes=-1, rows=None):
        self._log = Logging.getLogger("kbmod.search.run_search")
        sci, var, psfs = list(sci_imgs), list(var_imgs), list(psf_kernels)
        times = [float(t) for t in zeroed_times]
        if len(sci) == 0:
            raise RuntimeError("No images in the to process.")
        if len(sci) != len(var):
            raise RuntimeError(f"The number of science and variance images do not match. Science: {len(sci)}, "
                               f"Variance: {len(var)}")
        if len(sci) != len(psfs):
            raise RuntimeError(f"The number of science and PSF kernel images do not match. Science: {len(sci)}, "
                               f"PSF Kernels: {len(psfs)}")
        if len(sci) != len(times):
            raise RuntimeError(f"The number of science images and zeroed times do not match. Science: {len(sci)}, "
                               f"Zeroed Times: {len(times)}")
        self._num_imgs = len(sci)
        self._height, self._width = np.shape(sci[0])
        self._zeroed_times = times
        self.params = SearchParameters()
        self.set_default_parameters(num_bytes)
        timer = DebugTimer("preparing Psi and Phi images", self._log)
        self._ppa = PsiPhiArray()
        fill_psi_phi_array_from_image_arrays(self._ppa, num_bytes, sci, var, psfs, times, rows)
        self._preloaded = False
        self._result_cache = None
        self.last_search_note = ""   # why the last search was slower than usual, if it was (kb_search_note)
        timer.stop()

    # -- getters ------------------------------------------------------------------------------------
    num_images = property(lambda s: s._num_imgs)
    height = property(lambda s: s._height)
    width = property(lambda s: s._width)
    zeroed_times = property(lambda s: list(s._zeroed_times))
    psi_phi_array = property(lambda s: s._ppa)

    def get_num_images(self): return self._num_imgs
    def get_image_width(self): return self._width
    def get_image_height(self): return self._height

    # -- configuration (stack_search.cpp:89-186) -----------------------------------------------------
    def set_default_parameters(self, num_bytes=-1):
        p = self.params
        p.min_observations, p.min_lh = 0, 0.0
        p.do_sigmag_filter, p.sgl_L, p.sgl_H, p.sigmag_coeff = False, 0.25, 0.75, -1.0
        if num_bytes in (1, 2):
            p.encode_num_bytes = num_bytes
        elif num_bytes in (-1, 4):
            p.encode_num_bytes = -1
        else:
            raise RuntimeError(f"Invalid encoding size. Must be -1, 1, 2 or 4. Got {num_bytes}")
        p.results_per_pixel = 8
        p.x_start_min, p.x_start_max, p.y_start_min, p.y_start_max = 0, self._width, 0, self._height

    def set_min_obs(self, new_value):
        if new_value < 0:
            raise RuntimeError(f"min_obs must be >= 0. Got {new_value}")
        if new_value > self._num_imgs:
            raise RuntimeError(f"min_obs cannot be greater than the number of images. min_obs = {new_value}, "
                               f"num_imgs = {self._num_imgs}.")
        self.params.min_observations = int(new_value)

    def set_min_lh(self, new_value):
        self.params.min_lh = _narrow(new_value)

    def set_results_per_pixel(self, new_value):
        if new_value <= 0:
            raise RuntimeError(f"Invalid results per pixel. Got {new_value}")
        self.params.results_per_pixel = int(new_value)

    def enable_gpu_sigmag_filter(self, percentiles, sigmag_coeff, min_lh):
        pc = [float(v) for v in percentiles]
        if len(pc) != 2 or pc[0] >= pc[1] or pc[0] <= 0.0 or pc[1] >= 1.0:
            raise RuntimeError(f"Invalid percentiles for sigma G filtering. Got {pc}.")
        if sigmag_coeff <= 0.0:
            raise RuntimeError(f"Invalid coefficient for sigma G filtering. Got {sigmag_coeff}.")
        p = self.params
        p.do_sigmag_filter, p.sgl_L, p.sgl_H = True, _narrow(pc[0]), _narrow(pc[1])
        p.sigmag_coeff, p.min_lh = _narrow(sigmag_coeff), _narrow(min_lh)

    def disable_gpu_sigmag_filter(self):
        self.params.do_sigmag_filter = False

    def enable_gpu_encoding(self, num_bytes):
        """Compatibility alias from older kbmod releases: the encoding is fixed at construction
        (5th ctor argument, stack_search.cpp:37-38); this only checks that the request matches."""
        want = num_bytes if num_bytes in (1, 2) else 4
        if want != self._ppa.num_bytes:
            raise RuntimeError("The psi/phi encoding is chosen at construction: StackSearch(..., num_bytes).")

    def set_start_bounds_x(self, x_min, x_max):
        if x_min >= x_max:
            raise RuntimeError(f"Invalid search bounds for the x pixel [{x_min}, {x_max}]")
        self.params.x_start_min, self.params.x_start_max = int(x_min), int(x_max)

    def set_start_bounds_y(self, y_min, y_max):
        if y_min >= y_max:
            raise RuntimeError(f"Invalid search bounds for the y pixel [{y_min}, {y_max}]")
        self.params.y_start_min, self.params.y_start_max = int(y_min), int(y_max)

    def preload_psi_phi_array(self):
        if not self._ppa.on_gpu:
            self._ppa.move_to_gpu()
            self._preloaded = True

    def unload_psi_phi_array(self):
        # also frees the copy the constructor left on the device when the stack was never "moved to the GPU"
        self._ppa.clear_from_gpu()
        self._preloaded = False

    def psi_phi_array_on_gpu(self):
        return self._ppa.on_gpu

    def set_kernel_mode(self, mode):
        """0 = automatic, 1 = generic kernel, 2 = tiled TMA kernel, 3 = candidate-parallel kernel (extension; used by
        the parity tests)."""
        check(lib.kb_set_kernel_mode(self._ppa._require(), C.c_int(mode)))

    # -- search -------------------------------------------------------------------------------------
    def compute_max_results(self):
        p = self.params
        if p.x_start_min >= p.x_start_max:
            raise RuntimeError(f"Invalid search bounds for the x pixel [{p.x_start_min}, {p.x_start_max}]")
        if p.y_start_min >= p.y_start_max:
            raise RuntimeError(f"Invalid search bounds for the y pixel [{p.y_start_min}, {p.y_start_max}]")
        return (p.x_start_max - p.x_start_min) * (p.y_start_max - p.y_start_min) * p.results_per_pixel

    def search_all(self, search_list, on_gpu=True):
        """stack_search.cpp:221-284.  ``search_list``: list[Trajectory] or a ``(vx, vy)`` pair of float32 arrays."""
        if not on_gpu:
            raise RuntimeError("kbmod_b200 has no CPU search path; call search_all(candidates, True).")
        if isinstance(search_list, tuple) and len(search_list) == 2:
            vx, vy = f32(search_list[0]), f32(search_list[1])
        else:
            for i, t in enumerate(search_list):
                if not t.is_valid():
                    raise RuntimeError(f"Invalid trajectory detected at index {i}: {t}")
            vx = np.array([t.vx for t in search_list], dtype=np.float32)
            vy = np.array([t.vy for t in search_list], dtype=np.float32)
        if vx.shape != vy.shape:
            raise RuntimeError("vx and vy must have the same length.")
        max_results = self.compute_max_results()
        p = self.params
        self._log.info(f"Searching X=[{p.x_start_min}, {p.x_start_max}] Y=[{p.y_start_min}, {p.y_start_max}]\n"
                       f"Allocating space for {max_results} results.")
        timer = DebugTimer("Running batch search", self._log)
        n = C.c_uint64(0)
        cp = p._c()
        self._result_cache = None
        check(lib.kb_search(self._ppa._require(), C.byref(cp), ptr(vx), ptr(vy), C.c_uint64(vx.size), C.byref(n)))
        note = C.c_char_p()
        check(lib.kb_search_note(self._ppa._require(), C.byref(note)))
        self.last_search_note = note.value.decode("utf-8", "replace") if note.value else ""
        if self.last_search_note:   # the search ran, but not at the throughput kernel's usual speed: say why
            self._log.warning("kbmod_b200: " + self.last_search_note)
        self._log.debug(f"Core search returned {n.value} results after filtering by LH and min_obs.")
        timer.stop()

    def timer_start(self):
        """Record a CUDA event on the search stream (device-side timing of a region of calls)."""
        check(lib.kb_timer_start(self._ppa._require()))

    def timer_stop(self):
        ms = C.c_float(0)
        check(lib.kb_timer_stop(self._ppa._require(), C.byref(ms)))
        return ms.value

    def last_timing(self):
        t = KbTiming()
        check(lib.kb_get_timing(self._ppa._require(), C.byref(t)))
        return {f[0]: getattr(t, f[0]) for f in KbTiming._fields_}

    def get_number_total_results(self):
        n = C.c_uint64(0)
        check(lib.kb_num_results(self._ppa._require(), C.byref(n)))
        return n.value

    def get_results_array(self, start=0, count=None):
        """Results [start, start+count) as a numpy structured array (28-byte Trajectory records)."""
        total = self.get_number_total_results()
        if count is None:
            count = max(total - start, 0)
        if count == 0:
            return np.zeros(0, dtype=TRAJECTORY_DTYPE)
        out = np.zeros(count, dtype=TRAJECTORY_DTYPE)
        n = C.c_uint64(0)
        check(lib.kb_get_results(self._ppa._require(), C.c_uint64(start), C.c_uint64(count), ptr(out), C.byref(n)))
        return out[: n.value]

    def get_results(self, start, count):
        if count == 0:
            raise RuntimeError("count must be greater than 0")
        self._log.debug(f"Reading results [{start}, {start + count})")
        return array_to_trajectories(self.get_results_array(start, count))

    def get_all_results(self):
        return array_to_trajectories(self.get_results_array())

    def get_slots_array(self):
        """Raw K*P per-pixel slots of the last search (kernels.cu:285-301,322-330), before the filter."""
        n = C.c_uint64(0)
        check(lib.kb_get_slots(self._ppa._require(), None, C.c_uint64(0), C.byref(n)))
        out = np.zeros(n.value, dtype=TRAJECTORY_DTYPE)
        check(lib.kb_get_slots(self._ppa._require(), ptr(out), C.c_uint64(n.value), C.byref(n)))
        return out

    def set_results(self, new_results):
        arr = trajectories_to_array(new_results)
        check(lib.kb_set_results(self._ppa._require(), ptr(arr), C.c_uint64(arr.size)))

    def clear_results(self):
        check(lib.kb_clear_results(self._ppa._require()))

    def evaluate_single_trajectory(self, trj, use_kernel=True):
        if not use_kernel:
            raise RuntimeError("kbmod_b200 has no CPU path; call evaluate_single_trajectory(trj, True).")
        arr = trajectories_to_array([trj])
        cp = self.params._c()
        check(lib.kb_evaluate_trajectories(self._ppa._require(), C.byref(cp), ptr(arr), C.c_uint64(1), None))
        trj.lh, trj.flux, trj.obs_count = float(arr[0]["lh"]), float(arr[0]["flux"]), int(arr[0]["obs_count"])

    def evaluate_trajectories_array(self, arr, want_mask=False):
        """Vectorised evaluate_single_trajectory; optionally returns the per-observation keep mask [n, T]."""
        arr = trajectories_to_array(arr).copy()
        mask = np.zeros((arr.size, self._num_imgs), dtype=np.uint8) if want_mask else None
        cp = self.params._c()
        check(lib.kb_evaluate_trajectories(self._ppa._require(), C.byref(cp), ptr(arr), C.c_uint64(arr.size),
                                           ptr(mask) if want_mask else None))
        return (arr, mask) if want_mask else arr

    def search_linear_trajectory(self, x, y, vx, vy, use_kernel=True):
        t = Trajectory(x=x, y=y, vx=vx, vy=vy)
        self.evaluate_single_trajectory(t, use_kernel)
        return t

    def get_all_psi_phi_curves(self, trajectories):
        arr = trajectories_to_array(trajectories)
        out = np.zeros((arr.size, 2 * self._num_imgs), dtype=np.float32)
        if arr.size:
            check(lib.kb_psi_phi_curves(self._ppa._require(), ptr(arr), C.c_uint64(arr.size), ptr(out)))
        return out


# ---- logging bridge and timers (logging.h, debug_timer.{h,cpp}) -----------------------------------------
class Logging:
    """Stand-in for the C++<->Python logging registry (logging.h:135-237): hands out Python loggers."""

    @staticmethod
    def getLogger(name="kbmod"):
        return _pylogging.getLogger(name)

    @staticmethod
    def registerLogger(logger):
        return logger

    @staticmethod
    def setConfig(config):
        return None


class DebugTimer:
    """debug_timer.cpp:13-54: wall-clock timer that logs at DEBUG."""

    def __init__(self, name, logger=None):
        self._name = name
        self._log = Logging.getLogger(logger) if isinstance(logger, str) else (logger or Logging.getLogger("kbmod"))
        self.start()

    def start(self):
        self._t0, self._t1, self._running = _time.perf_counter(), None, True
        self._log.debug(f"Starting {self._name} timer.")

    def stop(self):
        self._t1, self._running = _time.perf_counter(), False
        self._log.debug(f"Finished {self._name} in {self.read():.6f} seconds.")

    def read(self):
        end = _time.perf_counter() if self._running else self._t1
        return end - self._t0
