"""ctypes binding of ``libkbmod_b200.so`` (C-ABI declared in ``include/kbmod_b200.h``).

There is no CPU fallback: if the shared library is missing this module raises at import time, and every
compute entry point raises ``RuntimeError`` when no CUDA device is present.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# KBMOD_B200_LIB: load another build of the same library (kernel-tuning experiments); default is the in-tree build
LIB_PATH = os.environ.get("KBMOD_B200_LIB") or os.path.join(_HERE, "libkbmod_b200.so")


class KbTrajectory(C.Structure):
    _fields_ = [("vx", C.c_float), ("vy", C.c_float), ("lh", C.c_float), ("flux", C.c_float),
                ("x", C.c_int32), ("y", C.c_int32), ("obs_count", C.c_int32)]


class KbParams(C.Structure):
    _fields_ = [("min_observations", C.c_int32), ("min_lh", C.c_float), ("do_sigmag_filter", C.c_int32),
                ("sgl_L", C.c_float), ("sgl_H", C.c_float), ("sigmag_coeff", C.c_float),
                ("encode_num_bytes", C.c_int32), ("x_start_min", C.c_int32), ("x_start_max", C.c_int32),
                ("y_start_min", C.c_int32), ("y_start_max", C.c_int32), ("results_per_pixel", C.c_uint32),
                ("total_results", C.c_uint64)]


class KbMeta(C.Structure):
    _fields_ = [("num_times", C.c_uint64), ("width", C.c_uint64), ("height", C.c_uint64),
                ("pixels_per_image", C.c_uint64), ("num_entries", C.c_uint64), ("block_size", C.c_uint64),
                ("total_array_size", C.c_uint64), ("num_bytes", C.c_int32), ("psi_min_val", C.c_float),
                ("psi_max_val", C.c_float), ("psi_scale", C.c_float), ("phi_min_val", C.c_float),
                ("phi_max_val", C.c_float), ("phi_scale", C.c_float)]


class KbTiming(C.Structure):
    _fields_ = [("ms_h2d", C.c_float), ("ms_prepare", C.c_float), ("ms_search", C.c_float),
                ("ms_filter_sort", C.c_float), ("ms_total", C.c_float), ("evals", C.c_uint64),
                ("search_launches", C.c_uint32), ("total_launches", C.c_uint32), ("used_tiled", C.c_uint32),
                ("ms_main_kernel", C.c_float), ("h2d_stack_bytes", C.c_uint64)]


# numpy view of kb_trajectory (common.h:55-67; 28 bytes)
TRAJECTORY_DTYPE = np.dtype([("vx", "<f4"), ("vy", "<f4"), ("lh", "<f4"), ("flux", "<f4"), ("x", "<i4"),
                             ("y", "<i4"), ("obs_count", "<i4")])
assert TRAJECTORY_DTYPE.itemsize == 28 and C.sizeof(KbTrajectory) == 28

# Every symbol include/kbmod_b200.h declares (checked by tests/test_cabi_symbols.py against the header).
SYMBOLS = [
    "kb_last_error", "kb_version", "kb_device_count", "kb_gpu_memory", "kb_set_device", "kb_device_pci_bus_id", "kb_convolve", "kb_generate_psi_phi",
    "kb_create_from_images", "kb_create_from_images_rows", "kb_create_from_psi_phi", "kb_destroy", "kb_get_meta", "kb_host_array",
    "kb_read_psi_phi", "kb_read_time", "kb_preload", "kb_unload", "kb_on_gpu", "kb_search", "kb_get_results",
    "kb_num_results", "kb_set_results", "kb_clear_results", "kb_get_timing", "kb_get_slots",
    "kb_set_kernel_mode", "kb_search_note", "kb_evaluate_trajectories", "kb_psi_phi_curves", "kb_sigmag_filtered_indices",
    "kb_timer_start", "kb_timer_stop", "kb_measure_peaks", "kb_sort_trajectories_static", "kb_nccl_unique_id", "kb_nccl_init", "kb_nccl_finalize", "kb_merge_results_allgather", "kb_merge_sorted_lists",
    "kb_sigmag_clip_matrix", "kb_sigmag_clip_curve", "kb_extract_curve_values", "kb_update_likelihood", "kb_sigmag_filter_curves", "kb_coadd_stamps", "kb_grid_filter", "kb_extract_stamps",
]

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `make -C kbmod_b200/csrc` (or __graft_entry__.build()). "
        "kbmod_b200 has no CPU fallback.")

lib = C.CDLL(LIB_PATH)
lib.kb_last_error.restype = C.c_char_p
lib.kb_version.restype = C.c_char_p
for _name in SYMBOLS:
    if _name not in ("kb_last_error", "kb_version"):
        getattr(lib, _name).restype = C.c_int


def check(rc):
    """Raise RuntimeError (what the reference's std::runtime_error becomes in Python) on a non-zero code."""
    if rc != 0:
        raise RuntimeError(lib.kb_last_error().decode("utf-8", "replace"))


def device_count():
    n = C.c_int(0)
    check(lib.kb_device_count(C.byref(n)))
    return n.value


def require_gpu():
    if device_count() <= 0:
        raise RuntimeError("GPU is not available (kbmod_b200 has no CPU fallback).")


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def ptr_array(arrays):
    """C array of pointers to the given contiguous numpy arrays."""
    arr = (C.c_void_p * len(arrays))()
    for i, a in enumerate(arrays):
        arr[i] = a.ctypes.data
    return arr
