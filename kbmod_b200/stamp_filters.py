"""Coadded stamps of the search results on the device -- the reference's ``kbmod.filters.stamp_filters.append_coadds``
(filters/stamp_filters.py:72-160) over ``kbmod.core.stamp_utils`` (extract_stamp_stack, coadd_sum / coadd_mean /
coadd_median / coadd_weighted, core/stamp_utils.py:16-349), through ``kb_coadd_stamps`` (csrc/stamps.cu).

No CPU path.  Per-night coadds (``nightly=True``) are one device pass per calendar day.  Results are float32 ``R x S x S`` arrays (S = 2 * radius + 1), bit-identical to the reference's.
"""
import ctypes as C

import numpy as np

from ._lib import check, f32, lib, ptr, ptr_array, require_gpu

COADD_TYPES = ("sum", "mean", "median", "weighted")


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def coadd_stamps(sci, var, zeroed_times, x, y, vx, vy, radius, coadd_types=COADD_TYPES, obs_valid=None):
    """Coadds for R trajectories.  ``sci`` / ``var``: T images (list or T x H x W array); ``x, y, vx, vy``: length-R
    arrays (the float64 columns of the reference's result table); ``obs_valid``: optional R x T booleans selecting the
    images of each trajectory (``valid_only``).  Returns {type: float32[R, S, S]}."""
    if radius <= 0:
        raise ValueError(f"Invalid stamp radius {radius}")
    for t in coadd_types:
        if t not in COADD_TYPES:
            raise ValueError(f"Unknown coadd type {t}")
    sci = [f32(a) for a in sci]
    T = len(sci)
    if T == 0:
        raise ValueError("empty image stack")
    H, W = sci[0].shape
    need_var = "weighted" in coadd_types
    if need_var:
        var = [f32(a) for a in var]
        if len(var) != T:
            raise ValueError("science and variance stacks differ in length")
    times = _f64(zeroed_times)
    if times.shape != (T,):
        raise ValueError("one time per image is required")
    x, y, vx, vy = _f64(x), _f64(y), _f64(vx), _f64(vy)
    R = x.size
    if not (y.size == R and vx.size == R and vy.size == R):
        raise ValueError("x, y, vx, vy must have the same length")
    ov = None
    if obs_valid is not None:
        ov = np.ascontiguousarray(obs_valid, dtype=np.uint8)
        if ov.shape != (R, T):
            raise ValueError("Wrong shape for obs_valid")
    S = 2 * radius + 1
    out = {t: np.zeros((R, S, S), dtype=np.float32) for t in coadd_types}
    require_gpu()
    if R:
        check(lib.kb_coadd_stamps(ptr_array(sci), ptr_array(var) if need_var else None, C.c_int(T), C.c_int(H), C.c_int(W),
                                  ptr(times), ptr(x), ptr(y), ptr(vx), ptr(vy), C.c_uint64(R), C.c_int(radius),
                                  ptr(ov) if ov is not None else None,
                                  *[ptr(out[t]) if t in out else None for t in COADD_TYPES]))
    return out


def extract_all_stamps(sci, zeroed_times, x, y, vx, vy, radius):
    """Every stamp of every trajectory: float32 ``R x T x S x S``, NaN where there is no image
    (extract_stamp_stack, core/stamp_utils.py:16-84, for all results at once)."""
    if radius < 1:
        raise ValueError(f"Invalid stamp radius: {radius}")
    sci = [f32(a) for a in sci]
    T = len(sci)
    if T == 0:
        raise ValueError("empty image stack")
    H, W = sci[0].shape
    times = _f64(zeroed_times)
    if times.shape != (T,):
        raise ValueError("one time per image is required")
    x, y, vx, vy = _f64(x), _f64(y), _f64(vx), _f64(vy)
    R = x.size
    if not (y.size == R and vx.size == R and vy.size == R):
        raise ValueError("x, y, vx, vy must have the same length")
    S = 2 * radius + 1
    out = np.zeros((R, T, S, S), dtype=np.float32)
    require_gpu()
    if R:
        check(lib.kb_extract_stamps(ptr_array(sci), C.c_int(T), C.c_int(H), C.c_int(W), ptr(times), ptr(x), ptr(y),
                                    ptr(vx), ptr(vy), C.c_uint64(R), C.c_int(radius), ptr(out)))
    return out


def extract_curve_values(imgs, x_vals, y_vals):
    """core/stamp_utils.py:87-141: the values of a stack of T images (a T x H x W array or a list of H x W arrays) at
    predicted integer positions -- a length-T array for one trajectory, R x T for R of them; NaN off the image."""
    imgs = [f32(a) for a in imgs]
    num_times = len(imgs)
    x_vals = np.asanyarray(x_vals, dtype=int)
    if x_vals.ndim == 1:
        x_vals = x_vals[np.newaxis, :]
    if x_vals.shape[1] != num_times:
        raise ValueError(f"X values must have the same length as times ({num_times}).")
    y_vals = np.asanyarray(y_vals, dtype=int)
    if y_vals.ndim == 1:
        y_vals = y_vals[np.newaxis, :]
    if y_vals.shape[1] != num_times:
        raise ValueError(f"Y values must have the same length as times ({num_times}).")
    if x_vals.shape[0] != y_vals.shape[0]:
        raise ValueError("X and Y values must have the same number of results.")
    require_gpu()
    H, W = imgs[0].shape
    xv, yv = np.ascontiguousarray(x_vals, dtype=np.int64), np.ascontiguousarray(y_vals, dtype=np.int64)
    out = np.full(xv.shape, np.nan)
    if xv.size:
        check(lib.kb_extract_curve_values(ptr_array(imgs), C.c_int(num_times), C.c_int(H), C.c_int(W), ptr(xv), ptr(yv),
                                          C.c_uint64(xv.shape[0]), ptr(out)))
    return out.flatten() if x_vals.shape[0] == 1 else out


def append_all_stamps(result_data, im_stack, stamp_radius):
    """filters/stamp_filters.py:163-211: the ``all_stamps`` column (R x T x S x S float32)."""
    if stamp_radius < 1:
        raise ValueError(f"Invalid stamp radius: {stamp_radius}")
    result_data.table["all_stamps"] = extract_all_stamps(im_stack.sci, im_stack.zeroed_times, result_data["x"],
                                                         result_data["y"], result_data["vx"], result_data["vy"],
                                                         stamp_radius)


def append_coadds(result_data, im_stack, coadd_types, radius, valid_only=True, nightly=False):
    """filters/stamp_filters.py:72-160 for a reference ``Results``-like table (columns x, y, vx, vy[, obs_valid]; new
    columns are written through ``result_data.table[name] = array``) and an ``ImageStackPy``-like stack (``sci``,
    ``var``, ``zeroed_times``)."""
    if radius <= 0:
        raise ValueError(f"Invalid stamp radius {radius}")
    valid_only = valid_only and "obs_valid" in result_data.colnames
    ov = np.asarray(result_data["obs_valid"], dtype=bool) if valid_only else None
    args = (im_stack.sci, im_stack.var, im_stack.zeroed_times, result_data["x"], result_data["y"],
            result_data["vx"], result_data["vy"], radius, tuple(coadd_types))
    for t, a in coadd_stamps(*args, ov).items():
        result_data.table[f"coadd_{t}"] = a
    if nightly:
        # one more set of coadds per calendar day (stamp_filters.py:101-106,151-168): the images of that day only
        day_strs = np.array([f"_{mjd_to_day(t)}" for t in im_stack.times])
        num_res = len(result_data["x"])
        for day in np.unique(day_strs):
            day_mask = np.broadcast_to(day_strs == day, (num_res, len(day_strs)))
            mask = day_mask if ov is None else (ov & day_mask)
            for t, a in coadd_stamps(*args, mask).items():
                result_data.table[f"coadd_{t}{day}"] = a


def mjd_to_day(mjd):
    """util_functions.py:52-65 without astropy: the UTC calendar date (YYYY-MM-DD) of a modified Julian date."""
    import datetime
    import math
    return (datetime.date(1858, 11, 17) + datetime.timedelta(days=math.floor(float(mjd)))).strftime("%Y-%m-%d")
