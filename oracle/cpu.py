"""TEST INFRASTRUCTURE: ctypes front-end of the CPU restatement ``kbmod_oracle.c`` (numpy in / numpy out).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` leg may import this.  The
product package never does; it fails loudly when its CUDA library is missing instead.

Parity pinned against the reference: see the header of ``kbmod_oracle.c`` and ``tests/test_oracle_pins.py``.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "kbmod_oracle.c")
_LIB = os.path.join(_HERE, "_build", "libkbmod_oracle.so")
_lib = None


def build(force=False):
    """Compile kbmod_oracle.c with gcc (no fast-math, no FMA contraction)."""
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(_SRC):
        os.makedirs(os.path.dirname(_LIB), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-std=gnu11", "-fopenmp", "-ffp-contract=off", "-fno-fast-math",
                               "-fvisibility=hidden", "-shared", "-fPIC", _SRC, "-o", _LIB, "-lm"])
    return _LIB


class _Stack(C.Structure):
    _fields_ = [("data", C.c_void_p), ("times", C.c_void_p), ("T", C.c_int), ("H", C.c_int), ("W", C.c_int),
                ("num_bytes", C.c_int), ("psi_min", C.c_float), ("psi_scale", C.c_float),
                ("phi_min", C.c_float), ("phi_scale", C.c_float)]


class _Params(C.Structure):
    _fields_ = [("min_obs", C.c_int), ("min_lh", C.c_float), ("do_sigmag", C.c_int), ("sgl_L", C.c_float),
                ("sgl_H", C.c_float), ("coeff", C.c_float), ("x0", C.c_int), ("x1", C.c_int), ("y0", C.c_int),
                ("y1", C.c_int), ("K", C.c_int)]


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def convolve(img, psf):
    img, psf = _f32(img), _f32(psf)
    out = np.empty_like(img)
    lib().orc_convolve(_p(img), C.c_int(img.shape[0]), C.c_int(img.shape[1]), _p(psf), C.c_int(psf.shape[0]), _p(out))
    return out


def generate_psi(sci, var, psf):
    sci, var, psf = _f32(sci), _f32(var), _f32(psf)
    out = np.empty_like(sci)
    lib().orc_generate_psi(_p(sci), _p(var), C.c_int(sci.shape[0]), C.c_int(sci.shape[1]), _p(psf),
                           C.c_int(psf.shape[0]), _p(out))
    return out


def generate_phi(var, psf):
    var, psf = _f32(var), _f32(psf)
    out = np.empty_like(var)
    lib().orc_generate_phi(_p(var), C.c_int(var.shape[0]), C.c_int(var.shape[1]), _p(psf), C.c_int(psf.shape[0]), _p(out))
    return out


def scale_params(imgs, num_bytes):
    """(min, max, scale) over the finite pixels of all images (float32 triple)."""
    flat = _f32(np.asarray(imgs)).reshape(-1)
    out = np.zeros(3, dtype=np.float32)
    lib().orc_scale_params(_p(flat), C.c_uint64(flat.size), C.c_int(num_bytes), _p(out))
    return out


class PsiPhiStack:
    """psi/phi array in the reference layout [t][row][col][psi,phi] plus its meta data."""

    def __init__(self, data, times, T, H, W, num_bytes, psi_params=(0, 0, 1), phi_params=(0, 0, 1)):
        self.data = np.ascontiguousarray(data)
        self.times = np.ascontiguousarray(times, dtype=np.float64)
        self.T, self.H, self.W, self.num_bytes = T, H, W, num_bytes
        self.psi_params = np.asarray(psi_params, dtype=np.float32)
        self.phi_params = np.asarray(phi_params, dtype=np.float32)

    def _c(self):
        return _Stack(self.data.ctypes.data, self.times.ctypes.data, self.T, self.H, self.W, self.num_bytes,
                      float(self.psi_params[0]), float(self.psi_params[2]), float(self.phi_params[0]),
                      float(self.phi_params[2]))

    def read(self, t, row, col):
        out = np.zeros(2, dtype=np.float32)
        s = self._c()
        lib().orc_read_psi_phi(C.byref(s), C.c_int(t), C.c_int(row), C.c_int(col), _p(out))
        return out


def fill_from_psi_phi(psi_imgs, phi_imgs, times, num_bytes=-1):
    """fill_psi_phi_array (psi_phi_array.cpp:321-372)."""
    psi = _f32(np.asarray(psi_imgs))
    phi = _f32(np.asarray(phi_imgs))
    T, H, W = psi.shape
    nb = 4 if num_bytes in (-1, 4) else num_bytes
    if nb == 4:
        data = np.empty((T, H, W, 2), dtype=np.float32)
        data[..., 0] = psi
        data[..., 1] = phi
        return PsiPhiStack(data, times, T, H, W, 4)
    pp, fp = scale_params(psi, nb), scale_params(phi, nb)
    data = np.empty((T, H, W, 2), dtype=np.uint8 if nb == 1 else np.uint16)
    lib().orc_encode(_p(psi), _p(phi), C.c_uint64(psi.size), C.c_int(nb), _p(pp), _p(fp), _p(data))
    return PsiPhiStack(data, times, T, H, W, nb, pp, fp)


def fill_from_images(sci, var, psfs, times, num_bytes=-1):
    """fill_psi_phi_array_from_image_arrays (psi_phi_array.cpp:374-410), CPU convolution."""
    psi = [generate_psi(s, v, k) for s, v, k in zip(sci, var, psfs)]
    phi = [generate_phi(v, k) for v, k in zip(var, psfs)]
    return fill_from_psi_phi(psi, phi, times, num_bytes)


def make_params(min_obs=0, min_lh=0.0, sigmag=None, bounds=None, K=8, stack=None):
    """sigmag = (sgl_L, sgl_H, coeff) or None; bounds = (x0, x1, y0, y1) or None (full frame of `stack`)."""
    if bounds is None:
        bounds = (0, stack.W, 0, stack.H)
    sg = sigmag if sigmag is not None else (0.25, 0.75, -1.0)
    return _Params(int(min_obs), float(min_lh), int(sigmag is not None), float(sg[0]), float(sg[1]), float(sg[2]),
                   int(bounds[0]), int(bounds[1]), int(bounds[2]), int(bounds[3]), int(K))


def sigmag_filtered_indices(values, sgl0, sgl1, coeff, width):
    v = _f32(values)
    n = v.size
    idx = np.zeros(max(n, 1), dtype=np.int32)
    lo, hi = C.c_int(0), C.c_int(n - 1)
    lib().orc_sigmag_indices(_p(v), C.c_int(n), C.c_float(sgl0), C.c_float(sgl1), C.c_float(coeff), C.c_float(width),
                             _p(idx), C.byref(lo), C.byref(hi))
    return [int(i) for i in idx[lo.value : hi.value + 1]]


def eval_trajectory(stack, params, x, y, vx, vy, want_mask=False):
    s = stack._c()
    lh, fl, ob = C.c_float(), C.c_float(), C.c_int()
    mask = np.zeros(stack.T, dtype=np.uint8)
    lib().orc_eval_trajectory(C.byref(s), C.byref(params), C.c_int(x), C.c_int(y), C.c_float(vx), C.c_float(vy),
                              C.byref(lh), C.byref(fl), C.byref(ob), _p(mask) if want_mask else None)
    res = (np.float32(lh.value), np.float32(fl.value), ob.value)
    return res + (mask,) if want_mask else res


def eval_many(stack, params, x, y, vx, vy):
    """Vectorised eval_trajectory: returns (lh, flux, obs) arrays."""
    x = np.ascontiguousarray(x, dtype=np.int32)
    y = np.ascontiguousarray(y, dtype=np.int32)
    vx, vy = _f32(vx), _f32(vy)
    n = x.size
    lh, fl, ob = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32)
    s = stack._c()
    lib().orc_eval_many(C.byref(s), C.byref(params), C.c_int64(n), _p(x), _p(y), _p(vx), _p(vy), _p(lh), _p(fl), _p(ob))
    return lh, fl, ob


RESULT_DTYPE = np.dtype([("vx", "<f4"), ("vy", "<f4"), ("lh", "<f4"), ("flux", "<f4"), ("x", "<i4"), ("y", "<i4"),
                         ("obs_count", "<i4")])


def search_slots(stack, params, vx, vy):
    """Raw K*P slots exactly as searchFilterImages leaves them (plus the winning candidate index)."""
    vx, vy = _f32(vx), _f32(vy)
    n = (params.x1 - params.x0) * (params.y1 - params.y0) * params.K
    out = np.zeros(n, dtype=RESULT_DTYPE)
    ox, oy, oo, oc = (np.zeros(n, dtype=np.int32) for _ in range(4))
    ovx, ovy, olh, ofl = (np.zeros(n, dtype=np.float32) for _ in range(4))
    s = stack._c()
    lib().orc_search_slots(C.byref(s), C.byref(params), _p(vx), _p(vy), C.c_int64(vx.size), _p(ox), _p(oy), _p(ovx),
                           _p(ovy), _p(olh), _p(ofl), _p(oo), _p(oc))
    out["x"], out["y"], out["vx"], out["vy"], out["lh"], out["flux"], out["obs_count"] = ox, oy, ovx, ovy, olh, ofl, oo
    return out, oc


def filter_and_sort(slots, min_lh, min_obs):
    """Host post-processing of search_all (stack_search.cpp:268-277): drop lh<min_lh, drop obs<min_obs,
    sort by lh descending.  The reference's sort is unstable; ties are ordered here by (y, x, slot)."""
    keep = ~(slots["lh"] < np.float32(min_lh))
    keep &= ~(slots["obs_count"] < min_obs)
    r = slots[keep]
    order = np.argsort(-r["lh"].astype(np.float64), kind="stable")
    return r[order]


def search_all(stack, params, vx, vy):
    slots, _ = search_slots(stack, params, vx, vy)
    return filter_and_sort(slots, params.min_lh, params.min_obs)


def psi_phi_curve(stack, x, y, vx, vy):
    out = np.zeros(2 * stack.T, dtype=np.float32)
    s = stack._c()
    lib().orc_psi_phi_curve(C.byref(s), C.c_int(x), C.c_int(y), C.c_float(vx), C.c_float(vy), _p(out))
    return out
