"""TEST INFRASTRUCTURE -- CPU restatement (numpy, float32) of the step that follows the search in the reference
pipeline: sigma-G clipping of the per-result likelihood curves and the likelihood update.

Follows /root/reference/src/kbmod (kbmod @ 617cd33):
  * ``lh_curves``                        results.py:568-606   (Results.compute_likelihood_curves)
  * ``clipped_sigma_g_matrix``           filters/sigma_g_filter.py:114-168 (SigmaGClipping.compute_clipped_sigma_g_matrix),
                                         whose percentiles are ``torch.nanquantile`` (pytorch 2.x, linear interpolation:
                                         rank = q*(n_valid-1) in float32, lerp(below, above, rank - floor(rank)))
  * ``update_likelihood``                results.py:608-645   (Results._update_likelihood; numpy's float32 pairwise sum)
  * ``find_sigma_g_coeff``               filters/sigma_g_filter.py:47-85

Pinned by tests/test_oracle_pins.py against fixtures produced by the reference's own class
(tests/golden/make_golden_filters.py -> tests/golden/sigmag_filter_case.npz) and, where /root/reference exists,
live against it.  Only tests/ may import this module.
"""
import math

import numpy as np

f32 = np.float32


def find_sigma_g_coeff(low_bnd, high_bnd):
    """sigma_g_filter.py:47-85 -- 1 / (Phi^-1(high) - Phi^-1(low)), percentiles on a [0, 100] scale."""
    from scipy.special import erfinv
    if (high_bnd <= low_bnd) or (low_bnd < 0) or (high_bnd > 100):
        raise ValueError(f"Invalid percentiles for sigma G coefficient [{low_bnd}, {high_bnd}]")

    def inv(z):
        sign = -1 if z < 0.5 else 1
        return float(sign * np.sqrt(2) * erfinv(sign * (2 * z - 1)))
    return 1 / (inv(high_bnd / 100.0) - inv(low_bnd / 100.0))


def lh_curves(psi, phi, obs_valid=None, mask_value=np.nan):
    """results.py:596-606."""
    psi = np.asarray(psi, dtype=f32)
    phi = np.asarray(phi, dtype=f32)
    valid = (phi != 0) & np.isfinite(psi) & np.isfinite(phi)
    if obs_valid is not None:
        valid = valid & np.asarray(obs_valid, dtype=bool)
    out = np.full(psi.shape, mask_value, dtype=f32)
    out[valid] = psi[valid] / np.sqrt(phi[valid])
    return out


def _lerp(a, b, w):
    """torch lerp as its vectorised CPU kernel and its CUDA kernel evaluate it (one fused multiply-add):
    w < 0.5 -> fma(w, b - a, a), else fma(w - 1, b - a, b).  Emulated in float64: the product of two float32
    numbers is exact in float64 and the sum is rounded once (double rounding cannot occur for these magnitudes
    except in sub-ulp corner cases the fixtures would expose)."""
    a, b, w = f32(a), f32(b), f32(w)
    d = f32(b - a)
    if abs(w) < 0.5:
        return f32(np.float64(w) * np.float64(d) + np.float64(a))
    return f32(np.float64(f32(w - f32(1))) * np.float64(d) + np.float64(b))


def nanquantile_row(row, qs):
    """torch.nanquantile(row, qs) for one float32 row: sort with NaN last, rank = float32(q * (n - 1))."""
    v = np.sort(row[~np.isnan(row)])
    n = v.size
    out = []
    for q in qs:
        if n == 0:
            out.append(f32(np.nan))
            continue
        rank = f32(f32(q) * f32(n - 1))
        if rank < 0:
            rank = f32(0)
        below = int(math.floor(rank))
        above = int(math.ceil(rank))
        w = f32(rank - f32(below))
        out.append(_lerp(v[below], v[above], w))
    return out


def clipped_sigma_g_matrix(lh, low_bnd=25, high_bnd=75, n_sigma=2, clip_negative=False, coeff=None):
    """sigma_g_filter.py:114-168.  Returns the boolean keep matrix."""
    lh = np.asarray(lh, dtype=f32)
    if coeff is None:
        coeff = find_sigma_g_coeff(low_bnd, high_bnd)
    qs = [f32(low_bnd / 100.0), f32(0.5), f32(high_bnd / 100.0)]
    mult = f32(n_sigma * coeff)          # python double product, narrowed when it meets the float32 tensor
    out = np.zeros(lh.shape, dtype=bool)
    with np.errstate(invalid="ignore"):
        for r in range(lh.shape[0]):
            row = lh[r]
            masked = np.where(row > 0, row, f32(np.nan)).astype(f32) if clip_negative else row
            lo, med, hi = nanquantile_row(masked, qs)
            delta = f32(hi - lo)
            if delta < f32(1e-5):
                delta = f32(1e-5)
            ns = f32(mult * delta)
            lower, upper = f32(med - ns), f32(med + ns)
            out[r] = np.isfinite(row) & (row < upper) & (row > lower)
    return out


def update_likelihood(psi, phi, obs_valid=None):
    """results.py:630-645: (likelihood, flux, obs_count) per row; float32 sums exactly as numpy adds them."""
    psi = np.asarray(psi, dtype=f32)
    phi = np.asarray(phi, dtype=f32)
    R, T = psi.shape
    if obs_valid is not None:
        ov = np.asarray(obs_valid, dtype=bool)
        phi_sum = (phi * ov).sum(axis=1)
        psi_sum = (psi * ov).sum(axis=1)
        num_obs = ov.sum(axis=1)
    else:
        phi_sum = phi.sum(axis=1)
        psi_sum = psi.sum(axis=1)
        num_obs = np.full(R, T)
    nz = phi_sum != 0
    lh = np.zeros(R)
    flux = np.zeros(R)
    with np.errstate(invalid="ignore", divide="ignore"):
        lh[nz] = psi_sum[nz] / np.sqrt(phi_sum[nz])
        flux[nz] = psi_sum[nz] / phi_sum[nz]
    return lh, flux, num_obs.astype(np.int64)


def pairwise_sum_f32(a):
    """numpy's float32 pairwise summation of a contiguous 1-D array (numpy/_core/src/umath/loops_utils.h.src,
    ``@TYPE@_pairwise_sum``), restated so the CUDA kernel's order can be checked without numpy internals."""
    a = np.asarray(a, dtype=f32)
    n = a.size
    if n < 8:
        res = f32(0.0) if n == 0 else None
        # numpy starts from -0.0 so that the sign of a lone -0.0 survives
        res = f32(-0.0)
        for x in a:
            res = f32(res + x)
        return res
    if n <= 128:
        r = [f32(a[i]) for i in range(8)]
        i = 8
        while i < n - (n % 8):
            for k in range(8):
                r[k] = f32(r[k] + a[i + k])
            i += 8
        res = f32(f32(f32(r[0] + r[1]) + f32(r[2] + r[3])) + f32(f32(r[4] + r[5]) + f32(r[6] + r[7])))
        while i < n:
            res = f32(res + a[i])
            i += 1
        return res
    n2 = n // 2
    n2 -= n2 % 8
    return f32(pairwise_sum_f32(a[:n2]) + pairwise_sum_f32(a[n2:]))


# ---- coadded stamps ------------------------------------------------------------------------
# core/stamp_utils.py:352-398 (extract_stamp), :403-439 (stack), :241-349 (coadds); positions as
# filters/stamp_filters.py:110-112 -> trajectory_utils.py:64-73 (float64, +0.5, astype(int) = truncation).
def stamp_positions(times, x0, v):
    pos = np.asarray(v, dtype=np.float64)[:, None] * np.asarray(times, dtype=np.float64)[None, :] \
        + np.asarray(x0, dtype=np.float64)[:, None]
    return (pos + 0.5).astype(int)


def extract_stamp_stack(imgs, xs, ys, radius, to_include=None):
    """T' x S x S float64 stack, NaN where there is no image (np.full(..., np.nan) is float64 in the reference)."""
    S = 2 * radius + 1
    sel = range(len(imgs)) if to_include is None else [t for t in range(len(imgs)) if to_include[t]]
    out = np.full((len(sel), S, S), np.nan)
    for k, t in enumerate(sel):
        img = imgs[t]
        H, W = img.shape
        for dy in range(S):
            y = int(ys[t]) - radius + dy
            if y < 0 or y >= H:
                continue
            for dx in range(S):
                x = int(xs[t]) - radius + dx
                if 0 <= x < W:
                    out[k, dy, dx] = img[y, x]
    return out


def coadd_sum(stack):
    out = np.zeros(stack.shape[1:])
    for k in range(stack.shape[0]):                      # np.nansum(axis=0): NaN -> 0, added image by image
        out = out + np.where(np.isnan(stack[k]), 0.0, stack[k])
    return out


def coadd_mean(stack):
    if stack.shape[0] == 0:
        return np.zeros(stack.shape[1:])
    cnt = (~np.isnan(stack)).sum(axis=0)
    s = coadd_sum(stack)
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.where(cnt > 0, s / cnt, 0.0)           # all-NaN pixels are zeroed first (stamp_utils.py:214-238)


def coadd_median(stack):
    """torch.nanmedian(dim=0): the LOWER of the two middle values for an even count; all-NaN -> 0."""
    out = np.zeros(stack.shape[1:])
    for i in range(stack.shape[1]):
        for j in range(stack.shape[2]):
            v = np.sort(stack[:, i, j][~np.isnan(stack[:, i, j])])
            out[i, j] = v[(v.size - 1) // 2] if v.size else 0.0
    return out


def coadd_weighted(stack, var_stack):
    if stack.shape[0] == 0:
        return np.zeros(stack.shape[1:])
    allnan = np.all(np.isnan(stack), axis=0)
    stack = np.where(allnan[None], 0.0, stack)
    ok = ~(np.isnan(stack) | np.isnan(var_stack) | (var_stack == 0.0))
    wsum = np.zeros(stack.shape[1:])
    ssum = np.zeros(stack.shape[1:])
    with np.errstate(invalid="ignore", divide="ignore"):
        for k in range(stack.shape[0]):
            w = np.where(ok[k], 1.0 / np.where(ok[k], var_stack[k], 1.0), 0.0)
            ssum = ssum + np.where(ok[k], stack[k] * w, 0.0)
            wsum = wsum + w
    wsum[wsum == 0.0] = 1e24
    return ssum / wsum


def coadds_for_results(sci, var, times, x, y, vx, vy, radius, obs_valid=None):
    """append_coadds (filters/stamp_filters.py:72-160, nightly=False): float32 R x S x S arrays per coadd type."""
    xs, ys = stamp_positions(times, x, vx), stamp_positions(times, y, vy)
    R, S = len(x), 2 * radius + 1
    out = {k: np.zeros((R, S, S), dtype=np.float32) for k in ("sum", "mean", "median", "weighted")}
    for r in range(R):
        inc = None if obs_valid is None else obs_valid[r]
        st = extract_stamp_stack(sci, xs[r], ys[r], radius, inc)
        vs = extract_stamp_stack(var, xs[r], ys[r], radius, inc)
        out["sum"][r] = coadd_sum(st)
        out["mean"][r] = coadd_mean(st)
        out["median"][r] = coadd_median(st)
        out["weighted"][r] = coadd_weighted(st, vs)
    return out


# ---- near-duplicate grid filter (filters/clustering_grid.py) --------------------------------------
def grid_filter_indices(x, y, vx, vy, lh, bin_width, max_time, vectorised=False):
    """Indices kept by TrajectoryClusterGrid: best lh per (start bin, end bin), strict '>' so the earlier of two equal
    trajectories stays, bins in first-seen order.  add_trajectory (clustering_grid.py:72-75) multiplies python
    doubles; so does the ``vectorised`` add_trajectory_list (:109-113), because extract_all_trajectory_vx returns a
    std::vector<float> that pybind11 turns into a list of Python floats, which numpy makes float64."""
    table, order = {}, []
    for i in range(len(x)):
        xi, yi = int(x[i]), int(y[i])
        ex, ey = max_time * float(vx[i]), max_time * float(vy[i])
        key = (int(xi / bin_width), int(yi / bin_width), int((xi + ex) / bin_width), int((yi + ey) / bin_width))
        if key not in table:
            table[key] = i
            order.append(key)
        elif float(lh[i]) > float(lh[table[key]]):
            table[key] = i
    return np.array([table[k] for k in order], dtype=np.int64)


# ---- single-curve sigma-G clip (filters/sigma_g_filter.py:86-112) ------------------------------------------------------
def _np_percentile_sorted(v, pct):
    """numpy.percentile, method 'linear', on an ascending float64 array, spelled out: virtual index (n - 1) * q,
    the two neighbours, and numpy's _lerp (from the nearer end)."""
    m = len(v)
    virt = (m - 1) * (pct / 100.0)
    if virt >= m - 1:
        return v[m - 1]
    lo = int(np.floor(virt))
    t = virt - np.floor(virt)
    d = v[lo + 1] - v[lo]
    return v[lo + 1] - d * (1 - t) if t >= 0.5 else v[lo] + d * t


def sigmag_clip_curve(lh, low_bnd, high_bnd, n_sigma, coeff, clip_negative):
    """Indices kept by SigmaGClipping.compute_clipped_sigma_g for one float64 likelihood curve."""
    lh = np.asarray(lh, dtype=np.float64)
    sel = lh[lh > 0] if clip_negative else lh
    if len(sel) == 0 or np.isnan(sel).any():
        return np.zeros(0, dtype=np.int64)
    v = np.sort(sel, kind="stable")
    lower, median, upper = (_np_percentile_sorted(v, q) for q in (low_bnd, 50, high_bnd))
    n_sigma_g = n_sigma * (coeff * max(upper - lower, 1e-8))
    return np.where(np.logical_and(lh > median - n_sigma_g, lh < median + n_sigma_g))[0].astype(np.int64)


def extract_curve_values(imgs, x_vals, y_vals):
    """core/stamp_utils.py:477-512 (_extract_curve_values): R x T float64, NaN off the image."""
    x_vals, y_vals = np.atleast_2d(np.asarray(x_vals, dtype=int)), np.atleast_2d(np.asarray(y_vals, dtype=int))
    out = np.full(x_vals.shape, np.nan)
    for t, img in enumerate(imgs):
        h, w = img.shape
        ok = (x_vals[:, t] >= 0) & (x_vals[:, t] < w) & (y_vals[:, t] >= 0) & (y_vals[:, t] < h)
        out[ok, t] = img[y_vals[ok, t], x_vals[ok, t]]
    return out
