"""Seeded synthetic inputs shared by the golden-fixture generator and the tests (numpy only)."""
import numpy as np


def gaussian_psf(std):
    r = int(3 * std)
    x = np.arange(-r, r + 1)
    xx, yy = np.meshgrid(x, x)
    k = np.exp(-0.5 * (xx**2 + yy**2) / std**2)
    k /= k.sum()
    return k.astype(np.float32)


def make_case(seed=11, T=10, H=24, W=32, n_cand=40, mask_frac=0.05, psf_std=1.0, vmax=6.0, objects=2,
              odd_pixels=True):
    """A small stack with masks, a zero-variance pixel, an infinite pixel, injected movers and a candidate
    list that contains duplicates / near-duplicates (exact lh ties) and one far off-image velocity."""
    rng = np.random.default_rng(seed)
    times = np.sort(rng.random(T) * 3.0)
    times = times - times[0]
    sci = [rng.normal(0, 2, (H, W)).astype(np.float32) for _ in range(T)]
    var = [(4.0 * (1 + 0.1 * rng.random((H, W)))).astype(np.float32) for _ in range(T)]
    for t in range(T):
        m = rng.random((H, W)) < mask_frac
        sci[t][m] = np.nan
        var[t][m] = np.nan
        if odd_pixels:
            var[t][3 % H, 4 % W] = 0.0          # var == 0 -> invalid psi and phi (image_utils_cpp.cpp:142)
            sci[t][7 % H, 7 % W] = np.inf       # non-finite science pixel
            m2 = rng.random((H, W)) < 0.01      # science-only masks: psi invalid, phi valid
            sci[t][m2] = np.nan
    psf = gaussian_psf(psf_std)
    vx = rng.uniform(-vmax, vmax, n_cand).astype(np.float32)
    vy = rng.uniform(-vmax, vmax, n_cand).astype(np.float32)
    objs = []
    for k in range(objects):
        ox, oy = int(rng.integers(4, W - 4)), int(rng.integers(4, H - 4))
        ovx, ovy = vx[3 + k], vy[3 + k]
        objs.append((ox, oy, float(ovx), float(ovy)))
        for t in range(T):
            x = int(ox + ovx * times[t] + 0.5)
            y = int(oy + ovy * times[t] + 0.5)
            for dy in range(-1, 2):
                for dx in range(-1, 2):
                    if 0 <= x + dx < W and 0 <= y + dy < H and np.isfinite(sci[t][y + dy, x + dx]):
                        sci[t][y + dy, x + dx] += np.float32(150.0 * psf[psf.shape[0] // 2 + dy, psf.shape[0] // 2 + dx])
    if n_cand >= 12:
        vx[0], vy[0] = 0.0, 0.0
        vx[1], vy[1] = 0.01, 0.0      # same pixels as candidate 0 -> exact lh tie
        vx[2], vy[2] = 0.0, 0.01      # ditto
        vx[8], vy[8] = vx[5], vy[5]   # exact duplicate
        vx[9], vy[9] = 500.0, -300.0  # leaves the image after t=0
        vx[10], vy[10] = 0.5, -0.5    # half-integer products when times are multiples of 1
    return dict(times=times, sci=sci, var=var, psfs=[psf] * T, vx=vx, vy=vy, objects=objs, T=T, H=H, W=W)


# (min_obs, min_lh, sigmag (L, H, coeff) or None, bounds (x0, x1, y0, y1) or None -> full frame, K)
SEARCH_CONFIGS = [
    (0, -1.0, None, None, 4),
    (5, 2.0, None, (-5, 37, -3, 27), 8),
    (10, 0.0, None, (3, 20, 2, 9), 1),
    (0, -1.0, (0.25, 0.75, 0.7413), None, 4),
    (5, 2.0, (0.25, 0.75, 0.7413), (-5, 37, -3, 27), 8),
    (7, 0.5, (0.15, 0.85, 0.4824), (3, 20, 2, 9), 2),
    (0, -3.0e38, None, (0, 8, 0, 8), 3),          # nothing filtered: empty slots survive too
    (3, 1.0, None, None, 12),                     # K > 8: list kept in global memory
]
