"""Timing of the N2 step (sigma-G clip of the results' likelihood curves + likelihood update) through the C-ABI with
host buffers, beside the reference's formulation (filters/sigma_g_filter.py:114-168 restated with the same torch
calls, run on the CPU and -- as the reference does when a GPU is present -- on torch-CUDA) and numpy's
_update_likelihood (results.py:630-645).  Checks that all keep matrices agree.

    python scripts/bench_filters.py [R=200000] [T=100]
"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from kbmod_b200.sigma_g_filter import SigmaGClipping, apply_clipped_sigma_g_arrays, update_likelihood

R = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
T = int(sys.argv[2]) if len(sys.argv) > 2 else 100
rng = np.random.default_rng(3)
phi = (0.2 + 0.1 * rng.random((R, T))).astype(np.float32)
psi = (rng.normal(0, 1, (R, T)) * np.sqrt(phi) + 1.5 * phi).astype(np.float32)
bad = rng.random((R, T)) < 0.05
psi[bad] = 0; phi[bad] = 0
clip = SigmaGClipping(25, 75, 2)


def lh_curves(psi, phi):
    valid = (phi != 0) & np.isfinite(psi) & np.isfinite(phi)
    out = np.full(psi.shape, np.nan, dtype=np.float32)
    out[valid] = psi[valid] / np.sqrt(phi[valid])
    return out


def torch_clip(lh, device):
    t = torch.tensor(lh, device=device, dtype=torch.float32)
    q = torch.tensor([0.25, 0.5, 0.75], device=device, dtype=torch.float32)
    lo, med, hi = torch.nanquantile(t, q, dim=1)
    delta = hi - lo
    delta[delta < 1e-5] = 1e-5
    ns = clip.n_sigma * clip.coeff * delta
    lower = torch.reshape(med - ns, (lh.shape[0], 1)).expand(-1, T)
    upper = torch.reshape(med + ns, (lh.shape[0], 1)).expand(-1, T)
    return (torch.isfinite(t) & (t < upper) & (t > lower)).cpu().numpy().astype(bool)


def np_update(psi, phi, ov):
    ps, ph = (psi * ov).sum(axis=1), (phi * ov).sum(axis=1)
    nz = ph != 0
    lh = np.zeros(R); lh[nz] = ps[nz] / np.sqrt(ph[nz])
    return lh


def best(f, n=3):
    ts = []
    for _ in range(n):
        t0 = time.perf_counter(); r = f(); ts.append(time.perf_counter() - t0)
    return min(ts), r

t_lh, lh = best(lambda: lh_curves(psi, phi))
t_ref_cpu, m_cpu = best(lambda: torch_clip(lh, "cpu"))
t_upd, _ = best(lambda: np_update(psi, phi, m_cpu))
t_ours_clip, m_ours = best(lambda: clip.compute_clipped_sigma_g_matrix(lh))
t_ours_fused, (ov, l2, f2, o2) = best(lambda: apply_clipped_sigma_g_arrays(clip, psi, phi))
print(f"R={R} T={T}: curves (numpy) {1e3*t_lh:.1f} ms + clip (torch CPU, {torch.get_num_threads()} threads) {1e3*t_ref_cpu:.1f} ms "
      f"+ update (numpy) {1e3*t_upd:.1f} ms = {1e3*(t_lh+t_ref_cpu+t_upd):.1f} ms")
if torch.cuda.is_available():
    torch_clip(lh[:1000], "cuda")
    t_ref_gpu, m_gpu = best(lambda: torch_clip(lh, "cuda"))
    print(f"  reference clip on torch-CUDA (host in/out): {1e3*t_ref_gpu:.1f} ms; mask equal to torch CPU: "
          f"{bool(np.array_equal(m_gpu, m_cpu))} ({int((m_gpu != m_cpu).sum())} differ)")
print(f"  ours: clip only {1e3*t_ours_clip:.1f} ms; fused curves+clip+update {1e3*t_ours_fused:.1f} ms "
      f"(host in/out, {psi.nbytes*2/1e6:.0f} MB H2D) -> {(t_lh+t_ref_cpu+t_upd)/t_ours_fused:.1f}x the CPU pipeline")
print(f"  keep matrix equal to torch CPU: clip {bool(np.array_equal(m_ours, m_cpu))}, fused {bool(np.array_equal(ov, m_cpu))}")
