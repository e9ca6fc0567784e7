"""Timing of the "next" rows around the search through the C-ABI with host buffers:
  N1  get_all_psi_phi_curves for R results x T images, gathered from the reference-layout bytes the constructor
      keeps on the device (stack_search.cpp:14-31,302-318; OpenMP + host memory in the reference);
  N3  TrajectoryExplorer's shape: an 11 x 11 start window, 10 000 candidates, results_per_pixel = 10 000
      (trajectory_explorer.py:221-255), which runs on the candidate-parallel kernel.

    python scripts/bench_curves.py [R=100000] [T=100] [side=1024]
"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench
from kbmod_b200 import search as kb
from kbmod_b200._lib import TRAJECTORY_DTYPE

R = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
T = int(sys.argv[2]) if len(sys.argv) > 2 else 100
side = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
c = dict(bench.CONFIGS["C4"], H=side, W=side, T=T)
stack, vx, vy = bench.make_inputs(c, 1)
s = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, -1)
rng = np.random.default_rng(4)
trj = np.zeros(R, dtype=TRAJECTORY_DTYPE)
trj["x"], trj["y"] = rng.integers(0, side, R), rng.integers(0, side, R)
k = rng.integers(0, len(vx), R)
trj["vx"], trj["vy"] = vx[k], vy[k]
s.get_all_psi_phi_curves(trj[:10])
best = 1e9
for _ in range(3):
    t0 = time.perf_counter()
    out = s.get_all_psi_phi_curves(trj)
    best = min(best, time.perf_counter() - t0)
gb = (R * T * 8 + out.nbytes) / 1e9      # pairs gathered + curves written (and copied to the host)
print(f"N1 curves: R={R} T={T}: {1e3 * best:.2f} ms with host buffers ({R * T / best:.3e} values pairs/s, "
      f"{gb / best:.1f} GB/s of algorithmic bytes incl. the {out.nbytes / 1e6:.0f} MB result copy)")

n = 10000
ang = rng.uniform(-0.3, 0.3, n) + 0.4
vel = rng.uniform(1.0, 4.0, n)
pvx, pvy = (np.cos(ang) * vel).astype(np.float32), (np.sin(ang) * vel).astype(np.float32)
s.set_min_obs(1)
s.set_min_lh(-1e25)
s.set_results_per_pixel(n)
s.set_start_bounds_x(500, 511)
s.set_start_bounds_y(400, 411)
for mode, name in ((0, "automatic (candidate-parallel)"), (1, "generic (one thread per start pixel)")):
    if mode == 1 and len(sys.argv) > 4 and sys.argv[4] == "nogeneric":
        continue
    s.set_kernel_mode(mode)
    s.search_all((pvx, pvy), True)
    t0 = time.perf_counter()
    s.search_all((pvx, pvy), True)
    res = s.get_results_array()
    dt = time.perf_counter() - t0
    tm = s.last_timing()
    print(f"N3 11x11 window, {n} candidates, results_per_pixel {n}, T={T}: {name}: {1e3 * dt:.1f} ms "
          f"(device {tm['ms_total']:.1f} ms, kernel id {tm['used_tiled']}), {len(res)} results, "
          f"{121 * n * T / dt:.3e} evals/s")
