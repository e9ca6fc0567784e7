"""Profiling driver: C2-shaped search (20 images, 2000-candidate velocity grid, sigma-G on) on a WxW frame."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kbmod_b200 import fake_data as fd, search as kb
W = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
nb = int(sys.argv[2]) if len(sys.argv) > 2 else -1
mode = int(sys.argv[3]) if len(sys.argv) > 3 else 0
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
nobj = int(sys.argv[5]) if len(sys.argv) > 5 else 5
sig = int(sys.argv[6]) if len(sys.argv) > 6 else 1
times = fd.create_fake_times(20, 57130.2, obs_per_day=4)
ds = fd.FakeDataSet(W, W, times, mask_fraction=0.01, psf_val=1.0, use_seed=101)
gen = fd.VelocityGridSearch(50, -40, 40, 40, -32, 32)
if nobj: ds.insert_random_objects_from_generator(nobj, gen, 500.0)
vx, vy = gen.to_arrays()
st = ds.stack_py
s = kb.StackSearch(st.sci, st.var, st.psfs, st.zeroed_times, nb)
s.set_min_obs(10)
if sig: s.enable_gpu_sigmag_filter([0.25, 0.75], 0.7413, 10.0)
else: s.set_min_lh(10.0)
s.set_kernel_mode(mode)
s.preload_psi_phi_array()
for _ in range(reps):
    s.search_all((vx, vy), True)
t = s.last_timing()
print("nobj", nobj, "sigmag", sig, "W", W, "nb", nb, "ms_search %.3f" % t["ms_search"], "evals/s %.4e" % (t["evals"] / t["ms_search"] * 1e3), "tiled", t["used_tiled"], "n", s.get_number_total_results())
