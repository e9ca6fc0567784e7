import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kbmod_b200 import fake_data as fd, search as kb
W = int(sys.argv[1]) if len(sys.argv) > 1 else 512
nb = int(sys.argv[2]) if len(sys.argv) > 2 else -1
times = fd.create_fake_times(20, 57130.2, obs_per_day=4)
ds = fd.FakeDataSet(W, W, times, mask_fraction=0.01, psf_val=1.0, use_seed=101)
gen = fd.VelocityGridSearch(50, -40, 40, 40, -32, 32)
ds.insert_random_objects_from_generator(5, gen, 500.0)
vx, vy = gen.to_arrays()
st = ds.stack_py
s = kb.StackSearch(st.sci, st.var, st.psfs, st.zeroed_times, nb)
s.set_min_obs(10); s.enable_gpu_sigmag_filter([0.25, 0.75], 0.7413, 10.0)
s.preload_psi_phi_array()
res = {}
for mode in (1, 2):
    s.set_kernel_mode(mode)
    for _ in range(2):
        s.search_all((vx, vy), True)
    t = s.last_timing()
    res[mode] = s.get_results_array()
    print("mode", mode, "ms_search %.2f prepare %.2f post %.2f" % (t["ms_search"], t["ms_prepare"], t["ms_filter_sort"]), "evals/s %.3e" % (t["evals"] / t["ms_search"] * 1e3), "n", len(res[mode]), "tiled", t["used_tiled"])
a, b = res[1], res[2]
print("same lh sequence:", len(a) == len(b) and np.array_equal(a["lh"].view(np.uint32), b["lh"].view(np.uint32)))
