import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from cases import make_case, SEARCH_CONFIGS
from kbmod_b200 import search as kb
from oracle import cpu as orc
case = make_case()
g = np.load(os.path.join(ROOT, "tests/golden/small_case.npz"))
want = g["slots_f32_0"]
for mode in (1, 2):
    s = kb.StackSearch(case["sci"], case["var"], case["psfs"], case["times"], -1)
    s.set_kernel_mode(mode); s.set_min_obs(0); s.set_min_lh(-1.0); s.set_results_per_pixel(4)
    s.search_all((case["vx"], case["vy"]), True)
    got = s.get_slots_array()
    bad = np.flatnonzero((got["vx"].view(np.uint32) != want["vx"].view(np.uint32)) | (got["lh"].view(np.uint32) != want["lh"].view(np.uint32)))
    print("mode", mode, "nbad", bad.size, bad[:10])
    for i in (28, 64):
        print(" got ", got[i:i+4])
        print(" want", want[i:i+4])
print("cands 0-2,5,8:", [(case["vx"][i], case["vy"][i]) for i in (0,1,2,5,8)])
