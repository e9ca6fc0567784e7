"""Smallest tiled-kernel invocation (debug aid): one search on a 24x32x10 stack with the tiled kernel forced."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from cases import make_case
from kbmod_b200 import search as kb
nb = int(sys.argv[1]) if len(sys.argv) > 1 else -1
case = make_case()
s = kb.StackSearch(case["sci"], case["var"], case["psfs"], case["times"], nb)
s.set_kernel_mode(2)
s.set_min_lh(-1.0); s.set_results_per_pixel(4)
s.search_all((case["vx"], case["vy"]), True)
print("ok", s.get_number_total_results(), s.last_timing())
