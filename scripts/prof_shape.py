"""Profiling driver: a bench config at reduced size.  python scripts/prof_shape.py C5 512 512 [grid-divisor=2] [reps=2]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from kbmod_b200 import search as kb
name = sys.argv[1] if len(sys.argv) > 1 else "C5"
H = int(sys.argv[2]) if len(sys.argv) > 2 else 512
W = int(sys.argv[3]) if len(sys.argv) > 3 else 512
div = int(sys.argv[4]) if len(sys.argv) > 4 else 2
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 2
c = dict(bench.CONFIGS[name], H=H, W=W)
c.pop("band_div", None)
if "grid" in c:
    g = c["grid"]
    c["grid"] = (g[0] // div, g[1], g[2], g[3] // div, g[4], g[5])
else:
    e = c["ecliptic"]
    c["ecliptic"] = ([e[0][0], e[0][1], e[0][2] // div], [e[1][0], e[1][1], e[1][2] // div])
stack, vx, vy = bench.make_inputs(c, 1)
s = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, c["nb"])
bench.configure(s, c)
s.set_kernel_mode(2)
s.preload_psi_phi_array()
for _ in range(reps):
    s.search_all((vx, vy), True)
t = s.last_timing()
print(name, H, W, len(vx), "ms_search %.3f ms_main %.3f" % (t["ms_search"], t["ms_main_kernel"]),
      "evals/s %.4e" % (t["evals"] / t["ms_search"] * 1e3), "n", s.get_number_total_results())
