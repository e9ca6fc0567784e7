import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from kbmod_b200 import search as kb
c = bench.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "C2"]
stack, vx, vy = bench.make_inputs(c, 1)
s = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, c["nb"])
bench.configure(s, c)
for i in range(4):
    t0 = time.perf_counter(); s.search_all((vx, vy), True); t1 = time.perf_counter(); r = s.get_results_array(); t2 = time.perf_counter()
    t = s.last_timing()
    print("wall search %.1f ms, get %.1f ms | h2d %.1f prep %.1f search %.1f (main %.1f) post %.1f total %.1f" % ((t1-t0)*1e3, (t2-t1)*1e3, t["ms_h2d"], t["ms_prepare"], t["ms_search"], t["ms_main_kernel"], t["ms_filter_sort"], t["ms_total"]))
