"""One preloaded search of a bench config with the timing breakdown (use KB_DEBUG=1 for plan / clip statistics)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from kbmod_b200 import search as kb
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
c = dict(bench.CONFIGS[name])
if len(sys.argv) > 3: c["n_obj"] = int(sys.argv[3])
stack, vx, vy = bench.make_inputs(c, 1)
s = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, c["nb"])
band = None
if c.get("band_div"):
    rows = c["H"] // c["band_div"]; band = (c["H"] // 2 - rows // 2, c["H"] // 2 - rows // 2 + rows)
bench.configure(s, c, band)
s.preload_psi_phi_array()
for i in range(reps):
    t0 = time.perf_counter(); s.search_all((vx, vy), True); t1 = time.perf_counter()
    t = s.last_timing()
    print(f"{name} nobj={c['n_obj']}: wall {1e3*(t1-t0):.1f} ms | prep {t['ms_prepare']:.1f} search {t['ms_search']:.1f} "
          f"(gather kernel {t['ms_main_kernel']:.1f}) post {t['ms_filter_sort']:.1f} | {t['evals']/t['ms_main_kernel']*1e3:.3e} evals/s "
          f"kernel-only, tiled={t['used_tiled']}, results={s.get_number_total_results()}", flush=True)
