"""Full-size parity on start-pixel windows (C4 / C5 and friends): build the stack of a bench config with the
library, hand the SAME psi/phi bytes to the CPU oracle (they are memcmp-equal to the reference's by the unit
tests), and compare the search of a few start-pixel windows -- raw slots and the filtered, sorted list --
modulo exact lh ties (tests/parity.py).

    python scripts/check_window_parity.py C4 [window=24] [n_windows=3]
"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench
from kbmod_b200 import search as kb
from oracle import cpu as orc
from parity import assert_lists_equal_modulo_ties, assert_slots_equal_modulo_ties

name = sys.argv[1] if len(sys.argv) > 1 else "C4"
win = int(sys.argv[2]) if len(sys.argv) > 2 else 24
nwin = int(sys.argv[3]) if len(sys.argv) > 3 else 3
c = bench.CONFIGS[name]
t0 = time.perf_counter()
stack, vx, vy = bench.make_inputs(c, 1)
t1 = time.perf_counter()
s = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, c["nb"])
t2 = time.perf_counter()
print(f"{name}: inputs {t1 - t0:.1f} s, StackSearch ctor (H2D + psi/phi on device + D2H) {t2 - t1:.2f} s", flush=True)
bench.configure(s, c)
s.preload_psi_phi_array()
arr = s.psi_phi_array
nb = arr.num_bytes
pp = (arr.psi_min_val, arr.psi_max_val, arr.psi_scale) if nb != 4 else (0, 0, 1)
fp = (arr.phi_min_val, arr.phi_max_val, arr.phi_scale) if nb != 4 else (0, 0, 1)
ost = orc.PsiPhiStack(arr.host_array(), stack.zeroed_times, c["T"], c["H"], c["W"], nb, pp, fp)
sg = (0.25, 0.75, 0.7413) if c["sigmag"] else None
H, W = c["H"], c["W"]
objs = []
corners = [(int(o.x) - win // 2, int(o.y) - win // 2) for o in stack.objects[:max(nwin - 2, 1)]]
corners += [(-win // 2, H // 3), (W - win, H - win), (W // 2, H // 2)]
ok = True
for (x0, y0) in corners[:nwin]:
    b = (x0, x0 + win, y0, y0 + win)
    s.set_start_bounds_x(b[0], b[1]); s.set_start_bounds_y(b[2], b[3])
    ta = time.perf_counter()
    s.search_all((vx, vy), True)
    tb = time.perf_counter()
    got_slots, got = s.get_slots_array(), s.get_results_array()
    p = orc.make_params(c["min_obs"], c["min_lh"], sg, b, c["K"], ost)
    want_slots, _ = orc.search_slots(ost, p, vx, vy)
    tc = time.perf_counter()
    want = orc.filter_and_sort(want_slots, c["min_lh"], c["min_obs"])
    try:
        assert_slots_equal_modulo_ties(got_slots, want_slots, c["K"], orc, ost, p, f"{name} window {b}")
        assert_lists_equal_modulo_ties(got, want, orc, ost, p, f"{name} window {b}")
        verdict = "identical (modulo exact lh ties)"
    except AssertionError as e:
        ok = False
        verdict = f"MISMATCH: {e}"
    print(f"window x[{b[0]},{b[1]}) y[{b[2]},{b[3]}): {len(want)} results, tiled={s.last_timing()['used_tiled']}, "
          f"gpu {1e3 * (tb - ta):.1f} ms, oracle {tc - tb:.1f} s -> {verdict}", flush=True)
sys.exit(0 if ok else 1)
