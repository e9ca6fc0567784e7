"""O4: the UNMODIFIED reference's own GPU path -- StackSearch.search_all(candidates, True) from
oracle/_ref (its kernels.cu compiled for sm_100a) -- run on the same B200 as our library, on a bench config at FULL
frame, and compared record by record; both are timed the same way (wall clock around search_all, psi/phi
preloaded, best of `reps`).  Test infrastructure: never imported by the product.

    python scripts/check_ref_gpu.py C2 [reps=3] [mode=identity|psf]

mode=identity (default): the science images are convolved beforehand (by our library) and both sides get a
1x1 PSF [[1.0]], for which the reference's GPU convolution (FFMA-contracted, image_kernels.cu:29-66) and its CPU
convolution are the same function, so both sides search bit-identical psi/phi and the lists must be identical
modulo exact lh ties (tests/parity.py).  mode=psf: both sides get the raw images and the real PSF; the reference
then builds psi/phi with its FFMA-contracted GPU convolution (last-bit differences from its own CPU convolution,
which is what we reproduce), so records are matched by (x, y, vx, vy) and compared at the north star's 1e-4.
"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench
from kbmod_b200 import search as kb
from kbmod_b200._lib import TRAJECTORY_DTYPE
from oracle import cpu as orc, ref as oref
from parity import assert_lists_equal_modulo_ties, bits


def ref_results_array(S, ss):
    n = ss.get_number_total_results()
    out = np.zeros(n, dtype=TRAJECTORY_DTYPE)
    if n == 0:
        return out
    lst = ss.get_results(0, n)
    for f in ("x", "y", "vx", "vy", "lh", "flux", "obs_count"):
        out[f] = getattr(S, "extract_all_trajectory_" + f)(lst)
    return out


def run(name, reps=3, mode="identity", log=print):
    S, Hn = oref.load()
    if not S.kb_has_gpu():
        raise RuntimeError("the reference build sees no GPU")
    c = bench.CONFIGS[name]
    stack, vx, vy = bench.make_inputs(c, 1)
    T, H, W = c["T"], c["H"], c["W"]
    times = [float(t) for t in stack.zeroed_times]
    if mode == "identity":
        sci = [kb.convolve_image(np.ascontiguousarray(a, dtype=np.float32), p) for a, p in zip(stack.sci, stack.psfs)]
        psfs = [np.ones((1, 1), dtype=np.float32)] * T
    else:
        sci, psfs = list(stack.sci), list(stack.psfs)
    var = list(stack.var)
    evals = H * W * len(vx) * T

    # ---- ours
    t0 = time.perf_counter()
    s = kb.StackSearch(sci, var, psfs, times, c["nb"])
    t_ctor = time.perf_counter() - t0
    bench.configure(s, c)
    s.preload_psi_phi_array()
    ours_ms = []
    for _ in range(reps):
        t0 = time.perf_counter(); s.search_all((vx, vy), True); got = s.get_results_array()
        ours_ms.append(1e3 * (time.perf_counter() - t0))
    s.unload_psi_phi_array()
    t0 = time.perf_counter(); s.search_all((vx, vy), True); s.get_results_array()
    ours_cold = 1e3 * (time.perf_counter() - t0)

    # ---- the reference on the same GPU
    t0 = time.perf_counter()
    ss = S.StackSearch(sci, var, psfs, times, c["nb"])
    r_ctor = time.perf_counter() - t0
    bench.configure(ss, c)
    cands = [S.Trajectory(vx=float(a), vy=float(b)) for a, b in zip(vx, vy)]
    t0 = time.perf_counter(); ss.search_all(cands, True)
    ref_cold = 1e3 * (time.perf_counter() - t0)
    ss.preload_psi_phi_array()
    ref_ms = []
    for _ in range(reps):
        t0 = time.perf_counter(); ss.search_all(cands, True); ref_ms.append(1e3 * (time.perf_counter() - t0))
    want = ref_results_array(S, ss)
    ss.unload_psi_phi_array()

    log(f"{name} full frame {T}x{H}x{W}, {len(vx)} candidates, {evals:.3e} evals, mode={mode}")
    log(f"  ctor (psi/phi generation incl. copies): ours {t_ctor:.2f} s, reference {r_ctor:.2f} s")
    log(f"  search_all + results on host, psi/phi preloaded: ours {min(ours_ms):.1f} ms ({evals / min(ours_ms) * 1e3:.3e} evals/s), "
        f"reference GPU {min(ref_ms):.1f} ms ({evals / min(ref_ms) * 1e3:.3e} evals/s) -> {min(ref_ms) / min(ours_ms):.1f}x")
    log(f"  search_all, psi/phi uploaded inside the call: ours {ours_cold:.1f} ms, reference GPU {ref_cold:.1f} ms "
        f"-> {ref_cold / ours_cold:.1f}x")
    log(f"  results: ours {len(got)}, reference {len(want)}")

    arr = s.psi_phi_array
    nb = arr.num_bytes
    ok = True
    if mode == "identity":
        ra = S.PsiPhiArray()
        S.fill_psi_phi_array_from_image_arrays(ra, c["nb"], sci, var, psfs, times)
        same = np.array_equal(Hn.psi_phi_bytes(ra), arr.host_array().view(np.uint8).ravel())
        log(f"  psi/phi bytes identical to the reference's (built on its GPU path): {same}")
        ok &= bool(same)
        del ra
        pp = (arr.psi_min_val, arr.psi_max_val, arr.psi_scale) if nb != 4 else (0, 0, 1)
        fp = (arr.phi_min_val, arr.phi_max_val, arr.phi_scale) if nb != 4 else (0, 0, 1)
        ost = orc.PsiPhiStack(arr.host_array(), stack.zeroed_times, T, H, W, nb, pp, fp)
        sg = (0.25, 0.75, 0.7413) if c["sigmag"] else None
        p = orc.make_params(c["min_obs"], c["min_lh"], sg, None, c["K"], ost)
        try:
            assert_lists_equal_modulo_ties(got, want, orc, ost, p, f"{name} vs reference GPU")
            ident = ((got["x"] == want["x"]) & (got["y"] == want["y"]) & (bits(got["vx"]) == bits(want["vx"]))
                     & (bits(got["vy"]) == bits(want["vy"])) & (bits(got["flux"]) == bits(want["flux"]))
                     & (got["obs_count"] == want["obs_count"]))
            log(f"  ranked list identical to the reference GPU's modulo exact lh ties: lh sequence bit-exact, "
                f"{int(ident.sum())}/{len(got)} records identical in place, every record re-evaluated exactly")
        except AssertionError as e:
            ok = False
            log(f"  MISMATCH: {e}")
    else:
        def keyed(a):
            k = np.stack([a["x"].astype(np.int64), a["y"].astype(np.int64), bits(a["vx"]).astype(np.int64),
                          bits(a["vy"]).astype(np.int64)], axis=1)
            o = np.lexsort(k.T[::-1])
            return k[o], a[o]
        kg, g = keyed(got); kw, w = keyed(want)
        sg_, sw_ = set(map(tuple, kg)), set(map(tuple, kw))
        both = sg_ & sw_
        log(f"  records in both lists {len(both)}, only ours {len(sg_ - sw_)}, only reference {len(sw_ - sg_)}")
        mg = np.array([tuple(k) in both for k in kg]); mw = np.array([tuple(k) in both for k in kw])
        g, w = g[mg], w[mw]
        if len(g) == len(w) and len(g):
            rl = np.abs(g["lh"] - w["lh"]) / np.maximum(np.abs(w["lh"]), 1e-30)
            rf = np.abs(g["flux"] - w["flux"]) / np.maximum(np.abs(w["flux"]), 1e-30)
            log(f"  common records: max rel diff lh {rl.max():.2e}, flux {rf.max():.2e}, obs_count equal "
                f"{bool(np.array_equal(g['obs_count'], w['obs_count']))}")
            ok &= bool(rl.max() <= 1e-4 and rf.max() <= 1e-4)
    return ok


if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    mode = sys.argv[3] if len(sys.argv) > 3 else "identity"
    sys.exit(0 if run(name, reps, mode) else 1)
