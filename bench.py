#!/usr/bin/env python
"""bench.py -- headline benchmark of the shift-and-stack search (BASELINE.json metric:
pixel-trajectory-observation evaluations per second).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C2|C3u16|C3u8|C4|C5|C1]

Workload at N=1 (BASELINE.json configs[1], "C2"): FakeDataSet 20 x 2048 x 2048 float32 psi/phi (seed 101, 1 % masks,
7x7 Gaussian PSF, 20 injected movers), VelocityGridSearch 50 x 40 = 2000 candidates, sigma-G filter on
(percentiles 25/75, coeff 0.7413, min_lh 10), min_obs 10, results_per_pixel 8.  A "step" is one search_all over
every start pixel.  For N > 1 (weak scaling) the image is N times taller (2048 x 2048*N), rank r searches the
start-pixel row band [2048 r, 2048 (r+1)) and builds psi/phi only for the image rows that band can reach; the
per-GPU sorted results are merged by an NCCL all-gather + rank merge inside the timed step.

value  : whole-job evals/s with psi/phi resident in HBM when the timed region starts (preloaded stack).
e2e    : the same metric through the public API with host buffers: every step uploads the psi/phi stack from the
         (pinned) host copy -- search_all on a StackSearch whose device copy was dropped with
         unload_psi_phi_array(), exactly what the reference does in PsiPhiArray::move_to_gpu -- and reads the
         filtered, sorted result list back to the host.
e2e_with_ctor / e2e_first_search (flat keys): the one-shot pipeline as KBMOD drives the API -- StackSearch(host
         images) + search_all + get_results, and the search + read part of it.
roofline: the gather kernel against the shared-memory load peak, measured in this process by kb_measure_peaks
         together with the L2-resident gather and HBM streaming figures (flat keys frac_* / peak_*).
parity / reference_gpu (N = 1): the unmodified reference's own CUDA search_all from oracle/_ref on the same GPU --
         result lists identical modulo exact lh ties, and its time.  merge_check (N > 1): the NCCL-merged list against
         one GPU searching the whole frame.  named_config (N > 1): BASELINE.json's C4, strong-scaled, ctor included.
other_configs (N = 1, default run): C3 (uint16, uint8), C1, C4 and one band of C5, each constructed, searched once
         and timed for a few resident steps; their headline numbers also as flat keys (c4_value, c5_band_ms_per_step ...).
The stack (671 MB at N=1) is larger than the 126 MB L2, so no L2 flush is needed between iterations.
Only the cpu_baseline, parity / reference_gpu legs and --impl reference touch oracle/ (the checker and the
reference's own code), all outside the timed regions of our arm.
"""
import argparse
import json
import math
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pixel-trajectory-obs evals/sec"
UNIT = "evals/s"

CONFIGS = {
    # name: (T, H, W, psf_std, mask_frac, generator args, num_bytes, sigmag, min_obs, min_lh, K, obs_per_day)
    "C2": dict(T=20, H=2048, W=2048, psf=1.0, mask=0.01, grid=(50, -40, 40, 40, -32, 32), nb=-1, sigmag=True,
               min_obs=10, min_lh=10.0, K=8, obs_per_day=4, seed=101, n_obj=20),
    "C3u16": dict(T=20, H=2048, W=2048, psf=1.0, mask=0.01, grid=(50, -40, 40, 40, -32, 32), nb=2, sigmag=True,
                  min_obs=10, min_lh=10.0, K=8, obs_per_day=4, seed=101, n_obj=20),
    "C3u8": dict(T=20, H=2048, W=2048, psf=1.0, mask=0.01, grid=(50, -40, 40, 40, -32, 32), nb=1, sigmag=True,
                 min_obs=10, min_lh=10.0, K=8, obs_per_day=4, seed=101, n_obj=20),
    # C4: BASELINE.json configs[3] (psi/phi generation reported as ctor_seconds); C5: configs[4] is an 8-GPU
    # config -- on one GPU the bench searches ONE of its 8 start-pixel row bands ("band_div": 8) of the full stack.
    "C4": dict(T=100, H=4096, W=4096, psf=1.0, mask=0.01, ecliptic=([0.5, 8.0, 200], [-math.pi / 15, math.pi / 15, 100]),
               nb=-1, sigmag=True, min_obs=50, min_lh=10.0, K=8, obs_per_day=4, seed=102, n_obj=20),
    "C5": dict(T=400, H=2048, W=2048, psf=1.0, mask=0.0, grid=(400, -4.0, 4.0, 250, -2.5, 2.5), nb=-1, sigmag=False,
               min_obs=200, min_lh=10.0, K=8, obs_per_day=8, intra=0.005, seed=103, n_obj=20, band_div=8),
    "C1": dict(T=10, H=512, W=512, psf=0.5, mask=0.0, kbmodv1=(5, 0, 4, 5, -0.1, 0.1), nb=-1, sigmag=False,
               min_obs=7, min_lh=0.0, K=8, obs_per_day=1, seed=0, n_obj=1),
}


def workload_name(cfg_name, c, n_gpus):
    if "grid" in c:
        gen = f"VelocityGridSearch {c['grid'][0]}x{c['grid'][3]} ({c['grid'][0] * c['grid'][3]} candidates)"
    elif "ecliptic" in c:
        gen = f"EclipticCenteredSearch {c['ecliptic'][0][2]}x{c['ecliptic'][1][2]} ({c['ecliptic'][0][2] * c['ecliptic'][1][2]} candidates)"
    else:
        gen = "KBMODV1Search 5x5 (25 candidates)"
    enc = {-1: "float32", 2: "uint16", 1: "uint8"}[c["nb"]]
    return (f"{cfg_name}: FakeDataSet {c['T']}x{c['H'] * n_gpus}x{c['W']} {enc} psi/phi, {gen}, "
            f"sigma-G {'on' if c['sigmag'] else 'off'}, min_obs {c['min_obs']}, results_per_pixel {c['K']}"
            + (f", one of {c['band_div']} start-pixel row bands per GPU" if c.get("band_div") else ""))


def host_threads():
    """Host threads this process may run on (its CPU affinity mask, not the machine's core count)."""
    try:
        return max(len(os.sched_getaffinity(0)), 1)
    except AttributeError:
        return os.cpu_count() or 1


def loaded_native_libraries():
    """Shared objects of this repository that are mapped into the process (from /proc/self/maps)."""
    out = set()
    try:
        for ln in open("/proc/self/maps"):
            path = ln.split()[-1] if "/" in ln else ""
            if path.endswith(".so") and path.startswith(ROOT):
                out.add(os.path.relpath(path, ROOT))
    except OSError:
        pass
    return sorted(out)


def make_inputs(c, n_gpus):
    # host-only modules: neither loads libkbmod_b200.so, so the reference arm's process never maps our library
    from kbmod_b200 import fake_data as fd
    from kbmod_b200.trajectory import Trajectory
    times = fd.create_fake_times(c["T"], 57130.2, obs_per_day=c["obs_per_day"], intra_night_gap=c.get("intra", 0.01))
    ds = fd.FakeDataSet(c["W"], c["H"] * n_gpus, times, mask_fraction=c["mask"], psf_val=c["psf"], use_seed=c["seed"])
    if "grid" in c or "ecliptic" in c:
        gen = fd.VelocityGridSearch(*c["grid"]) if "grid" in c else fd.EclipticCenteredSearch(*c["ecliptic"])
        ds.insert_random_objects_from_generator(c["n_obj"] * n_gpus, gen, 500.0)
    else:
        gen = fd.KBMODV1Search(*c["kbmodv1"])
        ds.insert_object(Trajectory(x=2, y=0, vx=10.7, vy=15.3, flux=275.0))  # README.md:104
    vx, vy = gen.to_arrays()
    ds.stack_py.objects = list(ds.trajectories)  # the injected movers (for window-parity checks)
    return ds.stack_py, vx, vy


class BandStack:
    pass


def make_band_inputs(c, need_rows):
    """Inputs of config `c` for the image rows [need_rows[0], need_rows[1]) only (what one rank of a strong-scaling run
    reads: its band of start rows + the candidates' reach + the PSF radius).  Every (image, 64-row block) has its own
    seeded generator and the movers come from one shared generator, so ranks that need different rows see the same
    frames without any of them generating a whole 13 GB stack.  The arrays are full-size (the C-ABI takes whole
    images); rows outside `need_rows` are never written, and never read."""
    from kbmod_b200 import fake_data as fd
    T, H, W, noise = c["T"], c["H"], c["W"], 2.0
    times = np.asarray(fd.create_fake_times(T, 57130.2, obs_per_day=c["obs_per_day"],
                                            intra_night_gap=c.get("intra", 0.01)), dtype=float)
    zeroed = times - times[0]
    psf = fd.make_gaussian_kernel(c["psf"])
    gen = fd.VelocityGridSearch(*c["grid"]) if "grid" in c else fd.EclipticCenteredSearch(*c["ecliptic"])
    vx, vy = gen.to_arrays()
    lo, hi = max(0, need_rows[0]), min(H, need_rows[1])
    B = 64
    sci = [np.empty((H, W), dtype=np.float32) for _ in range(T)]
    var = [np.empty((H, W), dtype=np.float32) for _ in range(T)]
    for t in range(T):
        for b in range(lo // B, (hi + B - 1) // B):
            r0, r1 = b * B, min((b + 1) * B, H)
            rng = np.random.default_rng([c["seed"], t, b])
            blk = rng.standard_normal((r1 - r0, W), dtype=np.float32) * np.float32(noise)
            vblk = np.full((r1 - r0, W), noise * noise, dtype=np.float32)
            if c["mask"] > 0:
                m = rng.random((r1 - r0, W), dtype=np.float32) < c["mask"]
                blk[m] = np.nan
                vblk[m] = np.nan
            a, e = max(r0, lo), min(r1, hi)
            sci[t][a:e] = blk[a - r0:e - r0]
            var[t][a:e] = vblk[a - r0:e - r0]
    orng = np.random.default_rng([c["seed"], 9999])
    rad = psf.shape[0] // 2
    objects = []
    for _ in range(c["n_obj"]):
        k = int(orng.integers(0, len(vx)))
        ox, oy = int(orng.random() * W), int(orng.random() * H)
        objects.append((ox, oy, float(vx[k]), float(vy[k])))
        for t in range(T):
            x, y = int(ox + vx[k] * zeroed[t] + 0.5), int(oy + vy[k] * zeroed[t] + 0.5)
            for dy in range(-rad, rad + 1):
                yy = y + dy
                if yy < lo or yy >= hi:
                    continue
                x0, x1 = max(0, x - rad), min(W, x + rad + 1)
                if x0 < x1:
                    sci[t][yy, x0:x1] += np.float32(500.0) * psf[dy + rad, x0 - (x - rad):x1 - (x - rad)]
    st = BandStack()
    st.sci, st.var, st.psfs, st.zeroed_times, st.objects = sci, var, [psf] * T, zeroed, objects
    return st, vx, vy


def configure(s, c, band=None):
    s.set_min_obs(c["min_obs"])
    s.set_results_per_pixel(c["K"])
    if c["sigmag"]:
        s.enable_gpu_sigmag_filter([0.25, 0.75], 0.7413, c["min_lh"])
    else:
        s.set_min_lh(c["min_lh"])
    if band is not None:
        s.set_start_bounds_y(band[0], band[1])


# ---- clocks during the timed region ---------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop = [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.th = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        if self.nv:
            self.th.start()

    def stop(self):
        self._stop.set()
        if self.nv and self.th.is_alive():
            self.th.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---- the reference's own CPU search_all on a bounded sample (cpu_baseline / --impl reference) -----------------
def reference_cpu_setup(c, stack, vx, vy, target_seconds):
    """Build the reference StackSearch on a crop of the same synthetic stack and choose a start-pixel window so
    one search_all(use_gpu=False) takes about `target_seconds` on this box's cores."""
    from oracle import ref
    S, H = ref.load()
    # torchrun exports OMP_NUM_THREADS=1 to its workers; the reference's CPU search is an OpenMP loop and gets every
    # host thread this process may use.  `cores` is what OpenMP reports afterwards, not a guess.
    H.set_omp_threads(host_threads())
    cores = int(H.omp_threads())
    n_cand, T = len(vx), c["T"]
    tmax = float(stack.zeroed_times[-1])
    reach = int(np.ceil(max(np.abs(vx).max(), np.abs(vy).max()) * tmax)) + 1
    # evals/s of the reference on these cores: measured on a small probe window of a small crop, not assumed
    rate = 3.0e7 * cores
    try:
        pside = min(min(c["H"], c["W"]), 48 + 2 * reach)
        pw = max(8, min(48, pside))
        plo = max((pside - pw) // 2, 0)
        ps = S.StackSearch([np.ascontiguousarray(a[:pside, :pside]) for a in stack.sci],
                           [np.ascontiguousarray(a[:pside, :pside]) for a in stack.var], list(stack.psfs),
                           [float(t) for t in stack.zeroed_times], c["nb"])
        ps.set_min_obs(c["min_obs"])
        ps.set_min_lh(c["min_lh"])
        ps.set_results_per_pixel(c["K"])
        ps.set_start_bounds_x(plo, min(plo + pw, pside))
        ps.set_start_bounds_y(plo, min(plo + pw, pside))
        pc = [S.Trajectory(vx=float(a), vy=float(b)) for a, b in zip(vx, vy)]
        pwx = min(plo + pw, pside) - plo
        ps.search_all(pc, False)
        t0 = time.perf_counter()
        ps.search_all(pc, False)
        rate = max(pwx * pwx * n_cand * T / max(time.perf_counter() - t0, 1e-4), 1e6)
        del ps
    except Exception:
        pass
    w = int(np.sqrt(max(target_seconds * rate / (n_cand * T), 16.0 * 16.0)))
    w = max(16, min(w, min(c["H"], c["W"]) - 2 * reach if min(c["H"], c["W"]) > 2 * reach + 64 else w))
    side = min(min(c["H"], c["W"]), w + 2 * reach)
    lo = max((side - w) // 2, 0)
    w = min(w, side)
    sci = [np.ascontiguousarray(a[:side, :side]) for a in stack.sci]
    var = [np.ascontiguousarray(a[:side, :side]) for a in stack.var]
    ss = S.StackSearch(sci, var, list(stack.psfs), [float(t) for t in stack.zeroed_times], c["nb"])
    ss.set_min_obs(c["min_obs"])
    ss.set_min_lh(c["min_lh"])
    ss.set_results_per_pixel(c["K"])
    ss.set_start_bounds_x(lo, min(lo + w, side))
    ss.set_start_bounds_y(lo, min(lo + w, side))
    wx = min(lo + w, side) - lo
    cands = [S.Trajectory(vx=float(a), vy=float(b)) for a, b in zip(vx, vy)]
    evals = wx * wx * n_cand * T
    sample = (f"reference kbmod search_all(use_gpu=False) (OpenMP, {cores} host threads) on a {wx}x{wx} start-pixel "
              f"window of a {side}x{side} crop of the same stack, all {n_cand} candidates x {T} images "
              f"({evals:.3e} evals); the CPU path has no sigma-G and filters after top-K")
    return ss, cands, evals, cores, sample


def run_reference(args, c, cfg_name):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    stack, vx, vy = make_inputs(c, 1)
    ss, cands, evals, cores, sample = reference_cpu_setup(c, stack, vx, vy, target_seconds=4.0)
    for _ in range(args.warmup):
        ss.search_all(cands, False)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ss.search_all(cands, False)
    dt = time.perf_counter() - t0
    value = evals * args.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(cfg_name, c, 1), "timed_on": "host cores", "sample": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "native_so_loaded": loaded_native_libraries()}
    print(json.dumps(line), flush=True)


def measure_peaks():
    """Roofline denominators measured on this device, in this process (kb_measure_peaks), in GB/s."""
    import ctypes as C
    from kbmod_b200._lib import check, lib
    buf = (C.c_double * 6)()
    check(lib.kb_measure_peaks(buf, C.c_int(6)))
    names = ("hbm_stream_read", "l2_resident_read", "l2_row_gather_256B", "smem_lds64", "smem_lds32", "smem_lds16")
    return dict(zip(names, [float(v) for v in buf]))


def results_digest(arr):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(arr).view(np.uint8).tobytes()).hexdigest()[:16]


def reference_gpu_leg(c, stack, vx, vy):
    """The checker, outside every timed region of our arm: the UNMODIFIED reference's own CUDA search_all from
    oracle/_ref on this GPU, same workload.  Returns (parity record, timing record).  Parity is taken on the pair of
    stacks both sides can build bit-identically (science images convolved beforehand + a 1x1 PSF: the reference's
    FFMA-contracted GPU convolution then equals its CPU convolution, which is what we reproduce)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from parity import assert_lists_equal_modulo_ties
    from kbmod_b200 import search as kb
    from kbmod_b200._lib import TRAJECTORY_DTYPE
    from oracle import cpu as orc, ref as oref
    S, Hn = oref.load()
    if not S.kb_has_gpu():
        raise RuntimeError("the reference build sees no GPU")
    T, H, W = c["T"], c["H"], c["W"]
    times = [float(t) for t in stack.zeroed_times]
    sci = [kb.convolve_image(np.ascontiguousarray(a, dtype=np.float32), p) for a, p in zip(stack.sci, stack.psfs)]
    one = [np.ones((1, 1), dtype=np.float32)] * T
    var = list(stack.var)
    evals = H * W * len(vx) * T
    ss = S.StackSearch(sci, var, one, times, c["nb"])
    configure(ss, c)
    cands = [S.Trajectory(vx=float(a), vy=float(b)) for a, b in zip(vx, vy)]
    t0 = time.perf_counter()
    ss.search_all(cands, True)                       # uploads psi/phi inside the call, like our e2e leg
    ref_e2e_ms = 1e3 * (time.perf_counter() - t0)
    ss.preload_psi_phi_array()
    ref_ms = []
    for _ in range(2):
        t0 = time.perf_counter()
        ss.search_all(cands, True)
        ref_ms.append(1e3 * (time.perf_counter() - t0))
    n = ss.get_number_total_results()
    want = np.zeros(n, dtype=TRAJECTORY_DTYPE)
    if n:
        lst = ss.get_results(0, n)
        for f in ("x", "y", "vx", "vy", "lh", "flux", "obs_count"):
            want[f] = getattr(S, "extract_all_trajectory_" + f)(lst)
    ss.unload_psi_phi_array()
    s = kb.StackSearch(sci, var, one, times, c["nb"])
    configure(s, c)
    s.search_all((vx, vy), True)
    got = s.get_results_array()
    arr = s.psi_phi_array
    ra = S.PsiPhiArray()
    S.fill_psi_phi_array_from_image_arrays(ra, c["nb"], sci, var, one, times)
    same_bytes = bool(np.array_equal(Hn.psi_phi_bytes(ra), arr.host_array().view(np.uint8).ravel()))
    del ra, ss
    nb = arr.num_bytes
    pp = (arr.psi_min_val, arr.psi_max_val, arr.psi_scale) if nb != 4 else (0, 0, 1)
    fp = (arr.phi_min_val, arr.phi_max_val, arr.phi_scale) if nb != 4 else (0, 0, 1)
    ost = orc.PsiPhiStack(arr.host_array(), stack.zeroed_times, T, H, W, nb, pp, fp)
    prm = orc.make_params(c["min_obs"], c["min_lh"], (0.25, 0.75, 0.7413) if c["sigmag"] else None, None, c["K"], ost)
    parity = {"vs": "reference_gpu", "what": "oracle/_ref StackSearch.search_all(candidates, True), full frame, same GPU",
              "psi_phi_bytes_identical": same_bytes, "n": int(len(got)), "n_reference": int(len(want))}
    try:
        assert_lists_equal_modulo_ties(got, want, orc, ost, prm, "bench parity leg")
        parity["identical_modulo_ties"] = bool(same_bytes)
    except AssertionError as e:
        parity["identical_modulo_ties"] = False
        parity["mismatch"] = str(e)[:300]
    timing = {"ms_per_search": min(ref_ms), "value": evals / (min(ref_ms) * 1e-3), "unit": UNIT,
              "e2e_ms": ref_e2e_ms, "e2e_value": evals / (ref_e2e_ms * 1e-3),
              "what": "the reference's own kernels.cu (built for sm_100a) through its StackSearch on this GPU: "
                      "psi/phi preloaded (ms_per_search, best of 2) and uploaded inside the call (e2e_ms)"}
    return parity, timing


def named_config_leg(kb, dist, comm, rank, world, c, name):
    """BASELINE.json's own multi-GPU workload next to the weak-scaling curve: the named frame, its start-pixel rows
    split over the ranks (strong scaling), each rank building psi/phi for its own rows only; construction, two
    searches with the NCCL merge and the read-back of the merged list, timed from the host images on."""
    H = c["H"]
    band = tuple(dist.row_bands(0, H, world)[rank])
    import kbmod_b200.fake_data as fd
    times = np.asarray(fd.create_fake_times(c["T"], 57130.2, obs_per_day=c["obs_per_day"],
                                            intra_night_gap=c.get("intra", 0.01)), dtype=float)
    gen = fd.VelocityGridSearch(*c["grid"]) if "grid" in c else fd.EclipticCenteredSearch(*c["ecliptic"])
    vy_all = gen.to_arrays()[1]
    rows = dist.band_rows(band, vy_all, times - times[0], H)
    rad = fd.make_gaussian_kernel(c["psf"]).shape[0] // 2
    stack, vx, vy = make_band_inputs(c, (rows[0] - rad, rows[1] + rad))
    comm.barrier()
    t0 = time.perf_counter()
    s = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, c["nb"], rows=rows)
    ctor_s = time.perf_counter() - t0
    configure(s, c, band)
    s.search_all((vx, vy), True)
    dist.merge_results(s, 0)
    n = len(s.get_results_array())
    first_ms = 1e3 * (time.perf_counter() - t0)
    comm.barrier()
    steps = 2
    s.timer_start()
    k_ms = 0.0
    for _ in range(steps):
        s.search_all((vx, vy), True)
        dist.merge_results(s, 0)
        k_ms += s.last_timing()["ms_main_kernel"]
    ms = comm.max_float(s.timer_stop() / steps)
    evals = c["W"] * H * len(vx) * c["T"]
    rec = {"workload": workload_name(name, c, 1), "scaling": "strong", "n_gpus": world, "steps": steps,
           "ms_per_step": ms, "value": evals / (ms * 1e-3), "unit": UNIT,
           "gather_kernel_ms_slowest_rank": comm.max_float(k_ms / steps),
           "ctor_seconds_slowest_rank": comm.max_float(ctor_s),
           "ctor_plus_first_search_ms": comm.max_float(first_ms),
           "ctor_plus_first_search_value": evals / (comm.max_float(first_ms) * 1e-3),
           "results": int(n), "rows_per_rank": int(rows[1] - rows[0]),
           "note": "every rank builds psi/phi for its band of start rows plus the candidates' reach only; inputs are "
                   "generated per rank for those rows (seeded per 64-row block, same frames on every rank)"}
    del s
    return rec


def other_config_leg(kb, dist, name, shared=None):
    """One more BASELINE.json config measured in the same single-GPU run (the driver runs only the default line):
    construction from host images, one search + read-back, then a few timed resident steps.  C3 reuses C2's images
    (they differ in the encoding only); C4 / C5 generate theirs block-seeded (make_band_inputs); C5, an 8-GPU config,
    is measured as one of its 8 start-pixel row bands, with psi/phi built for the rows that band can reach."""
    c = CONFIGS[name]
    band = rows = None
    if shared is not None:
        stack, vx, vy = shared
    elif "kbmodv1" in c:
        stack, vx, vy = make_inputs(c, 1)
    else:
        import kbmod_b200.fake_data as fd
        need = (0, c["H"])
        if c.get("band_div"):
            n = c["H"] // c["band_div"]
            band = (c["H"] // 2 - n // 2, c["H"] // 2 - n // 2 + n)
            times = np.asarray(fd.create_fake_times(c["T"], 57130.2, obs_per_day=c["obs_per_day"],
                                                    intra_night_gap=c.get("intra", 0.01)), dtype=float)
            gen = fd.VelocityGridSearch(*c["grid"]) if "grid" in c else fd.EclipticCenteredSearch(*c["ecliptic"])
            rows = dist.band_rows(band, gen.to_arrays()[1], times - times[0], c["H"])
            rad = fd.make_gaussian_kernel(c["psf"]).shape[0] // 2
            need = (rows[0] - rad, rows[1] + rad)
        stack, vx, vy = make_band_inputs(c, need)
    t0 = time.perf_counter()
    s = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, c["nb"], rows=rows)
    ctor_s = time.perf_counter() - t0
    configure(s, c, band)
    s.search_all((vx, vy), True)
    n_res = len(s.get_results_array())
    first_ms = 1e3 * (time.perf_counter() - t0)
    heavy = c["T"] >= 100
    steps = 2 if heavy else 5
    for _ in range(1 if heavy else 3):
        s.search_all((vx, vy), True)
    s.timer_start()
    k_ms = 0.0
    for _ in range(steps):
        s.search_all((vx, vy), True)
        k_ms += s.last_timing()["ms_main_kernel"]
    ms = s.timer_stop() / steps
    tiled = s.last_timing()["used_tiled"]
    start_rows = (band[1] - band[0]) if band else c["H"]
    evals = c["W"] * start_rows * len(vx) * c["T"]
    rec = {"workload": workload_name(name, c, 1), "steps": steps, "ms_per_step": ms, "value": evals / (ms * 1e-3),
           "unit": UNIT, "gather_kernel_ms": k_ms / steps, "ctor_seconds": ctor_s,
           "ctor_plus_first_search_ms": first_ms, "results": int(n_res),
           "kernel": ("generic", "tiled", "candidate-parallel")[tiled]}
    del s
    return rec


def run_ours(args, c, cfg_name):
    from kbmod_b200 import search as kb
    from kbmod_b200 import dist

    rank, world, local_rank = dist.env_rank()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    dist.set_device(local_rank)
    # One process per GPU: keep this rank's host buffers (and the threads that fill them) on its GPU's own socket.
    bound = dist.bind_to_device_cpus() if world > 1 and not os.environ.get("KB_NO_CPU_BINDING") else None
    comm = dist.Comm(rank, world) if world > 1 else None

    strong = args.scaling == "strong" and world > 1
    frames = 1 if strong else world            # strong: the named frame itself, start-pixel rows split over the ranks
    stack, vx, vy = make_inputs(c, frames)
    if strong:
        band = tuple(dist.row_bands(0, c["H"], world)[rank])
        rows = c["H"] / world                  # per-rank average (the job total is what is reported)
    else:
        band = (rank * c["H"], (rank + 1) * c["H"])
        rows = c["H"]
    if c.get("band_div") and not strong:  # an N-GPU config measured band by band: this GPU takes a middle band of its frame
        rows = c["H"] // c["band_div"]
        mid = rank * c["H"] + c["H"] // 2
        band = (mid - rows // 2, mid - rows // 2 + rows)
    use_band = band if (world > 1 or c.get("band_div")) else None

    # ---- one-shot pipeline, as KBMOD drives the API (run_search.py:363-378): construct, search once, read results --
    # Timed on the host clock from the first byte of the host images to the result list on the host; the device is
    # idle before and after.  `e2e_first_search` is the search + read part of the same pass.  The PROCESS is warm
    # (CUDA context, the library's device code, the pinned staging ring: one 64 x 64 search on other data first);
    # the stack, the handle and every buffer sized by it are cold.
    if True:
        wrng = np.random.default_rng(1)
        wsci = [wrng.normal(0, 2, (64, 64)).astype(np.float32) for _ in range(c["T"])]
        wvar = [np.full((64, 64), 4.0, dtype=np.float32) for _ in range(c["T"])]
        ws = kb.StackSearch(wsci, wvar, list(stack.psfs), stack.zeroed_times, c["nb"])
        configure(ws, c)
        ws.set_kernel_mode(2)
        ws.search_all((vx[:64], vy[:64]), True)
        ws.get_results_array()
        del ws
    # N > 1: every rank builds psi/phi only for the image rows its band of start rows can reach (float stacks; the
    # scale of an encoded stack depends on every row, so those are built whole on every rank)
    rows_held = None
    if world > 1 and c["nb"] not in (1, 2):
        rows_held = dist.band_rows(band, vy, stack.zeroed_times, c["H"] * frames)
    t0 = time.perf_counter()
    s = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, c["nb"], rows=rows_held)
    ctor_s = time.perf_counter() - t0
    configure(s, c, use_band)
    t1 = time.perf_counter()
    s.search_all((vx, vy), True)
    first = s.get_results_array()
    t2 = time.perf_counter()
    first_tm = s.last_timing()
    first_ms, with_ctor_ms = 1e3 * (t2 - t1), 1e3 * (t2 - t0)
    first_digest = results_digest(first)
    evals_rank = c["W"] * (band[1] - band[0] if strong else rows) * len(vx) * c["T"]
    evals_job = c["W"] * c["H"] * len(vx) * c["T"] if strong else evals_rank * world
    stack_bytes = s.psi_phi_array.total_array_size
    top_n = 0  # merge everything that survives the filter

    def step():
        s.search_all((vx, vy), True)
        if comm is not None:
            dist.merge_results(s, top_n)
        return s.last_timing()

    # ---- resident leg: psi/phi preloaded ---------------------------------------------------------------------
    s.preload_psi_phi_array()
    for _ in range(max(args.warmup, 3)):
        step()
    sampler = ClockSampler(local_rank)
    if comm is not None:
        comm.barrier()
    sampler.start()
    s.timer_start()
    k_ms, launches, n_kernel, tiled = 0.0, 0, 0, 0
    for _ in range(args.steps):
        tm = step()
        k_ms += tm["ms_main_kernel"]   # the one gather kernel of this step, CUDA events on its own stream
        n_kernel += 1
        launches += tm["total_launches"]
        tiled = tm["used_tiled"]
    ms_dev = s.timer_stop()
    clocks = sampler.stop()
    if comm is not None:
        ms_dev = comm.max_float(ms_dev)
        k_ms = comm.max_float(k_ms)
    n_results = s.get_number_total_results()
    resident_digest = results_digest(s.get_results_array())
    ms_step = ms_dev / args.steps
    value = evals_job / (ms_step * 1e-3)

    # ---- end-to-end leg: host buffers, stack uploaded every step, results read back ------------------------------
    s.unload_psi_phi_array()
    e2e_steps = max(1, min(args.steps, 5))
    for _ in range(2):
        step()
        s.get_results_array()
    if comm is not None:
        comm.barrier()
    d2h = 0
    h2d = 0
    t0 = time.perf_counter()
    s.timer_start()
    for _ in range(e2e_steps):
        tm = step()
        h2d = tm["h2d_stack_bytes"]   # the rows this rank's band can reach (the whole stack for a full-frame search)
        res = s.get_results_array()
        d2h = res.nbytes
    e2e_tm = tm
    ms_e2e_dev = s.timer_stop()
    wall_e2e = (time.perf_counter() - t0) * 1e3
    ms_e2e = max(ms_e2e_dev, wall_e2e) / e2e_steps  # results are read after the last kernel: use the wall clock
    if comm is not None:
        ms_e2e = comm.max_float(ms_e2e)
        first_ms = comm.max_float(first_ms)
        with_ctor_ms = comm.max_float(with_ctor_ms)
    e2e_value = evals_job / (ms_e2e * 1e-3)

    # ---- roofline of the dominant kernel (the gather kernel), CUDA events inside the library around that launch -----
    # The gathers are served from shared memory (the stack is staged by TMA), so the binding roof is the SM's
    # shared-memory pipe; its peak for this encoding's load width, the L2-resident gather figure and the HBM
    # streaming figure are measured here, in this process, by kb_measure_peaks.
    pk = measure_peaks()
    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_copy = float(json.load(open(peaks_file))["hbm_gbs"]) if os.path.exists(peaks_file) else 6650.0
    hbm_src = "MEASURED_PEAKS.json hbm_gbs (measured copy)" if os.path.exists(peaks_file) else "fallback (B200_PROFILING.md)"
    nbytes = {-1: 4, 4: 4, 2: 2, 1: 1}[c["nb"]]
    bytes_per_eval = 2 * nbytes
    # whatever the stack's encoding, the gather loop reads one 32-bit word of the 16-bit fixed-point screening copy per
    # evaluation (LDS.32): that is the load whose measured peak bounds it
    smem_key, actual_bytes_per_eval = "smem_lds32", 4
    kernel_ms = k_ms / max(n_kernel, 1)
    achieved = evals_rank * bytes_per_eval / (kernel_ms * 1e-3) / 1e9
    traffic_tab = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    ncu = json.load(open(traffic_tab)).get(cfg_name) if os.path.exists(traffic_tab) else None
    kname = "search_tiled_kernel" if tiled else "search_generic_kernel"
    if ncu is not None and kname not in ncu["kernel"]:
        ncu = None
    bound_peak = pk[smem_key] if tiled else pk["l2_row_gather_256B"]
    roofline = {"bound": "shared_memory" if tiled else "l2", "achieved": achieved, "peak": bound_peak, "unit": "GB/s",
                "frac": achieved / bound_peak,
                "traffic": ncu["dram_bytes"] if ncu else None, "traffic_unit": "bytes/launch (dram read + write)",
                "traffic_source": ncu["source"] if ncu else None,
                "kernel": kname, "kernel_ms": kernel_ms, "bytes_per_eval": bytes_per_eval,
                "actual_smem_bytes_per_eval": actual_bytes_per_eval if tiled else None,
                "frac_actual_bytes": (achieved * actual_bytes_per_eval / bytes_per_eval / bound_peak) if tiled else None,
                "peak_source": f"kb_measure_peaks in this run: {smem_key} (shared-memory loads of this encoding's "
                               f"pair width on all SMs)" if tiled else "kb_measure_peaks in this run: L2-resident 256 B row gather",
                "peaks_measured_in_run_gbs": pk, "hbm_copy_peak_gbs": hbm_copy, "hbm_copy_peak_source": hbm_src,
                "note": "achieved = algorithmic gather bytes (P*N*T pairs x bytes per pair of this encoding) / gather-kernel "
                        "time.  The kernel screens every evaluation with a 4-byte word of a 16-bit fixed-point copy of "
                        "the stack (exact integer sums, rigorous bounds) and evaluates exactly only what the bounds "
                        "cannot rule out, so for float32 stacks (8-byte pairs) the algorithmic fraction can exceed 1; "
                        "frac_actual_bytes counts the 4 bytes the loop really reads"}
    flat = {"kernel_ms": kernel_ms, "achieved_gbs": achieved,
            "frac_shared_memory": achieved / pk[smem_key], "peak_shared_memory": pk[smem_key],
            "frac_l2_gather": achieved / pk["l2_row_gather_256B"], "peak_l2_gather": pk["l2_row_gather_256B"],
            "frac_hbm": achieved / hbm_copy, "peak_hbm": hbm_copy,
            "wavefronts_per_gather": (ncu["smem_wavefronts"] / (evals_rank / 32.0)) if ncu else None,
            "e2e_first_search_ms": first_ms, "e2e_first_search": evals_job / (first_ms * 1e-3),
            "e2e_with_ctor_ms": with_ctor_ms, "e2e_with_ctor": evals_job / (with_ctor_ms * 1e-3),
            "ctor_seconds": ctor_s, "storage_dtype": {4: "f32", 2: "u16", 1: "u8"}[nbytes],
            "first_search_device_ms": {k: first_tm[k] for k in ("ms_h2d", "ms_prepare", "ms_search", "ms_filter_sort", "ms_total")},
            "e2e_step_device_ms": {k: e2e_tm[k] for k in ("ms_h2d", "ms_prepare", "ms_search", "ms_filter_sort", "ms_total")}}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong" if strong else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(cfg_name, c, frames), "candidates": int(len(vx)),
                       "start_pixels_per_gpu": int(c["W"] * rows), "evals_per_step": evals_job,
                       "l2": "psi/phi stack (%.0f MB) larger than the 126 MB L2; no flush" % (stack_bytes / 1e6),
                       "parallelism": f"start-pixel row bands x{world}, psi/phi replicated, NCCL all-gather merge"
                       if world > 1 else "single GPU", "results": int(n_results), "ctor_seconds": ctor_s},
            "roofline": roofline,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d + 8 * len(vx)),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e, "steps": e2e_steps,
                    "first_search_h2d_bytes": int(first_tm["h2d_stack_bytes"])},
            "gpu_launches": int(launches), "clocks": clocks}
    line.update(flat)
    # every leg returned the same list (one-shot, resident, and -- below -- the reference's)
    line["legs_identical"] = bool(first_digest == resident_digest) if comm is None else None

    if comm is not None:
        # NCCL-merged list of all ranks against one GPU searching the whole frame alone (rank 0, outside the timing)
        merged = results_digest(s.get_results_array())
        line["ctor_seconds_max_over_ranks"] = comm.max_float(ctor_s)
        line["host_cpus_rank0"] = (f"{len(bound)} CPUs local to the GPU (sysfs local_cpulist)" if bound
                                   else "unbound (nothing narrower than the process mask reported)")
        if rank == 0:
            ymin, ymax = (0, c["H"]) if strong else (0, c["H"] * world)
            if c.get("band_div") and not strong:
                line["merge_check"] = "skipped (bands of a band_div config do not tile the frame)"
            elif args.no_merge_check:
                line["merge_check"] = "skipped (--no-merge-check)"
            else:
                sw_ = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, c["nb"])   # the whole frame
                configure(sw_, c, (ymin, ymax))
                sw_.search_all((vx, vy), True)
                whole = results_digest(sw_.get_results_array())
                line["merge_check"] = "identical" if whole == merged else "differs"
                line["merge_check_n"] = int(sw_.get_number_total_results())
                del sw_
        comm.barrier()
        if args.named_config and args.named_config in CONFIGS and not strong:
            del s
            line["named_config"] = named_config_leg(kb, dist, comm, rank, world, CONFIGS[args.named_config], args.named_config)

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            ss, cands, evals, cores, sample = reference_cpu_setup(c, stack, vx, vy, target_seconds=10.0)
            t0 = time.perf_counter()
            ss.search_all(cands, False)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": evals / dt, "unit": UNIT, "cores": cores, "kind": "reference",
                                    "sample": sample, "seconds": dt}
            del ss
        except Exception as e:  # the reference build did not travel: fall back to the C port of the oracle
            line["cpu_baseline"] = cpu_baseline_port(c, stack, vx, vy, str(e))
    if rank == 0 and world == 1 and not args.no_parity and c["T"] <= 200 and not c.get("band_div"):
        try:
            del s
            line["parity"], line["reference_gpu"] = reference_gpu_leg(c, stack, vx, vy)
            line["vs_reference_gpu"] = value / line["reference_gpu"]["value"]
            line["vs_reference_gpu_e2e"] = e2e_value / line["reference_gpu"]["e2e_value"]
        except Exception as e:
            line["parity"] = {"vs": "reference_gpu", "unavailable": str(e)[:200]}
    if world == 1 and cfg_name == "C2" and args.other_configs:
        # BASELINE.json's other configs in the same driver-run line (flat copies of the headline numbers: a parser
        # that drops nested objects still keeps them)
        others = {}
        for name in [n for n in args.other_configs.split(",") if n in CONFIGS and n != cfg_name]:
            try:
                rec = other_config_leg(kb, dist, name, shared=(stack, vx, vy) if name.startswith("C3") else None)
                key = name.lower() + ("_band" if CONFIGS[name].get("band_div") else "")
                for f in ("value", "ms_per_step", "gather_kernel_ms", "ctor_seconds"):
                    line[f"{key}_{f}"] = rec[f]
            except Exception as e:
                rec = {"error": str(e)[:300]}
            others[name] = rec
        line["other_configs"] = others
    if comm is not None:
        comm.finalize()
    if rank == 0:
        print(json.dumps(line), flush=True)


def cpu_baseline_port(c, stack, vx, vy, why):
    from oracle import cpu as orc
    side, w = 384, 96
    st = orc.fill_from_images([a[:side, :side] for a in stack.sci], [a[:side, :side] for a in stack.var], stack.psfs,
                              stack.zeroed_times, c["nb"])
    lo = (side - w) // 2
    p = orc.make_params(c["min_obs"], c["min_lh"], (0.25, 0.75, 0.7413) if c["sigmag"] else None,
                        (lo, lo + w, lo, lo + w), c["K"], st)
    t0 = time.perf_counter()
    orc.search_slots(st, p, vx, vy)
    dt = time.perf_counter() - t0
    evals = w * w * len(vx) * c["T"]
    return {"value": evals / dt, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
            "sample": f"oracle C port (OpenMP) on a {w}x{w} window of a {side}x{side} crop; reference build "
                      f"unavailable: {why}", "seconds": dt}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the reference-GPU parity / timing leg")
    ap.add_argument("--no-merge-check", action="store_true", help="N > 1: skip rank 0's whole-frame check of the merge")
    ap.add_argument("--named-config", default="C4", help="N > 1 (weak runs): also run this config strong-scaled, "
                    "construction included (BASELINE.json's multi-GPU workload); '' to skip")
    ap.add_argument("--other-configs", default="C3u16,C3u8,C1,C4,C5",
                    help="N = 1, config C2 (the default run): also measure these configs briefly and add them to the "
                         "line; '' to skip")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N > 1: weak = every GPU searches its own copy-sized band of an N-times taller frame (default, "
                         "the driver's scaling run); strong = the named frame, start-pixel rows split over the GPUs "
                         "(BASELINE configs C4 / C5)")
    args = ap.parse_args()
    c = CONFIGS[args.config]
    # torchrun exports OMP_NUM_THREADS=1; our host code (first touch of the host copy, schedule construction) is
    # OpenMP too.  Give every rank its share of the host threads before the library (and its libgomp) is loaded.
    if args.impl == "ours" and os.environ.get("OMP_NUM_THREADS") == "1" and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        os.environ["OMP_NUM_THREADS"] = str(max(host_threads() // int(os.environ.get("LOCAL_WORLD_SIZE", os.environ["WORLD_SIZE"])), 1))
    if args.impl == "reference":
        run_reference(args, c, args.config)
    else:
        run_ours(args, c, args.config)


if __name__ == "__main__":
    main()
