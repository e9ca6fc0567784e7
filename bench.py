#!/usr/bin/env python
"""bench.py -- headline benchmark of the shift-and-stack search (BASELINE.json metric:
pixel-trajectory-observation evaluations per second).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C2|C3u16|C3u8|C1]

Workload at N=1 (BASELINE.json configs[1], "C2"): FakeDataSet 20 x 2048 x 2048 float32 psi/phi (seed 101, 1 % masks,
7x7 Gaussian PSF, 20 injected movers), VelocityGridSearch 50 x 40 = 2000 candidates, sigma-G filter on
(percentiles 25/75, coeff 0.7413, min_lh 10), min_obs 10, results_per_pixel 8.  A "step" is one search_all over
every start pixel.  For N > 1 (weak scaling) the image is N times taller (2048 x 2048*N), psi/phi is replicated on
every GPU, rank r searches the start-pixel row band [2048 r, 2048 (r+1)) and the per-GPU sorted results are merged
by an NCCL all-gather + global sort inside the timed step.

value  : whole-job evals/s with psi/phi resident in HBM when the timed region starts (preloaded stack).
e2e    : the same metric through the public API with host buffers: every step uploads the psi/phi stack from
         pinned host memory (search_all on a non-preloaded StackSearch, exactly what the reference does in
         PsiPhiArray::move_to_gpu) and reads the filtered, sorted result list back to the host.
The stack (671 MB at N=1) is larger than the 126 MB L2, so no L2 flush is needed between iterations.
Only the cpu_baseline leg and --impl reference touch oracle/ (the reference's own CPU search_all).
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

# dram__bytes_read.sum + dram__bytes_write.sum of ONE search-kernel launch from an `ncu --set full` capture of this
# workload (profiles/r01_ncu_search_tiled_raw.csv: 4.146 GB read + 0.941 GB written; the algorithmic gather
# bytes of the same launch are 1342 GB, the stack itself is 0.671 GB).
NCU_DRAM_BYTES_PER_LAUNCH = {("C2", True): 5.09e9}

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


def run_ours(args, c, cfg_name):
    from kbmod_b200 import search as kb
    from kbmod_b200 import dist

    rank, world, local_rank = dist.env_rank()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    dist.set_device(local_rank)
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
    t0 = time.perf_counter()
    s = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, c["nb"])
    ctor_s = time.perf_counter() - t0
    configure(s, c, band if (world > 1 or c.get("band_div")) else None)
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
    ms_e2e_dev = s.timer_stop()
    wall_e2e = (time.perf_counter() - t0) * 1e3
    ms_e2e = max(ms_e2e_dev, wall_e2e) / e2e_steps  # results are read after the last kernel: use the wall clock
    if comm is not None:
        ms_e2e = comm.max_float(ms_e2e)
    e2e_value = evals_job / (ms_e2e * 1e-3)

    # ---- roofline of the dominant kernel (the search kernel), from CUDA events inside the library ----------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy)"
    else:
        peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
    bytes_per_eval = {-1: 8, 4: 8, 2: 4, 1: 2}[c["nb"]]
    kernel_ms = k_ms / max(n_kernel, 1)
    achieved = evals_rank * bytes_per_eval / (kernel_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": NCU_DRAM_BYTES_PER_LAUNCH.get((cfg_name, bool(tiled))), "traffic_unit": "bytes/launch",
                "kernel": "search_tiled_kernel" if tiled else "search_generic_kernel",
                "kernel_ms": kernel_ms, "bytes_per_eval": bytes_per_eval, "peak_source": peak_src,
                "note": "algorithmic gather bytes (P*N*T pairs) / kernel time; gathers are served from shared "
                        "memory / L2, so the fraction of the HBM copy peak can exceed 1"}

    # The gathers never go to HBM: the stack is staged once into shared memory by TMA, so the binding resource is the
    # shared-memory pipe.  Measured on this pool by scripts/peaks_micro.cu (profiles/r01_peaks_micro.jsonl):
    # L2-resident 256 B row gather and LDS.64 bandwidth -- the "L2 gather roofline" the metric names and the
    # roofline of the resource the kernel actually runs on.
    micro = os.path.join(ROOT, "profiles", "r01_peaks_micro.jsonl")
    if os.path.exists(micro):
        pk = {}
        for ln in open(micro):
            ln = ln.strip()
            if ln.startswith("{"):
                d = json.loads(ln)
                pk[d["what"]] = d["GBps"]
        if "l2_resident_row_gather_256B" in pk:
            roofline["l2_gather"] = {"peak": pk["l2_resident_row_gather_256B"], "unit": "GB/s",
                                     "frac": achieved / pk["l2_resident_row_gather_256B"],
                                     "peak_source": "scripts/peaks_micro.cu, L2-resident 256 B row gather, measured"}
        if "smem_lds64" in pk and tiled:
            roofline["shared_memory"] = {"peak": pk["smem_lds64"], "unit": "GB/s", "frac": achieved / pk["smem_lds64"],
                                         "peak_source": "scripts/peaks_micro.cu, LDS.64 on all SMs, measured",
                                         "note": "algorithmic gather bytes / time against the LDS.64 peak; pairs "
                                                 "shared by neighbouring candidates are read once, so this too "
                                                 "can exceed the bytes actually moved"}

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
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e, "steps": e2e_steps},
            "gpu_launches": int(launches), "clocks": clocks}

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            ss, cands, evals, cores, sample = reference_cpu_setup(c, stack, vx, vy, target_seconds=10.0)
            t0 = time.perf_counter()
            ss.search_all(cands, False)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": evals / dt, "unit": UNIT, "cores": cores, "kind": "reference",
                                    "sample": sample, "seconds": dt}
        except Exception as e:  # the reference build did not travel: fall back to the C port of the oracle
            line["cpu_baseline"] = cpu_baseline_port(c, stack, vx, vy, str(e))
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
