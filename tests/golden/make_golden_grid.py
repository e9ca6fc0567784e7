"""Generate tests/golden/grid_filter_case.npz with the REFERENCE's own TrajectoryClusterGrid
(src/kbmod/filters/clustering_grid.py, loaded unmodified; its only import, four ``extract_all_trajectory_*``
helpers of kbmod.search, comes from the reference build in oracle/_ref).

    python tests/golden/make_golden_grid.py      (build container: needs /root/reference and oracle/_ref)
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("KBMOD_REFERENCE", "/root/reference")


def load_reference_grid():
    from oracle import ref as oref
    S, _ = oref.load()
    sys.modules.setdefault("kbmod", types.ModuleType("kbmod"))
    sys.modules["kbmod.search"] = S
    spec = importlib.util.spec_from_file_location("ref_clustering_grid",
                                                  os.path.join(REF, "src", "kbmod", "filters", "clustering_grid.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod, S


def make_trajectories(rng, n):
    """Result-like records: clustered start pixels (near duplicates), negative coordinates, exact lh ties."""
    nc = max(n // 25, 1)
    cx, cy = rng.integers(-30, 400, nc), rng.integers(-30, 300, nc)
    cvx, cvy = rng.uniform(-25, 25, nc), rng.uniform(-25, 25, nc)
    k = rng.integers(0, nc, n)
    x = (cx[k] + rng.integers(-6, 7, n)).astype(np.int32)
    y = (cy[k] + rng.integers(-6, 7, n)).astype(np.int32)
    vx = (cvx[k] + rng.uniform(-1.5, 1.5, n)).astype(np.float32)
    vy = (cvy[k] + rng.uniform(-1.5, 1.5, n)).astype(np.float32)
    lh = np.round(rng.uniform(5, 40, n), 1).astype(np.float32)      # one decimal -> many exact ties
    lh[::17] = 0.0
    lh[5::29] = -0.0
    return x, y, vx, vy, lh


def main():
    mod, S = load_reference_grid()
    rng = np.random.default_rng(123)
    out = {}
    cases = [(400, 10, 1.0), (1500, 25, 4.03), (60, 1, 0.0), (300, 7.5, 2.5), (64, 25, 4.03)]
    out["cases"] = np.array(cases, dtype=np.float64)
    for i, (n, bw, mt) in enumerate(cases):
        n = int(n)
        x, y, vx, vy, lh = make_trajectories(rng, n)
        if i == 4:
            # A bin edge that tells a double product from a float32 one: (69 + 4.03 * 7.692307472229004) / 25 is
            # 3.99999... in double (bin 3) and 4.0 when 4.03 * vx is rounded to float32 first (bin 4).
            x[:2], y[:2], vx[:2], vy[:2] = 69, 10, np.float32(7.692307472229004), 0.0
            x[1] = 70
            lh[:2] = (20.0, 30.0)
        trjs = [S.Trajectory(x=int(a), y=int(b), vx=float(c), vy=float(d), lh=float(e), flux=1.0, obs_count=10)
                for a, b, c, d, e in zip(x, y, vx, vy, lh)]
        bw_arg = int(bw) if float(bw).is_integer() else bw
        _, idx = mod.apply_trajectory_grid_filter(trjs, bw_arg, mt)
        g = mod.TrajectoryClusterGrid(bin_width=bw_arg, max_time=mt)
        g.add_trajectory_list(trjs)
        for k, v in dict(x=x, y=y, vx=vx, vy=vy, lh=lh).items():
            out[f"c{i}_{k}"] = v
        out[f"c{i}_keep"] = np.array(idx, dtype=np.int64)
        out[f"c{i}_keep_list"] = np.array(g.get_indices(), dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "grid_filter_case.npz"), **out)
    print("wrote grid_filter_case.npz", [(int(c[0]), len(out[f"c{i}_keep"]), len(out[f"c{i}_keep_list"])) for i, c in enumerate(cases)])


if __name__ == "__main__":
    main()
