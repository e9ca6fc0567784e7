"""Near-duplicate grid filter (filters/clustering_grid.py).  CPU: the oracle restatement against
fixtures written by the reference's own TrajectoryClusterGrid (tests/golden/make_golden_grid.py).  GPU (-m gpu):
kb_grid_filter through the C-ABI against the same fixtures -- the kept indices and their order are exact."""
import os

import numpy as np
import pytest

from oracle import filters as of


@pytest.fixture(scope="module")
def golden(golden_dir):
    return np.load(os.path.join(golden_dir, "grid_filter_case.npz"))


def _cases(g):
    for i, (n, bw, mt) in enumerate(g["cases"]):
        bw = int(bw) if float(bw).is_integer() else float(bw)
        yield i, bw, float(mt), [g[f"c{i}_{k}"] for k in ("x", "y", "vx", "vy", "lh")]


def test_oracle_matches_reference_fixtures(golden):
    dup = 0
    for i, bw, mt, (x, y, vx, vy, lh) in _cases(golden):
        assert np.array_equal(of.grid_filter_indices(x, y, vx, vy, lh, bw, mt), golden[f"c{i}_keep"]), i
        assert np.array_equal(of.grid_filter_indices(x, y, vx, vy, lh, bw, mt, True), golden[f"c{i}_keep_list"]), i
        dup += len(x) - len(golden[f"c{i}_keep"])
    assert dup > 300      # the fixtures do contain near duplicates


def _records(x, y, vx, vy, lh):
    from kbmod_b200._lib import TRAJECTORY_DTYPE
    a = np.zeros(len(x), dtype=TRAJECTORY_DTYPE)
    a["x"], a["y"], a["vx"], a["vy"], a["lh"] = x, y, vx, vy, lh
    a["obs_count"] = 10
    return a


@pytest.mark.gpu
def test_gpu_grid_filter_matches_reference_fixtures(golden):
    from kbmod_b200 import search as kb
    from kbmod_b200.clustering_grid import TrajectoryClusterGrid, apply_trajectory_grid_filter
    for i, bw, mt, (x, y, vx, vy, lh) in _cases(golden):
        arr = _records(x, y, vx, vy, lh)
        kept, idx = apply_trajectory_grid_filter(arr, bw, mt)
        assert idx == list(golden[f"c{i}_keep"]), i
        assert np.array_equal(kept, arr[golden[f"c{i}_keep"]])
        trjs = kb.array_to_trajectories(arr)
        kept2, idx2 = apply_trajectory_grid_filter(trjs, bw, mt)
        assert idx2 == idx and all(a is trjs[j] for a, j in zip(kept2, idx2))
        g = TrajectoryClusterGrid(bin_width=bw, max_time=mt)
        g.add_trajectory_list(trjs)
        assert g.get_indices() == list(golden[f"c{i}_keep_list"]) and len(g) == len(golden[f"c{i}_keep_list"])
        assert g.total_count == len(trjs)
        g1 = TrajectoryClusterGrid(bin_width=bw, max_time=mt)
        for t in trjs[:50]:
            g1.add_trajectory(t)
        assert g1.get_indices() == list(of.grid_filter_indices(x[:50], y[:50], vx[:50], vy[:50], lh[:50], bw, mt))


@pytest.mark.gpu
def test_gpu_grid_filter_large_and_errors():
    from kbmod_b200.clustering_grid import TrajectoryClusterGrid, apply_trajectory_grid_filter
    rng = np.random.default_rng(2)
    n = 200000
    x = rng.integers(-50, 2100, n).astype(np.int32)
    y = rng.integers(-50, 2100, n).astype(np.int32)
    vx = rng.choice(np.linspace(-40, 40, 50), n).astype(np.float32)
    vy = rng.choice(np.linspace(-32, 32, 40), n).astype(np.float32)
    lh = np.round(rng.uniform(10, 30, n), 2).astype(np.float32)
    arr = _records(x, y, vx, vy, lh)
    _, idx = apply_trajectory_grid_filter(arr, 40, 4.0)
    want = of.grid_filter_indices(x, y, vx, vy, lh, 40, 4.0)
    assert len(want) < n and idx == list(want)
    assert apply_trajectory_grid_filter(arr[:0], 10, 1.0)[1] == []
    for bad in [(0, 1.0), (10, -1.0), (float("nan"), 1.0)]:
        with pytest.raises(ValueError):
            apply_trajectory_grid_filter(arr[:5], *bad)
        with pytest.raises(ValueError):
            TrajectoryClusterGrid(*bad)


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/kbmod/filters"), reason="reference sources not present")
def test_oracle_matches_reference_live():
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    import make_golden_grid as mg
    from oracle import ref as oref
    if not oref.available():
        pytest.skip("oracle/_ref not built")
    mod, S = mg.load_reference_grid()
    rng = np.random.default_rng(77)
    for trial, (bw, mt) in enumerate([(10, 1.0), (3, 7.7), (16.5, 0.3)]):
        x, y, vx, vy, lh = mg.make_trajectories(rng, 500)
        trjs = [S.Trajectory(x=int(a), y=int(b), vx=float(c), vy=float(d), lh=float(e), flux=1.0, obs_count=10)
                for a, b, c, d, e in zip(x, y, vx, vy, lh)]
        _, idx = mod.apply_trajectory_grid_filter(trjs, bw, mt)
        assert np.array_equal(of.grid_filter_indices(x, y, vx, vy, lh, bw, mt), np.array(idx)), trial
        g = mod.TrajectoryClusterGrid(bin_width=bw, max_time=mt)
        g.add_trajectory_list(trjs)
        assert np.array_equal(of.grid_filter_indices(x, y, vx, vy, lh, bw, mt, True), np.array(g.get_indices())), trial
