"""Coadded stamps.  CPU: the numpy oracle (oracle/filters.py) against fixtures produced by the
reference's own stamp_utils functions (tests/golden/make_golden_stamps.py).  GPU (-m gpu): kb_coadd_stamps through
the C-ABI against the same fixtures and the oracle, bit-exact (float32 compared as uint32)."""
import os

import numpy as np
import pytest

from oracle import filters as of

TYPES = ("sum", "mean", "median", "weighted")


@pytest.fixture(scope="module")
def golden(golden_dir):
    return np.load(os.path.join(golden_dir, "stamps_case.npz"))


def same_f32(a, b):
    a, b = np.ascontiguousarray(a, dtype=np.float32), np.ascontiguousarray(b, dtype=np.float32)
    na, nb = np.isnan(a), np.isnan(b)
    return a.shape == b.shape and np.array_equal(na, nb) and np.array_equal(a[~na].view(np.uint32), b[~nb].view(np.uint32))


def test_oracle_matches_reference_fixtures(golden):
    g = golden
    sci, var = list(g["sci"]), list(g["var"])
    for tag, ov in (("all", None), ("valid", g["obs_valid"])):
        co = of.coadds_for_results(sci, var, g["times"], g["x"], g["y"], g["vx"], g["vy"], int(g["radius"]), ov)
        for k in TYPES:
            assert same_f32(co[k], g[f"{tag}_{k}"]), (tag, k)
    xs = of.stamp_positions(g["times"], g["x"], g["vx"])
    ys = of.stamp_positions(g["times"], g["y"], g["vy"])
    st = of.extract_stamp_stack(sci, xs[2], ys[2], int(g["radius"]))
    assert same_f32(st.astype(np.float32), g["all_stamps_r2"])


def test_oracle_curve_values_match_reference_fixture(golden):
    g = golden
    xs = of.stamp_positions(g["times"], g["x"], g["vx"])
    ys = of.stamp_positions(g["times"], g["y"], g["vy"])
    got = of.extract_curve_values(list(g["sci"]), xs, ys)
    assert got.dtype == np.float64 and np.array_equal(got, g["curve_values"], equal_nan=True)
    assert np.isnan(got).any() and np.isfinite(got).any()


@pytest.mark.gpu
def test_gpu_extract_curve_values_matches_reference_fixture(golden):
    """extract_curve_values (core/stamp_utils.py:87-141) on the device: R x T float64, NaN off the image."""
    from kbmod_b200.stamp_filters import extract_curve_values
    g = golden
    xs = of.stamp_positions(g["times"], g["x"], g["vx"])
    ys = of.stamp_positions(g["times"], g["y"], g["vy"])
    got = extract_curve_values(g["sci"], xs, ys)
    assert got.dtype == np.float64 and np.array_equal(got, g["curve_values"], equal_nan=True)
    one = extract_curve_values(list(g["sci"]), xs[2], ys[2])
    assert one.shape == (len(g["times"]),) and np.array_equal(one, g["curve_values_one"], equal_nan=True)
    with pytest.raises(ValueError):
        extract_curve_values(g["sci"], xs[:, :-1], ys[:, :-1])


@pytest.mark.gpu
def test_gpu_coadds_bit_exact(golden):
    from kbmod_b200.stamp_filters import coadd_stamps
    g = golden
    for tag, ov in (("all", None), ("valid", g["obs_valid"])):
        co = coadd_stamps(g["sci"], g["var"], g["times"], g["x"], g["y"], g["vx"], g["vy"], int(g["radius"]), TYPES, ov)
        for k in TYPES:
            assert co[k].dtype == np.float32 and same_f32(co[k], g[f"{tag}_{k}"]), (tag, k)
    only = coadd_stamps(g["sci"], None, g["times"], g["x"], g["y"], g["vx"], g["vy"], int(g["radius"]), ("median",))
    assert list(only) == ["median"] and same_f32(only["median"], g["all_median"])


@pytest.mark.gpu
def test_gpu_coadds_random_against_oracle():
    from cases import make_case
    from kbmod_b200.stamp_filters import coadd_stamps
    case = make_case(seed=61, T=37, H=50, W=44, n_cand=20, mask_frac=0.1)
    rng = np.random.default_rng(4)
    R = 40
    x = rng.integers(-8, 52, R).astype(np.float64)
    y = rng.integers(-8, 58, R).astype(np.float64)
    vx = rng.uniform(-9, 9, R).astype(np.float32).astype(np.float64)
    vy = rng.uniform(-9, 9, R).astype(np.float32).astype(np.float64)
    ov = rng.random((R, case["T"])) < 0.7
    for radius in (1, 6):
        got = coadd_stamps(case["sci"], case["var"], case["times"], x, y, vx, vy, radius, TYPES, ov)
        want = of.coadds_for_results(case["sci"], case["var"], case["times"], x, y, vx, vy, radius, ov)
        for k in TYPES:
            assert same_f32(got[k], want[k]), (radius, k)


@pytest.mark.gpu
def test_gpu_coadds_argument_errors(golden):
    from kbmod_b200.stamp_filters import coadd_stamps
    g = golden
    with pytest.raises(ValueError):
        coadd_stamps(g["sci"], g["var"], g["times"], g["x"], g["y"], g["vx"], g["vy"], 0)
    with pytest.raises(ValueError):
        coadd_stamps(g["sci"], g["var"], g["times"], g["x"], g["y"], g["vx"], g["vy"], 2, ("mode",))
    with pytest.raises(ValueError):
        coadd_stamps(g["sci"], g["var"], g["times"], g["x"], g["y"], g["vx"], g["vy"], 2, TYPES, np.ones((3, 3), bool))
    empty = coadd_stamps(g["sci"], g["var"], g["times"], [], [], [], [], 2)
    assert empty["sum"].shape == (0, 5, 5)


@pytest.mark.gpu
def test_gpu_append_coadds_on_search_results():
    """search -> results -> sigma-G filter -> coadds of the valid observations, like run_search.py does."""
    from cases import make_case
    from kbmod_b200 import search as kb
    from kbmod_b200.sigma_g_filter import SigmaGClipping, apply_clipped_sigma_g_arrays
    from kbmod_b200.stamp_filters import append_coadds
    case = make_case(seed=43, T=20, H=64, W=72, n_cand=60, objects=3)
    s = kb.StackSearch(case["sci"], case["var"], case["psfs"], case["times"])
    s.set_min_obs(8)
    s.set_min_lh(4.0)
    s.search_all((case["vx"], case["vy"]), True)
    res = s.get_results_array(0, 200)
    assert len(res) > 5
    curves = s.get_all_psi_phi_curves(s.get_results(0, len(res)))
    T = case["T"]
    ov, _, _, _ = apply_clipped_sigma_g_arrays(SigmaGClipping(), np.ascontiguousarray(curves[:, :T]),
                                               np.ascontiguousarray(curves[:, T:]))

    class Table(dict):
        colnames = property(lambda self: list(self.keys()))
        table = property(lambda self: self)

        def __len__(self):
            return len(self["x"])
    tab = Table(x=res["x"].astype(np.float64), y=res["y"].astype(np.float64), vx=res["vx"].astype(np.float64),
                vy=res["vy"].astype(np.float64), obs_valid=ov)
    stack = type("Stack", (), dict(sci=case["sci"], var=case["var"], zeroed_times=case["times"]))()
    append_coadds(tab, stack, ["sum", "mean", "median", "weighted"], 4)
    want = of.coadds_for_results(case["sci"], case["var"], case["times"], tab["x"], tab["y"], tab["vx"], tab["vy"], 4, ov)
    for k in TYPES:
        assert same_f32(tab[f"coadd_{k}"], want[k]), k
    # per-night coadds: one column per calendar day, built from that day's valid images only
    stack.times = 60000.25 + np.asarray(case["times"])
    append_coadds(tab, stack, ["sum", "median"], 4, nightly=True)
    from kbmod_b200.stamp_filters import mjd_to_day
    assert mjd_to_day(60000) == "2023-02-25"           # util_functions.py:63
    days = sorted({mjd_to_day(t) for t in stack.times})
    assert len(days) >= 2
    total = np.zeros_like(tab["coadd_sum"], dtype=np.float64)
    for d in days:
        day_mask = np.array([mjd_to_day(t) == d for t in stack.times])
        want_d = of.coadds_for_results(case["sci"], case["var"], case["times"], tab["x"], tab["y"], tab["vx"], tab["vy"], 4,
                                       ov & day_mask[None, :])
        assert same_f32(tab[f"coadd_sum_{d}"], want_d["sum"]) and same_f32(tab[f"coadd_median_{d}"], want_d["median"])


@pytest.mark.gpu
def test_gpu_all_stamps(golden):
    from kbmod_b200.stamp_filters import append_all_stamps, extract_all_stamps
    g = golden
    radius = int(g["radius"])
    got = extract_all_stamps(g["sci"], g["times"], g["x"], g["y"], g["vx"], g["vy"], radius)
    assert got.shape == (len(g["x"]), len(g["times"]), 2 * radius + 1, 2 * radius + 1) and got.dtype == np.float32
    assert same_f32(got[2], g["all_stamps_r2"])                      # the reference's own extract_stamp_stack
    xs = of.stamp_positions(g["times"], g["x"], g["vx"])
    ys = of.stamp_positions(g["times"], g["y"], g["vy"])
    for r in range(len(g["x"])):
        assert same_f32(got[r], of.extract_stamp_stack(list(g["sci"]), xs[r], ys[r], radius).astype(np.float32)), r
    assert np.isnan(got[0]).all()                                     # trajectory 0 never touches the image

    class Table(dict):
        table = property(lambda self: self)
    tab = Table(x=g["x"], y=g["y"], vx=g["vx"], vy=g["vy"])
    stack = type("Stack", (), dict(sci=list(g["sci"]), zeroed_times=g["times"]))()
    append_all_stamps(tab, stack, radius)
    assert same_f32(tab["all_stamps"], got)
    with pytest.raises(ValueError):
        append_all_stamps(tab, stack, 0)


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/kbmod/core"), reason="reference sources not present")
def test_oracle_matches_reference_live():
    """Random stacks through the reference's own stamp_utils functions (loaded as in make_golden_stamps.py)."""
    import sys
    import warnings
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    import make_golden_stamps as mg
    su = mg.load_reference_stamp_utils()
    rng = np.random.default_rng(31)
    for trial in range(4):
        T, H, W, radius = int(rng.integers(1, 9)), 17, 13, int(rng.integers(1, 5))
        sci = rng.normal(0, 3, (T, H, W)).astype(np.float32)
        var = (1 + rng.random((T, H, W))).astype(np.float32)
        sci[rng.random((T, H, W)) < 0.15] = np.nan
        var[rng.random((T, H, W)) < 0.05] = 0.0
        times = np.sort(rng.random(T) * 4)
        x, y = rng.integers(-4, W + 4, 6).astype(float), rng.integers(-4, H + 4, 6).astype(float)
        vx, vy = rng.uniform(-3, 3, 6), rng.uniform(-3, 3, 6)
        inc = rng.random((6, T)) < 0.7
        xs, ys = mg.reference_positions(times, x, vx), mg.reference_positions(times, y, vy)
        assert np.array_equal(xs, of.stamp_positions(times, x, vx)) and np.array_equal(ys, of.stamp_positions(times, y, vy))
        want = of.coadds_for_results(list(sci), list(var), times, x, y, vx, vy, radius, inc)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            for r in range(6):
                s = np.asanyarray(su.extract_stamp_stack(sci, xs[r], ys[r], radius, to_include=inc[r]))
                v = np.asanyarray(su.extract_stamp_stack(var, xs[r], ys[r], radius, to_include=inc[r]))
                assert same_f32(want["sum"][r], su.coadd_sum(s).astype(np.float32))
                assert same_f32(want["mean"][r], su.coadd_mean(s).astype(np.float32))
                assert same_f32(want["median"][r], su.coadd_median(s, device="cpu").astype(np.float32))
                assert same_f32(want["weighted"][r], su.coadd_weighted(s, v).astype(np.float32))
