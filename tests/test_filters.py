"""Sigma-G clipping of likelihood curves + likelihood update.

CPU part: the numpy oracle (oracle/filters.py) against fixtures written by the reference's own SigmaGClipping class
(tests/golden/make_golden_filters.py) and, where /root/reference exists, live against that class.
GPU part (-m gpu): the CUDA kernels through the C-ABI against the same fixtures and the oracle -- keep matrices
bit-exact, likelihood / flux bit-exact (float32 sums in numpy's pairwise order), obs_count exact.
"""
import os

import numpy as np
import pytest

from oracle import filters as of


@pytest.fixture(scope="module")
def golden(golden_dir):
    return np.load(os.path.join(golden_dir, "sigmag_filter_case.npz"))


def same_f64(a, b):
    """Bit-exact where finite, NaN where NaN (a NaN's payload is not part of the contract)."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    na, nb = np.isnan(a), np.isnan(b)
    return a.shape == b.shape and np.array_equal(na, nb) and np.array_equal(a[~na].view(np.uint64), b[~nb].view(np.uint64))


def _cases(golden):
    for i, (R, T, lo, hi, ns, neg) in enumerate(golden["cases"]):
        yield i, int(R), int(T), float(lo), float(hi), float(ns), bool(neg)


def test_oracle_matches_reference_fixtures(golden):
    for i, R, T, lo, hi, ns, neg in _cases(golden):
        psi, phi, prior = golden[f"c{i}_psi"], golden[f"c{i}_phi"], golden[f"c{i}_prior"]
        with np.errstate(all="ignore"):
            lh = of.lh_curves(psi, phi, prior)
        assert np.array_equal(lh.view(np.uint32), golden[f"c{i}_lh_curves"].view(np.uint32))
        assert abs(of.find_sigma_g_coeff(lo, hi) - float(golden[f"c{i}_coeff"])) < 1e-15
        valid = of.clipped_sigma_g_matrix(lh, lo, hi, ns, neg)
        assert np.array_equal(valid, golden[f"c{i}_valid"]), f"case {i}"
        with np.errstate(all="ignore"):
            ulh, uflux, uobs = of.update_likelihood(psi, phi, valid)
        assert same_f64(ulh, golden[f"c{i}_lh"]) and same_f64(uflux, golden[f"c{i}_flux"])
        assert np.array_equal(uobs, golden[f"c{i}_obs"])


def _single_cases(golden):
    for k, (i, r, neg) in enumerate(golden["single_meta"]):
        R, T, lo, hi, ns, _ = golden["cases"][i]
        lh = np.nan_to_num(golden[f"c{i}_lh_curves"].astype(np.float64), nan=0.0)[r]
        yield k, lh, float(lo), float(hi), float(ns), bool(neg), float(golden[f"c{i}_coeff"])


def test_oracle_single_curve_clip_matches_reference_fixtures(golden):
    n = 0
    for k, lh, lo, hi, ns, neg, coeff in _single_cases(golden):
        assert np.array_equal(of.sigmag_clip_curve(lh, lo, hi, ns, coeff, neg), golden[f"single_{k}"]), k
        n += len(golden[f"single_{k}"])
    assert n > 100


def test_oracle_pairwise_sum_is_numpys():
    rng = np.random.default_rng(5)
    for n in [0, 1, 7, 8, 9, 64, 127, 128, 129, 255, 256, 257, 400, 1000, 4096]:
        a = rng.normal(0, 3, n).astype(np.float32)
        if n:
            assert of.pairwise_sum_f32(a).tobytes() == np.float32(a.sum()).tobytes(), n


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/kbmod/filters"), reason="reference sources not present")
def test_oracle_matches_reference_live():
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    import make_golden_filters as mg
    ref = mg.load_reference_filter()
    rng = np.random.default_rng(77)
    for trial in range(12):
        R, T = int(rng.integers(1, 40)), int(rng.integers(1, 90))
        lh = rng.normal(1, 3, (R, T)).astype(np.float32)
        lh[rng.random((R, T)) < 0.2] = np.nan
        lo, hi = [(25, 75), (10, 90), (15, 60)][trial % 3]
        for neg in (False, True):
            want = ref.SigmaGClipping(lo, hi, 2, neg).compute_clipped_sigma_g_matrix(lh)
            assert np.array_equal(of.clipped_sigma_g_matrix(lh, lo, hi, 2, neg), want)


def test_mirror_validates_like_the_reference():
    from kbmod_b200.sigma_g_filter import SigmaGClipping
    for bad in [(75, 25, 2), (0, 75, 2), (25, 100, 2), (25, 75, 0), (25, 75, -1)]:
        with pytest.raises(ValueError):
            SigmaGClipping(*bad)
    with pytest.raises(ValueError):
        SigmaGClipping.find_sigma_g_coeff(80, 20)
    c = SigmaGClipping(25, 75, 2)
    assert abs(c.coeff - 0.7413011092528009) < 1e-12      # tests/test_sigma_g_filter.py in the reference: 0.7413


# ---- GPU ------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_gpu_clip_matrix_bit_exact(golden):
    from kbmod_b200.sigma_g_filter import SigmaGClipping
    for i, R, T, lo, hi, ns, neg in _cases(golden):
        got = SigmaGClipping(lo, hi, ns, neg).compute_clipped_sigma_g_matrix(golden[f"c{i}_lh_curves"])
        assert got.dtype == bool and got.shape == (R, T)
        assert np.array_equal(got, golden[f"c{i}_valid"]), f"case {i}"


@pytest.mark.gpu
def test_gpu_single_curve_clip_matches_reference_fixtures(golden):
    """SigmaGClipping.compute_clipped_sigma_g (sigma_g_filter.py:86-112; float64, numpy percentiles) on the device."""
    from kbmod_b200.sigma_g_filter import SigmaGClipping
    for k, lh, lo, hi, ns, neg, _ in _single_cases(golden):
        got = SigmaGClipping(lo, hi, ns, neg).compute_clipped_sigma_g(lh)
        assert np.array_equal(got, golden[f"single_{k}"]), k
    rng = np.random.default_rng(3)
    for n in (1, 2, 3, 50, 1025, 5000):
        lh = np.round(rng.normal(2, 3, n), 2)     # ties
        for neg in (False, True):
            c = SigmaGClipping(20, 80, 2.5, neg)
            assert np.array_equal(c.compute_clipped_sigma_g(lh), of.sigmag_clip_curve(lh, 20, 80, 2.5, c.coeff, neg)), (n, neg)
    assert len(SigmaGClipping(25, 75, 2, True).compute_clipped_sigma_g(np.array([-1.0, -2.0, 0.0]))) == 0


@pytest.mark.gpu
def test_gpu_update_likelihood_bit_exact(golden):
    from kbmod_b200.sigma_g_filter import update_likelihood
    for i, R, T, lo, hi, ns, neg in _cases(golden):
        lh, flux, obs = update_likelihood(golden[f"c{i}_psi"], golden[f"c{i}_phi"], golden[f"c{i}_valid"])
        assert same_f64(lh, golden[f"c{i}_lh"]), f"case {i}"
        assert same_f64(flux, golden[f"c{i}_flux"]), f"case {i}"
        assert np.array_equal(obs, golden[f"c{i}_obs"])
    # no obs_valid column: every point counts (results.py:635-638)
    psi, phi = golden["c2_psi"], golden["c2_phi"]
    with np.errstate(all="ignore"):
        wl, wf, wo = of.update_likelihood(psi, phi, None)
    lh, flux, obs = update_likelihood(psi, phi, None)
    assert same_f64(lh, wl) and same_f64(flux, wf)
    assert np.array_equal(obs, wo)


@pytest.mark.gpu
def test_gpu_fused_filter_matches_the_three_steps(golden):
    from kbmod_b200.sigma_g_filter import SigmaGClipping, apply_clipped_sigma_g_arrays
    for i, R, T, lo, hi, ns, neg in _cases(golden):
        ov, lh, flux, obs = apply_clipped_sigma_g_arrays(SigmaGClipping(lo, hi, ns, neg), golden[f"c{i}_psi"],
                                                         golden[f"c{i}_phi"], golden[f"c{i}_prior"])
        assert np.array_equal(ov, golden[f"c{i}_valid"]), f"case {i}"
        assert same_f64(lh, golden[f"c{i}_lh"]) and same_f64(flux, golden[f"c{i}_flux"])
        assert np.array_equal(obs, golden[f"c{i}_obs"])


@pytest.mark.gpu
def test_gpu_clip_large_random_against_oracle():
    from kbmod_b200.sigma_g_filter import SigmaGClipping
    rng = np.random.default_rng(9)
    for (R, T) in [(3000, 20), (500, 100), (3, 4096), (1, 1), (64, 33)]:
        lh = rng.normal(2, 4, (R, T)).astype(np.float32)
        lh[rng.random((R, T)) < 0.1] = np.nan
        lh[0, :] = np.nan
        for neg in (False, True):
            got = SigmaGClipping(25, 75, 2, neg).compute_clipped_sigma_g_matrix(lh)
            want = of.clipped_sigma_g_matrix(lh, 25, 75, 2, neg)
            assert np.array_equal(got, want), (R, T, neg)
    with pytest.raises(RuntimeError):
        SigmaGClipping().compute_clipped_sigma_g_matrix(np.zeros((2, 5000), dtype=np.float32))
    assert SigmaGClipping().compute_clipped_sigma_g_matrix(np.zeros((0, 7), dtype=np.float32)).shape == (0, 7)


@pytest.mark.gpu
def test_gpu_filter_on_search_results():
    """End of the path: search -> psi/phi curves of the results -> fused sigma-G filter, against the oracle."""
    from cases import make_case
    from kbmod_b200 import search as kb
    from kbmod_b200.sigma_g_filter import SigmaGClipping, apply_clipped_sigma_g_arrays
    case = make_case(seed=41, T=24, H=64, W=80, n_cand=80, objects=3)
    s = kb.StackSearch(case["sci"], case["var"], case["psfs"], case["times"])
    s.set_min_obs(8)
    s.set_min_lh(3.0)
    s.search_all((case["vx"], case["vy"]), True)
    res = s.get_results(0, 500)
    assert len(res) > 10
    curves = s.get_all_psi_phi_curves(res)
    T = case["T"]
    psi, phi = np.ascontiguousarray(curves[:, :T]), np.ascontiguousarray(curves[:, T:])
    ov, lh, flux, obs = apply_clipped_sigma_g_arrays(SigmaGClipping(25, 75, 2), psi, phi)
    with np.errstate(all="ignore"):
        want_ov = of.clipped_sigma_g_matrix(of.lh_curves(psi, phi), 25, 75, 2, False)
        wl, wf, wo = of.update_likelihood(psi, phi, want_ov)
    assert np.array_equal(ov, want_ov)
    assert same_f64(lh, wl) and same_f64(flux, wf) and np.array_equal(obs, wo)
