"""Parity at the sizes BASELINE.json names, run by ``pytest -m gpu`` on the B200 box.

* C2 (float32) and C3 (uint16, uint8), full frame 20 x 2048 x 2048, 2000 candidates, sigma-G on: our search against
  the UNMODIFIED reference's own GPU ``search_all(candidates, True)`` from ``oracle/_ref`` on the same device
  (oracle tier O4).  Both sides get a 1x1 PSF over science images convolved beforehand, for which the
  reference's FFMA-contracted GPU convolution and its CPU convolution are the same function (see
  test_gpu_ref_crosscheck.py), so the psi/phi bytes must be identical and the ranked lists identical modulo exact
  lh ties (tests/parity.py: lh sequence bit-exact, records identical where lh is unique, every record re-evaluated).
* C4-shaped (T = 100, 20 000 ecliptic candidates, sigma-G on, min_obs 50) and C5-shaped (T = 400, 100 000 grid
  candidates, sigma-G off, min_lh 10, min_obs 200) stacks of 1024 x 1024 pixels with the real candidate grids:
  start-pixel windows against the CPU oracle -- raw slots and filtered lists.  These exercise the T > 64 sigma-G
  and the observation-counting instantiations of the tiled kernel, and the deferred clip queue.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from parity import assert_lists_equal_modulo_ties, assert_slots_equal_modulo_ties, bits  # noqa: E402

pytestmark = pytest.mark.gpu


def _ref_list(S, ss, dtype):
    n = ss.get_number_total_results()
    out = np.zeros(n, dtype=dtype)
    if n:
        lst = ss.get_results(0, n)
        for f in ("x", "y", "vx", "vy", "lh", "flux", "obs_count"):
            out[f] = getattr(S, "extract_all_trajectory_" + f)(lst)
    return out


def _oracle_stack(orc, arr, times, T, H, W):
    nb = arr.num_bytes
    pp = (arr.psi_min_val, arr.psi_max_val, arr.psi_scale) if nb != 4 else (0, 0, 1)
    fp = (arr.phi_min_val, arr.phi_max_val, arr.phi_scale) if nb != 4 else (0, 0, 1)
    return orc.PsiPhiStack(arr.host_array(), times, T, H, W, nb, pp, fp)


@pytest.fixture(scope="module")
def c2_inputs():
    import bench
    from kbmod_b200 import search as kb
    assert kb.kb_has_gpu()
    c = bench.CONFIGS["C2"]
    stack, vx, vy = bench.make_inputs(c, 1)
    sci = [kb.convolve_image(np.ascontiguousarray(a, dtype=np.float32), p) for a, p in zip(stack.sci, stack.psfs)]
    one = [np.ones((1, 1), dtype=np.float32)] * c["T"]
    return dict(c=c, sci=sci, var=list(stack.var), psfs=one, times=[float(t) for t in stack.zeroed_times], vx=vx, vy=vy)


@pytest.mark.parametrize("cfg_name", ["C2", "C3u16", "C3u8"])
def test_full_frame_matches_reference_gpu(c2_inputs, cfg_name):
    import bench
    from kbmod_b200 import search as kb
    from kbmod_b200._lib import TRAJECTORY_DTYPE
    from oracle import cpu as orc, ref as oref
    if not oref.available():
        pytest.skip("oracle/_ref (the reference build) is not present")
    S, Hn = oref.load()
    if not S.kb_has_gpu():
        pytest.skip("the reference build sees no GPU")
    c = bench.CONFIGS[cfg_name]
    d = c2_inputs
    T, H, W = c["T"], c["H"], c["W"]
    s = kb.StackSearch(d["sci"], d["var"], d["psfs"], d["times"], c["nb"])
    ss = S.StackSearch(d["sci"], d["var"], d["psfs"], d["times"], c["nb"])
    ra = S.PsiPhiArray()
    S.fill_psi_phi_array_from_image_arrays(ra, c["nb"], d["sci"], d["var"], d["psfs"], d["times"])
    ours = s.psi_phi_array
    assert np.array_equal(Hn.psi_phi_bytes(ra), ours.host_array().view(np.uint8).ravel()), "psi/phi bytes differ"
    del ra
    bench.configure(s, c)
    bench.configure(ss, c)
    cands = [S.Trajectory(vx=float(a), vy=float(b)) for a, b in zip(d["vx"], d["vy"])]
    ss.search_all(cands, True)
    want = _ref_list(S, ss, TRAJECTORY_DTYPE)
    del ss
    assert len(want) > 1000
    s.set_kernel_mode(2)   # the tiled kernel must take this config (an error otherwise, never the slow path)
    s.search_all((d["vx"], d["vy"]), True)
    got = s.get_results_array()
    assert s.last_timing()["used_tiled"] == 1
    ost = _oracle_stack(orc, ours, d["times"], T, H, W)
    prm = orc.make_params(c["min_obs"], c["min_lh"], (0.25, 0.75, 0.7413), None, c["K"], ost)
    assert_lists_equal_modulo_ties(got, want, orc, ost, prm, f"{cfg_name} full frame vs reference GPU")
    # the search of a stack that is not preloaded (chunked upload under the kernel) returns the same list
    s.set_kernel_mode(0)
    s.search_all((d["vx"], d["vy"]), True)
    again = s.get_results_array()
    assert np.array_equal(again.view(np.uint8), got.view(np.uint8))


def _window_case(name, side, windows, win):
    import bench
    from kbmod_b200 import search as kb
    from oracle import cpu as orc
    c = dict(bench.CONFIGS[name], H=side, W=side)
    c.pop("band_div", None)
    stack, vx, vy = bench.make_inputs(c, 1)
    s = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times, c["nb"])
    bench.configure(s, c)
    s.preload_psi_phi_array()
    ost = _oracle_stack(orc, s.psi_phi_array, stack.zeroed_times, c["T"], side, side)
    sg = (0.25, 0.75, 0.7413) if c["sigmag"] else None
    corners = [(int(o.x) - win // 2, int(o.y) - win // 2) for o in stack.objects[:max(windows - 1, 1)]]
    corners += [(-win // 2, side - win // 2)]      # hangs off two edges of the image
    n_results = 0
    for (x0, y0) in corners[:windows]:
        b = (x0, x0 + win, y0, y0 + win)
        s.set_start_bounds_x(b[0], b[1])
        s.set_start_bounds_y(b[2], b[3])
        s.set_kernel_mode(2)
        s.search_all((vx, vy), True)
        assert s.last_timing()["used_tiled"] == 1
        got_slots, got = s.get_slots_array(), s.get_results_array()
        p = orc.make_params(c["min_obs"], c["min_lh"], sg, b, c["K"], ost)
        want_slots, _ = orc.search_slots(ost, p, vx, vy)
        want = orc.filter_and_sort(want_slots, c["min_lh"], c["min_obs"])
        assert_slots_equal_modulo_ties(got_slots, want_slots, c["K"], orc, ost, p, f"{name}-shaped window {b}")
        assert_lists_equal_modulo_ties(got, want, orc, ost, p, f"{name}-shaped window {b}")
        n_results += len(want)
    assert n_results > 0   # the windows around the injected movers do hold results


def test_c4_shaped_windows_match_oracle():
    """T = 100 > 64: the large sigma-G instantiation, the 128-image clip kernel, min_obs 50."""
    _window_case("C4", 1024, windows=3, win=20)


def test_c5_shaped_windows_match_oracle():
    """T = 400, sigma-G off, min_obs 200: the observation-counting instantiation, 100 000 candidates."""
    _window_case("C5", 1024, windows=2, win=10)
