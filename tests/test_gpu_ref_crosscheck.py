"""Cross-check against the UNMODIFIED reference's own GPU path (oracle tier O4): its
``StackSearch.search_all(candidates, True)`` from ``oracle/_ref`` (kernels.cu compiled for sm_100a, see
oracle/build_ref.sh) runs on the same B200 and its filtered, sorted list must equal ours modulo exact lh ties.

Both sides get a 1x1 PSF [[1.0]] over science images convolved beforehand: for that PSF the reference's GPU
convolution (FFMA-contracted, image_kernels.cu:29-66) is the same function as its CPU convolution (which is the
one our library reproduces), so both search bit-identical psi/phi -- checked first, byte for byte.
"""
import numpy as np
import pytest

from cases import make_case
from parity import assert_lists_equal_modulo_ties

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    from kbmod_b200 import search as kb
    from kbmod_b200._lib import TRAJECTORY_DTYPE
    from oracle import cpu as orc, ref as oref
    assert kb.kb_has_gpu()
    if not oref.available():
        pytest.skip("oracle/_ref (the reference build) is not present")
    S, Hn = oref.load()
    if not S.kb_has_gpu():
        pytest.skip("the reference build sees no GPU")
    case = make_case(seed=29, T=16, H=160, W=192, n_cand=300, vmax=8.0, objects=3)
    sci = [kb.convolve_image(a, p) for a, p in zip(case["sci"], case["psfs"])]
    one = [np.ones((1, 1), dtype=np.float32)] * case["T"]
    return dict(kb=kb, orc=orc, S=S, Hn=Hn, case=case, sci=sci, var=case["var"], psfs=one,
                times=[float(t) for t in case["times"]], dtype=TRAJECTORY_DTYPE)


def _ref_list(S, ss, dtype):
    n = ss.get_number_total_results()
    out = np.zeros(n, dtype=dtype)
    if n:
        lst = ss.get_results(0, n)
        for f in ("x", "y", "vx", "vy", "lh", "flux", "obs_count"):
            out[f] = getattr(S, "extract_all_trajectory_" + f)(lst)
    return out


@pytest.mark.parametrize("nb", [-1, 2, 1])
@pytest.mark.parametrize("cfg", [
    (8, 1.0, (0.25, 0.75, 0.7413), None, 8),        # sigma-G on, full frame
    (8, 2.0, None, (-6, 150, -4, 170), 4),          # sigma-G off, start bounds hanging off the image
])
def test_matches_reference_gpu_search(env, nb, cfg):
    kb, S, Hn, orc, case = env["kb"], env["S"], env["Hn"], env["orc"], env["case"]
    mo, ml, sg, b, K = cfg
    s = kb.StackSearch(env["sci"], env["var"], env["psfs"], env["times"], nb)
    ss = S.StackSearch(env["sci"], env["var"], env["psfs"], env["times"], nb)
    ra = S.PsiPhiArray()
    S.fill_psi_phi_array_from_image_arrays(ra, nb, env["sci"], env["var"], env["psfs"], env["times"])
    ours = s.psi_phi_array
    assert np.array_equal(Hn.psi_phi_bytes(ra), ours.host_array().view(np.uint8).ravel()), "psi/phi bytes differ"
    for o in (s, ss):
        o.set_min_obs(mo)
        o.set_results_per_pixel(K)
        if sg is not None:
            o.enable_gpu_sigmag_filter([sg[0], sg[1]], sg[2], ml)
        else:
            o.set_min_lh(ml)
        if b is not None:
            o.set_start_bounds_x(b[0], b[1])
            o.set_start_bounds_y(b[2], b[3])
    cands = [S.Trajectory(vx=float(a), vy=float(c)) for a, c in zip(case["vx"], case["vy"])]
    ss.search_all(cands, True)
    want = _ref_list(S, ss, env["dtype"])
    assert len(want) > 0
    n4 = ours.num_bytes
    pp = (ours.psi_min_val, ours.psi_max_val, ours.psi_scale) if n4 != 4 else (0, 0, 1)
    fp = (ours.phi_min_val, ours.phi_max_val, ours.phi_scale) if n4 != 4 else (0, 0, 1)
    ost = orc.PsiPhiStack(ours.host_array(), case["times"], case["T"], case["H"], case["W"], n4, pp, fp)
    prm = orc.make_params(mo, ml, sg, b, K, ost)
    for mode in (1, 2):   # generic and tiled kernels
        s.set_kernel_mode(mode)
        s.search_all((case["vx"], case["vy"]), True)
        got = s.get_results_array()
        assert_lists_equal_modulo_ties(got, want, orc, ost, prm, f"vs reference GPU nb={nb} mode={mode}")
