"""CPU tests of the host-side mirror of kbmod.search (no compute calls) and of the synthetic-input generators."""
import numpy as np
import pytest

from kbmod_b200 import fake_data as fd
from kbmod_b200 import search as kb
from kbmod_b200 import _lib


def test_trajectory_fields_and_index():
    # tests/test_common.py:22-109 in the reference
    t = kb.Trajectory()
    assert (t.x, t.y, t.vx, t.vy, t.flux, t.lh, t.obs_count) == (0, 0, 0.0, 0.0, 0.0, 0.0, 0)
    t = kb.Trajectory(x=1, y=2, vx=3.0, vy=4.0, flux=5.0, lh=6.0, obs_count=7)
    assert t.get_x_pos(0.0) == 1.5 and t.get_y_pos(0.0) == 2.5
    assert t.get_x_pos(1.0, False) == 4.0 and t.get_x_index(1.0) == 4 and t.get_y_index(2.0) == 10
    t2 = kb.Trajectory(x=5, y=7, vx=-2.25, vy=1.1)
    assert t2.vy == float(np.float32(1.1))  # stored as float32
    assert t2.get_x_index(1.0) == 3 and t2.get_x_index(2.0) == 1
    assert t2.is_valid()
    t2.lh = float("nan")
    assert not t2.is_valid()


def test_trajectory_list_semantics():
    trjs = [kb.Trajectory(x=i, lh=float(i % 5), obs_count=i) for i in range(10)]
    tl = kb.TrajectoryList(trjs)
    tl.sort_by_likelihood()
    lhs = [t.lh for t in tl.get_list()]
    assert lhs == sorted(lhs, reverse=True)
    tl.filter_by_likelihood(2.0)
    assert all(t.lh >= 2.0 for t in tl.get_list())
    tl.filter_by_obs_count(4)
    assert all(t.obs_count >= 4 for t in tl.get_list())
    with pytest.raises(RuntimeError):
        tl.get_batch(0, 0)
    with pytest.raises(RuntimeError):
        kb.TrajectoryList([kb.Trajectory(lh=float("inf"))])


def test_encode_decode_scalars():
    # tests/test_psi_phi_array.py:87-114
    assert kb.decode_uint_scalar(1.0, 0.0, 5.0) == 0.0
    assert kb.decode_uint_scalar(3.0, 2.5, 3.0) == 8.5
    assert np.isnan(kb.decode_uint_scalar(0.0, 1.0, 5.0))
    assert abs(kb.encode_uint_scalar(0.1, 0.0, 10.0, 0.1) - 2.0) < 1e-5
    assert kb.encode_uint_scalar(float("nan"), 0.0, 10.0, 0.1) == 0.0
    assert abs(kb.encode_uint_scalar(-100.0, 0.0, 10.0, 0.1) - 1.0) < 1e-6


def test_fake_times_and_generators():
    t = fd.create_fake_times(6, 57130.2, obs_per_day=3, intra_night_gap=0.01)
    assert np.allclose(np.array(t) - 57130.2, [0, 0.01, 0.02, 1, 1.01, 1.02])
    g = fd.VelocityGridSearch(50, -40, 40, 40, -32, 32)
    vx, vy = g.to_arrays()
    assert len(g) == 2000 and vx.size == 2000 and vx[0] == -40 and vx[49] == 40 and vy[0] == -32 and vy[50] > vy[0]
    k = fd.KBMODV1Search(5, 0, 4, 5, -0.1, 0.1)
    kvx, kvy = k.to_arrays()
    assert kvx.size == 25 and kvx[0] == 0 and abs(kvx[1] - np.float32(np.cos(-0.1) * 0.8)) < 1e-7
    psf = fd.make_gaussian_kernel(1.0)
    assert psf.shape == (7, 7) and abs(psf.sum() - 1) < 1e-6


def test_fake_stack_is_seeded():
    a = fd.FakeDataSet(32, 16, fd.create_fake_times(4), use_seed=5, mask_fraction=0.1)
    b = fd.FakeDataSet(32, 16, fd.create_fake_times(4), use_seed=5, mask_fraction=0.1)
    for x, y in zip(a.stack_py.sci, b.stack_py.sci):
        assert np.array_equal(x, y, equal_nan=True)
    rng = np.random.default_rng(5)
    ref0 = rng.normal(0.0, 2.0, (16, 32)).astype(np.float32)
    ok = np.isfinite(a.stack_py.sci[0])
    assert np.array_equal(a.stack_py.sci[0][ok], ref0[ok])
    trj = kb.Trajectory(x=3, y=4, vx=1.0, vy=0.5, flux=100.0)
    before = a.stack_py.sci[0].copy()
    a.insert_object(trj)
    assert np.nanmax(a.stack_py.sci[0] - before) > 10


@pytest.mark.skipif(_lib.device_count() > 0, reason="checks the no-GPU error path")
def test_compute_calls_fail_loudly_without_gpu():
    img = np.zeros((4, 4), np.float32)
    with pytest.raises(RuntimeError):
        kb.convolve_image(img, np.ones((1, 1), np.float32))
    with pytest.raises(RuntimeError):
        kb.StackSearch([img], [img + 1], [np.ones((1, 1), np.float32)], [0.0])
    with pytest.raises(RuntimeError):
        kb.search_cpu_only(None, None, None, None)


def test_post_search_mirrors_validate_before_touching_the_gpu():
    """Argument errors of the filter / stamp mirrors are the reference's (ValueError) and need no device."""
    import numpy as np
    from kbmod_b200.clustering_grid import TrajectoryClusterGrid, apply_trajectory_grid_filter
    from kbmod_b200.sigma_g_filter import SigmaGClipping
    from kbmod_b200.stamp_filters import coadd_stamps, mjd_to_day
    for bad in [(0, 1.0), (10, -1.0), (float("inf"), 1.0)]:
        with pytest.raises(ValueError):
            TrajectoryClusterGrid(*bad)
        with pytest.raises(ValueError):
            apply_trajectory_grid_filter([], *bad)
    with pytest.raises(ValueError):
        SigmaGClipping(low_bnd=80, high_bnd=20)
    img = [np.zeros((8, 8), np.float32)] * 3
    with pytest.raises(ValueError):
        coadd_stamps(img, img, [0.0, 1.0, 2.0], [1], [1], [0.0], [0.0], 0)
    with pytest.raises(ValueError):
        coadd_stamps(img, img, [0.0, 1.0, 2.0], [1], [1], [0.0], [0.0], 2, ("mode",))
    with pytest.raises(ValueError):
        coadd_stamps(img, img, [0.0, 1.0], [1], [1], [0.0], [0.0], 2)          # one time per image
    with pytest.raises(ValueError):
        coadd_stamps(img, img, [0.0, 1.0, 2.0], [1, 2], [1], [0.0], [0.0], 2)   # ragged trajectory columns
    assert mjd_to_day(60000) == "2023-02-25" and mjd_to_day(60000.99) == "2023-02-25"


# ---- the bound behind clip_kernel's certain rejections (search_device.cuh: warp_sum, lh_interval) ------------------
def _warp_tree_sum(v):
    """float32 sum of v the way the kernel takes it: lane l adds elements l, l + 32, ... in order, then a butterfly."""
    f = np.float32
    lanes = np.zeros(32, dtype=f)
    for i, x in enumerate(v):
        lanes[i % 32] = f(lanes[i % 32] + f(x))
    d = 16
    while d:
        lanes = (lanes + lanes[np.arange(32) ^ d]).astype(f)
        d >>= 1
    return lanes[0]


def _lh_interval(sp, sa, sf, n):
    f = np.float32
    g = f(f(n) * f(1.5e-7))
    dp, df = f(f(g * sa) + f(1e-30)), f(f(g * sf) + f(1e-30))
    p_hi, p_lo, f_hi, f_lo = f(sp + dp), f(sp - dp), f(sf + df), f(sf - df)
    if not (f_lo > 0):
        return None
    s_lo, s_hi = np.sqrt(f_lo, dtype=f), np.sqrt(f_hi, dtype=f)
    hi = f(p_hi / s_lo) if p_hi >= 0 else f(p_hi / s_hi)
    lo = f(p_lo / s_hi) if p_lo >= 0 else f(p_lo / s_lo)
    hi = f(hi + f(f(abs(hi)) * f(4e-6) + f(1e-30)))
    lo = f(lo - f(f(abs(lo)) * f(4e-6) + f(1e-30)))
    return lo, hi


def test_bracketed_sums_contain_the_sequential_sum():
    """A float sum taken by warp reduction, widened by 1.5e-7 n sum|v|, brackets the reference's sequential float sum
    of the same numbers, and the lh interval built from two such sums contains the reference's lh -- for sums with
    heavy cancellation, outliers and every n up to 256."""
    rng = np.random.default_rng(5)
    f = np.float32
    for trial in range(400):
        n = int(rng.integers(1, 257))
        psi = rng.normal(0, 1, n).astype(f) * f(10.0 ** rng.uniform(-3, 3))
        if trial % 3 == 0:
            psi[rng.integers(0, n)] += f(10.0 ** rng.uniform(0, 6))       # one bright source on the trajectory
        if trial % 5 == 0:
            psi = np.concatenate([psi, -psi])[:n]                         # near-total cancellation
        phi = (rng.random(n).astype(f) + f(0.01)) * f(10.0 ** rng.uniform(-4, 4))
        seq_p = seq_f = f(0)
        for a, b in zip(psi, phi):                                        # kernels.cu:175-197, time order
            seq_p, seq_f = f(seq_p + a), f(seq_f + b)
        sp, sa, sf = _warp_tree_sum(psi), _warp_tree_sum(np.abs(psi)), _warp_tree_sum(phi)
        g = f(f(n) * f(1.5e-7))
        assert abs(float(sp) - float(seq_p)) <= float(g) * float(sa) + 1e-30
        assert abs(float(sf) - float(seq_f)) <= float(g) * float(sf) + 1e-30
        iv = _lh_interval(sp, sa, sf, n)
        assert iv is not None
        ref_lh = f(seq_p / np.sqrt(seq_f, dtype=f))                      # kernels.cu:195-197
        assert iv[0] <= ref_lh <= iv[1], (n, iv, ref_lh)
        # and in a sorted (ascending psi/phi) order, the order of the clipped re-sum (kernels.cu:233-237)
        order = np.argsort(psi / phi, kind="stable")
        seq_p = seq_f = f(0)
        for a, b in zip(psi[order], phi[order]):
            seq_p, seq_f = f(seq_p + a), f(seq_f + b)
        ref_lh = f(seq_p / np.sqrt(seq_f, dtype=f))
        assert iv[0] <= ref_lh <= iv[1], (n, iv, ref_lh)
