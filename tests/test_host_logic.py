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
