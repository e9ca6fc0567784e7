"""The reference's own known-answer tests for the post-search steps, run against the device implementations
(-m gpu): tests/test_sigma_g_filter.py:47-118,131-148, tests/test_clustering_grid.py:8-106 and
tests/test_stamp_utils.py:143-224 of kbmod @ 617cd33, restated with the mirror modules' names."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_sigma_g_clipping_matrix():                       # test_sigma_g_filter.py:47-67
    from kbmod_b200.sigma_g_filter import SigmaGClipping
    lh = np.array([[(10.0 + i * 0.05) for i in range(20)] for _ in range(5)])
    lh[1, 2] = 100.0
    lh[1, 14] = -100.0
    lh[2, 0] = 50.0
    lh[3, 2] = 100.0
    lh[3, 14] = -100.0
    lh[3, 0] = 50.0
    lh[4, 7] = np.nan
    lh[4, 8] = np.nan
    lh[4, 11] = np.nan
    expected = np.isfinite(lh) & (lh < 20.0) & (lh > 0.0)
    assert np.array_equal(SigmaGClipping().compute_clipped_sigma_g_matrix(lh), expected)


def test_sigma_g_clipping_matrix_same():                  # :69-76
    from kbmod_b200.sigma_g_filter import SigmaGClipping
    lh = np.array([[5 for i in range(10)], [5.1 for i in range(10)]])
    assert np.array_equal(SigmaGClipping().compute_clipped_sigma_g_matrix(lh), np.full(lh.shape, True))


def test_sigma_g_clipping_matrix_negative_clipping():     # :95-118
    from kbmod_b200.sigma_g_filter import SigmaGClipping
    n = 20
    lh = np.array([[5 for i in range(n)], [(-1.0 + i * 0.2) for i in range(n)], [(-100.0 + i * 0.2) for i in range(n)]])
    expected = np.array([[True] * n, [False, False, False] + [True] * (n - 3), [False] * n])
    assert np.array_equal(SigmaGClipping(clip_negative=True).compute_clipped_sigma_g_matrix(lh), expected)


def test_apply_clipped_sigma_g_results():                 # :120-143, on plain arrays
    from kbmod_b200.sigma_g_filter import SigmaGClipping, apply_clipped_sigma_g_arrays
    num_times, num_results = 20, 5
    phi_all = np.full((num_results, num_times), 0.1)
    psi_all = np.full((num_results, num_times), 1.0)
    for i in range(5):
        for j in range(i):
            psi_all[i, j] = 100.0
    valid, lh, flux, obs = apply_clipped_sigma_g_arrays(SigmaGClipping(10, 90), psi_all, phi_all)
    for i in range(num_results):
        assert not valid[i, :i].any() and valid[i, i:].all()
        assert obs[i] == num_times - i
    empty = apply_clipped_sigma_g_arrays(SigmaGClipping(10, 90), np.zeros((0, 10)), np.zeros((0, 10)))   # :145-158
    assert empty[0].shape == (0, 10) and len(empty[1]) == 0


def _trjs(kb):
    T = kb.Trajectory
    return [T(0, 0, 0.0, 0.0, 1.0, 10.0, 10), T(21, 21, 10.0, 10.0, 1.0, 10.0, 10), T(21, 21, 0.0, 0.0, 1.0, 10.0, 10),
            T(21, 21, 0.0, 0.0, 1.0, 100.0, 9), T(0, 0, 0.0, 0.0, 1.0, 5.0, 5), T(0, 0, 0.0, 0.0, 1.0, 15.0, 15)]


def test_trajectory_cluster_grid_basic():                 # test_clustering_grid.py:8-57
    from kbmod_b200 import search as kb
    from kbmod_b200.clustering_grid import TrajectoryClusterGrid
    t = _trjs(kb)
    table = TrajectoryClusterGrid(10, 1.0)
    assert len(table) == 0 and table.total_count == 0
    table.add_trajectory(t[0])
    assert len(table) == 1 and table.total_count == 1
    assert table.table.get((0, 0, 0, 0)) is not None and table.count.get((0, 0, 0, 0)) == 1
    assert table.get_indices() == [0]
    for k in (1, 2, 3):
        table.add_trajectory(t[k])
    assert len(table) == 3 and table.total_count == 4
    for key in [(0, 0, 0, 0), (2, 2, 3, 3), (2, 2, 2, 2)]:
        assert table.table.get(key) is not None
    assert table.count.get((0, 0, 0, 0)) == 1 and table.count.get((2, 2, 3, 3)) == 1 and table.count.get((2, 2, 2, 2)) == 2
    assert table.table[(2, 2, 2, 2)].obs_count == 9
    assert set(table.get_indices()) == {0, 1, 3} and len(table.get_trajectories()) == 3
    table.add_trajectory(t[4])                                        # a worse one for bin (0, 0, 0, 0)
    assert len(table) == 3 and table.count[(0, 0, 0, 0)] == 2 and table.table[(0, 0, 0, 0)].obs_count == 10
    assert set(table.get_indices()) == {0, 1, 3}
    table.add_trajectory(t[5], idx=10)                                # a better one
    assert len(table) == 3 and table.count[(0, 0, 0, 0)] == 3 and table.table[(0, 0, 0, 0)].obs_count == 15
    assert set(table.get_indices()) == {10, 1, 3} and len(table.get_trajectories()) == 3


def test_trajectory_cluster_add_list():                   # :59-86
    from kbmod_b200 import search as kb
    from kbmod_b200.clustering_grid import TrajectoryClusterGrid
    table = TrajectoryClusterGrid(10, 1.0)
    table.add_trajectory_list(_trjs(kb)[:5])
    assert len(table) == 3 and table.total_count == 5
    assert table.count.get((0, 0, 0, 0)) == 2 and table.count.get((2, 2, 3, 3)) == 1 and table.count.get((2, 2, 2, 2)) == 2
    assert table.table[(0, 0, 0, 0)].obs_count == 10 and table.table[(2, 2, 2, 2)].obs_count == 9
    assert set(table.get_indices()) == {0, 1, 3} and len(table.get_trajectories()) == 3


def test_apply_trajectory_grid_filter():                  # :88-102
    from kbmod_b200 import search as kb
    from kbmod_b200.clustering_grid import apply_trajectory_grid_filter
    results, indices = apply_trajectory_grid_filter(_trjs(kb), bin_width=10, max_dt=1.0)
    assert len(results) == 3 and len(indices) == 3 and set(indices) == {5, 1, 3}


def test_make_coadds_simple():                            # test_stamp_utils.py:143-214
    from kbmod_b200.stamp_filters import coadd_stamps
    sci1 = np.array([[0, np.nan, np.nan], [0, np.nan, 0.5], [0, 1, 0.5]]).astype(np.float32)
    sci2 = np.array([[1, np.nan, 0.5], [1, 2, 0.5], [1, 2, 0.5]]).astype(np.float32)
    sci3 = np.array([[2, 3, 0.5], [2, 3, 0.5], [2, 3, 0.5]]).astype(np.float32)
    sci = np.array([sci1, sci2, sci3])
    var = np.array([np.full((3, 3), v).astype(np.float32) for v in (0.1, 0.2, 0.5)])
    # one trajectory resting on the middle pixel
    args = (sci, var, [0.0, 1.0, 2.0], [1.0], [1.0], [0.0], [0.0], 1)
    co = coadd_stamps(*args)
    assert np.allclose(co["sum"][0], [[3.0, 3.0, 1.0], [3.0, 5.0, 1.5], [3.0, 6.0, 1.5]], atol=1e-5)
    assert np.allclose(co["mean"][0], [[1.0, 3.0, 0.5], [1.0, 2.5, 0.5], [1.0, 2.0, 0.5]], atol=1e-5)
    assert np.allclose(co["median"][0], [[1.0, 3.0, 0.5], [1.0, 2.0, 0.5], [1.0, 2.0, 0.5]], atol=1e-5)
    assert np.allclose(co["weighted"][0], [[0.5294117647058824, 3.0, 0.5], [0.5294117647058824, 2.2857142857142856, 0.5],
                                           [0.5294117647058824, 1.5294117647058822, 0.5]], atol=1e-5)
    # third image masked out: pixel (0, 1) has no valid value at all -> 0.0
    co = coadd_stamps(*args, obs_valid=np.array([[True, True, False]]))
    assert np.allclose(co["sum"][0], [[1.0, 0.0, 0.5], [1.0, 2.0, 1.0], [1.0, 3.0, 1.0]], atol=1e-5)
    assert np.allclose(co["mean"][0], [[0.5, 0.0, 0.5], [0.5, 2.0, 0.5], [0.5, 1.5, 0.5]], atol=1e-5)
    assert np.allclose(co["median"][0], [[0.0, 0.0, 0.5], [0.0, 2.0, 0.5], [0.0, 1.0, 0.5]], atol=1e-5)
    assert np.allclose(co["weighted"][0], [[0.3333333333333333, 0.0, 0.5], [0.3333333333333333, 2.0, 0.5],
                                           [0.3333333333333333, 1.3333333333333333, 0.5]], atol=1e-5)
    # nothing selected (test_make_coadds_simple_empty, :216-224): all zero
    co = coadd_stamps(*args, obs_valid=np.array([[False, False, False]]))
    for k in co:
        assert np.array_equal(co[k][0], np.zeros((3, 3), dtype=np.float32)), k
