"""The reference's own tests of the search path (tests/test_search.py and tests/test_stack_search_results.py of
kbmod @ 617cd33), restated with kbmod_b200's names and run on the device (-m gpu): same stack, same generator, same
thresholds.  Where the reference test exercises the CPU path (use_gpu=False) the device path is used, and where it
expects the 200-image limit the opposite is checked (kbmod_b200 has no such limit)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

PIXEL_ERR, VEL_ERR, FLUX_ERR = 1, 0.1, 0.15     # test_search.py:14-17
IMG_COUNT, DIM_Y, DIM_X, NOISE, FLUX = 20, 80, 60, 4.0, 250.0
START_X, START_Y, VX, VY = 17, 12, 21.0, 16.0


@pytest.fixture(scope="module")
def env():                                       # test_search.py:12-98 (setUp)
    from kbmod_b200 import fake_data as fd, search as kb
    times = np.array([i / IMG_COUNT for i in range(IMG_COUNT)])
    stack = fd.make_fake_image_stack(DIM_Y, DIM_X, times, noise_level=NOISE, psf_val=1.0, rng=np.random.default_rng(100))
    fd.add_fake_object(stack, START_X, START_Y, VX, VY, flux=FLUX)
    for i in range(IMG_COUNT):                   # one pixel masked in every other image
        if i % 2 == 0:
            stack.sci[i][5, 6] = np.nan
            stack.var[i][5, 6] = np.nan
    gen = fd.KBMODV1Search(150, 5.0, 40.0, 150, 0.0, 1.5)
    return dict(kb=kb, fd=fd, stack=stack, cands=[t for t in gen])


def new_search(env):
    s = env["stack"]
    return env["kb"].StackSearch(s.sci, s.var, s.psfs, s.zeroed_times)


def check_best(best, x=START_X, y=START_Y, vx=VX, vy=VY, flux=FLUX):
    assert abs(best.x - x) <= PIXEL_ERR and abs(best.y - y) <= PIXEL_ERR
    assert abs(best.vx / vx - 1) <= VEL_ERR and abs(best.vy / vy - 1) <= VEL_ERR
    if flux is not None:
        assert abs(best.flux / flux - 1) <= FLUX_ERR


def test_evaluate_single_trajectory(env):        # :100-112 and :114-127, on the device
    kb = env["kb"]
    search = new_search(env)
    trj = kb.Trajectory(x=START_X, y=START_Y, vx=VX, vy=VY)
    search.evaluate_single_trajectory(trj, True)
    assert trj.obs_count > 0 and trj.flux > 0.0 and trj.lh > 0.0
    trj2 = search.search_linear_trajectory(START_X, START_Y, VX, VY, True)
    assert trj2.obs_count > 0 and trj2.flux > 0.0 and trj2.lh > 0.0
    with pytest.raises(RuntimeError):            # no CPU path in the product
        search.evaluate_single_trajectory(trj, False)


def test_results_gpu(env):                       # :150-169
    search = new_search(env)
    search.search_all(env["cands"], True)
    expected_size = 8 * DIM_X * DIM_Y
    assert search.compute_max_results() == expected_size
    results = search.get_results(0, 10 * expected_size)
    assert 0 < len(results) <= expected_size
    check_best(results[0])
    with pytest.raises(RuntimeError):            # :129-148 is the CPU search: not in the product
        search.search_all(env["cands"], False)


def test_results_extended_bounds(env):           # :171-197
    search = new_search(env)
    search.set_results_per_pixel(5)
    search.set_start_bounds_x(-10, DIM_X + 10)
    search.set_start_bounds_y(-10, DIM_Y + 10)
    expected = (DIM_X + 20) * (DIM_Y + 20) * 5
    assert search.compute_max_results() == expected
    search.search_all(env["cands"], True)
    results = search.get_results(0, 10 * expected)
    assert 0 < len(results) <= expected
    check_best(results[0])


def test_results_reduced_bounds(env):            # :199-221 -- the exact result count
    search = new_search(env)
    search.set_results_per_pixel(10)
    search.set_start_bounds_x(5, DIM_X - 5)
    search.set_start_bounds_y(5, DIM_Y - 5)
    expected = (DIM_X - 10) * (DIM_Y - 10) * 10
    assert search.compute_max_results() == expected
    search.search_all(env["cands"], True)
    results = search.get_results(0, 10 * expected)
    assert len(results) == expected
    check_best(results[0])


def test_invalid_start_bounds_and_sigmag_config(env):   # :223-234
    search = new_search(env)
    with pytest.raises(RuntimeError):
        search.set_start_bounds_x(6, 5)
    with pytest.raises(RuntimeError):
        search.set_start_bounds_y(-1, -5)
    search.enable_gpu_sigmag_filter([0.25, 0.75], 0.5, 1.0)
    for bad in ([0.25], [0.75, 0.25], [-0.01, 0.75], [0.75, 1.10]):
        with pytest.raises(RuntimeError):
            search.enable_gpu_sigmag_filter(bad, 0.5, 1.0)
    with pytest.raises(RuntimeError):
        search.enable_gpu_sigmag_filter([0.25, 0.75], -0.5, 1.0)


def test_results_off_chip(env):                  # :236-268
    kb, fd = env["kb"], env["fd"]
    times = np.array([i / IMG_COUNT for i in range(IMG_COUNT)])
    stack = fd.make_fake_image_stack(DIM_Y, DIM_X, times, noise_level=NOISE, psf_val=1.0, rng=np.random.default_rng(100))
    fd.add_fake_object(stack, -3, 12, 25.0, 10.0, flux=FLUX)
    search = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times)
    search.set_start_bounds_x(-10, DIM_X + 10)
    search.set_start_bounds_y(-10, DIM_Y + 10)
    search.search_all(env["cands"], True)
    check_best(search.get_results(0, 10)[0], x=-3, y=12, vx=25.0, vy=10.0, flux=None)


def test_stack_search_set_min_obs(env):          # :270-277
    search = new_search(env)
    search.set_min_obs(1)
    search.set_min_obs(IMG_COUNT)
    with pytest.raises(RuntimeError):
        search.set_min_obs(-1)
    with pytest.raises(RuntimeError):
        search.set_min_obs(IMG_COUNT + 1)


def test_search_many_images(env):                # :279-304 expects RuntimeError at 1000 images; here it works
    kb, fd = env["kb"], env["fd"]
    num_times = 1000
    times = np.array([i / num_times for i in range(num_times)])
    stack = fd.make_fake_image_stack(20, 10, times, noise_level=0.5, psf_val=1.0, rng=np.random.default_rng(1))
    search = kb.StackSearch(stack.sci, stack.var, stack.psfs, stack.zeroed_times)
    trj = kb.Trajectory(x=0, y=0, vx=0.0, vy=0.0)
    search.search_all([trj], True)
    assert search.get_number_total_results() > 0
    search.evaluate_single_trajectory(trj, True)
    assert trj.obs_count == num_times


def test_psi_phi_curves_known(env):              # test_stack_search_results.py:87-118
    kb = env["kb"]
    height, width, num_times = 5, 4, 5
    times = np.arange(num_times, dtype=np.float32)
    sci = [np.full((height, width), float(i), dtype=np.float32) for i in range(num_times)]
    var = [np.full((height, width), 0.1, dtype=np.float32) for _ in range(num_times)]
    psf = [np.array([[1.0]], dtype=np.float32) for _ in range(num_times)]
    search = kb.StackSearch(sci, var, psf, times - times[0])
    psi_phi = search.get_all_psi_phi_curves([kb.Trajectory(x=2, y=2, vx=0.0, vy=0.0)])
    assert psi_phi.shape == (1, 2 * num_times)
    assert np.allclose(psi_phi[0, :num_times], [float(i) / 0.1 for i in range(num_times)])
    assert np.allclose(psi_phi[0, num_times:], [1.0 / 0.1] * num_times)


def test_set_get_results(env):                   # test_stack_search_results.py:40-72
    kb = env["kb"]
    search = new_search(env)
    trjs = [kb.Trajectory(x=i, y=i, vx=0.0, vy=0.0, lh=float(10 - i), obs_count=5) for i in range(10)]
    search.set_results(trjs)
    assert search.get_number_total_results() == 10
    res = search.get_results(2, 2)
    assert len(res) == 2 and res[0].x == 2 and res[1].x == 3
    res = search.get_results(8, 20)               # an over-run truncates
    assert len(res) == 2 and res[0].x == 8 and res[1].x == 9
    with pytest.raises(RuntimeError):
        search.get_results(0, 0)
    search.clear_results()
    assert len(search.get_all_results()) == 0
