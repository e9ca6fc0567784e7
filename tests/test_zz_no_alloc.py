"""GPU test of kb_sort_trajectories_static, the entry point for callers that cannot allocate device memory
(a service whose own pool owns all of HBM, or a multi-stream loop that cannot afford the device-wide
synchronisation of cudaMalloc/cudaFree).  It sorts a device-resident trajectory list by lh using a scratch
array that is part of the library image.

The check: take a real search, whose filtered and sorted result list comes from the existing entry points
(kb_search + kb_get_results, which allocate their sort scratch), rebuild the unsorted survivor list from the raw
slots, sort it with the new entry point and require the very same records in the very same order (both use the
same stable radix sort, so even ties must agree: tolerance zero).  Done twice in a row with different data,
because all calls share one scratch array that keeps the previous call's contents.

This module is collected last (see conftest.py) and must pass alone in a fresh process as well as after all
other tests.
"""
import ctypes as C

import numpy as np
import pytest

from cases import make_case

pytestmark = pytest.mark.gpu


def _device_copy(torch, arr):
    t = torch.from_numpy(arr.view(np.uint8).copy()).cuda()
    torch.cuda.synchronize()
    return t


def _sorted_by_new_entry_point(torch, lib, records):
    dev = _device_copy(torch, records)
    rc = lib.kb_sort_trajectories_static(C.c_void_p(dev.data_ptr()), C.c_uint64(len(records)), None)
    assert rc == 0, lib.kb_last_error().decode()
    torch.cuda.synchronize()
    return dev.cpu().numpy().view(records.dtype)


def test_static_sort_matches_search_results_twice():
    torch = pytest.importorskip("torch")
    from kbmod_b200 import search as kb
    from kbmod_b200._lib import lib
    assert kb.kb_has_gpu()

    for seed, min_lh, min_obs in ((31, 0.5, 4), (32, -1.0, 0)):  # two calls in a row, different sizes and data
        case = make_case(seed=seed, T=9, H=40, W=52, n_cand=30)
        s = kb.StackSearch(case["sci"], case["var"], case["psfs"], case["times"], -1)
        s.set_min_obs(min_obs)
        s.set_min_lh(min_lh)
        s.set_results_per_pixel(4)
        s.search_all((case["vx"], case["vy"]), True)
        want = s.get_results_array()                      # existing entry points: filtered + sorted on the device
        slots = s.get_slots_array()                       # raw per-pixel slots, unsorted
        keep = ~(slots["lh"] < np.float32(min_lh)) & ~(slots["obs_count"] < min_obs)
        unsorted = np.ascontiguousarray(slots[keep])
        assert len(unsorted) == len(want) and 1000 < len(want) <= (1 << 18)
        got = _sorted_by_new_entry_point(torch, lib, unsorted)
        assert got.tobytes() == want.tobytes(), "static-scratch sort differs from the search's own sorted list"
        assert np.all(np.diff(got["lh"]) <= 0)


def test_static_sort_rejects_oversized_lists_and_handles_trivial_ones():
    torch = pytest.importorskip("torch")
    from kbmod_b200._lib import TRAJECTORY_DTYPE, lib
    one = np.zeros(1, dtype=TRAJECTORY_DTYPE)
    one["lh"] = 3.0
    assert _sorted_by_new_entry_point(torch, lib, one)["lh"][0] == 3.0
    dummy = torch.zeros(64, dtype=torch.uint8, device="cuda")
    rc = lib.kb_sort_trajectories_static(C.c_void_p(dummy.data_ptr()), C.c_uint64((1 << 18) + 1), None)
    assert rc != 0 and b"exceed" in lib.kb_last_error()


def test_static_sort_treats_plus_and_minus_zero_as_a_tie():
    """+0.0 and -0.0 compare equal in the reference's sort, so they must keep their input order like any other tie
    (and not sort +0 ahead of -0 by their bit patterns)."""
    torch = pytest.importorskip("torch")
    from kbmod_b200._lib import TRAJECTORY_DTYPE, lib
    n = 5000
    rec = np.zeros(n, dtype=TRAJECTORY_DTYPE)
    rng = np.random.default_rng(7)
    rec["lh"] = rng.choice(np.array([0.0, -0.0, 1.5, -2.0], dtype=np.float32), n)
    rec["x"] = np.arange(n)                       # input position
    got = _sorted_by_new_entry_point(torch, lib, rec)
    order = np.argsort(-rec["lh"].astype(np.float64), kind="stable")   # -0.0 == 0.0 for numpy's comparison too
    assert np.array_equal(got["x"], rec["x"][order])
    zeros = got[got["lh"] == 0.0]
    assert np.all(np.diff(zeros["x"]) > 0) and np.signbit(zeros["lh"]).any() and (~np.signbit(zeros["lh"])).any()
