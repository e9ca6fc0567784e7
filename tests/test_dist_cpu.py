"""CPU tests of the multi-GPU host logic (world_size 2 over gloo): row-band sharding of the start pixels and the
merge rule.  Each rank computes its band with the oracle (standing in for its GPU) and the ranks exchange their
sorted lists over gloo; the merged list must equal the single-process full-frame list.  The NCCL data path itself
(kb_merge_results_allgather) is exercised on the GPU box by bench.py --gpus N."""
import os
import sys

import numpy as np
import pytest

from cases import make_case

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_row_bands_cover_exactly():
    from kbmod_b200.dist import row_bands
    for (y0, y1, w) in [(0, 2048, 8), (-10, 37, 4), (3, 5, 4), (0, 7, 2), (5, 6, 3)]:
        bands = row_bands(y0, y1, w)
        assert len(bands) == w and bands[0][0] == y0 and bands[-1][1] == y1
        assert all(bands[i][1] == bands[i + 1][0] for i in range(w - 1))
        sizes = [b - a for a, b in bands]
        assert min(sizes) >= 0 and max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    from kbmod_b200 import dist
    from oracle import cpu as orc
    comm = dist.Comm(rank, world, init_nccl=False, backend="gloo")
    case = make_case(seed=21, T=8, H=30, W=36, n_cand=24)
    st = orc.fill_from_images(case["sci"], case["var"], case["psfs"], case["times"], -1)
    y0, y1 = -2, case["H"] + 3
    band = dist.row_bands(y0, y1, world)[rank]
    p = orc.make_params(3, 0.5, (0.25, 0.75, 0.7413), (0, case["W"], band[0], band[1]), 4, st)
    mine = orc.search_all(st, p, case["vx"], case["vy"])
    lists = comm.all_gather_arrays(mine)
    merged = dist.merge_sorted_lists(lists, top_n=0)
    top = dist.merge_sorted_lists(lists, top_n=50)
    worst = comm.max_float(float(rank))
    comm.barrier()
    np.save(os.path.join(out_dir, f"merged_{rank}.npy"), merged)
    np.save(os.path.join(out_dir, f"top_{rank}.npy"), top)
    assert worst == world - 1
    comm.finalize()


def test_two_rank_shard_and_merge_equals_single_process(tmp_path):
    torch = pytest.importorskip("torch")
    import torch.multiprocessing as mp
    from oracle import cpu as orc
    world, port = 2, 29500 + (os.getpid() % 400)
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    case = make_case(seed=21, T=8, H=30, W=36, n_cand=24)
    st = orc.fill_from_images(case["sci"], case["var"], case["psfs"], case["times"], -1)
    p = orc.make_params(3, 0.5, (0.25, 0.75, 0.7413), (0, case["W"], -2, case["H"] + 3), 4, st)
    want = orc.search_all(st, p, case["vx"], case["vy"])
    for r in range(world):
        got = np.load(tmp_path / f"merged_{r}.npy")
        assert got.tobytes() == want.tobytes(), f"rank {r}: merged list differs from the single-process list"
        top = np.load(tmp_path / f"top_{r}.npy")
        assert top.tobytes() == want[:50].tobytes()


def test_parse_cpulist():
    from kbmod_b200.dist import parse_cpulist
    assert parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert parse_cpulist("5") == {5}
    assert parse_cpulist("") == set()
    with pytest.raises(ValueError):
        parse_cpulist("a-b")


def test_band_rows_cover_the_reach_of_a_band():
    from kbmod_b200.dist import band_rows
    vy = np.array([-3.5, 0.0, 2.0], dtype=np.float32)
    times = [0.0, 1.0, 4.0]
    lo, hi = band_rows((100, 200), vy, times, 1000)
    assert lo <= 100 - 14 and hi >= 200 + 14 and lo >= 0 and hi <= 1000     # 3.5 px/day * 4 days
    assert band_rows((0, 50), vy, times, 60) == (0, 60)                       # clamped to the image
    assert band_rows((10, 20), np.zeros(0, dtype=np.float32), times, 100)[0] <= 10


def test_bench_band_inputs_are_consistent_between_ranks():
    """bench.make_band_inputs (the strong-scaling named config): two ranks that need overlapping rows generate the same
    pixels for them, without either generating the whole frame."""
    sys.path.insert(0, ROOT)
    import bench
    c = dict(bench.CONFIGS["C4"], H=320, W=96, T=5)
    a, vx, vy = bench.make_band_inputs(c, (40, 200))
    b, vx2, _ = bench.make_band_inputs(c, (150, 320))
    assert len(vx) == 20000 and np.array_equal(vx, vx2)
    for t in range(5):
        assert np.array_equal(a.sci[t][150:200], b.sci[t][150:200], equal_nan=True)
        assert np.array_equal(a.var[t][150:200], b.var[t][150:200], equal_nan=True)
    assert np.isnan(a.sci[0][40:200]).mean() > 0.003 and a.objects == b.objects
