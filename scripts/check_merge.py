"""Run under torchrun on N GPUs: sharded search + NCCL all-gather merge must equal the single-GPU search."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from cases import make_case
from kbmod_b200 import dist, search as kb

rank, world, local = dist.env_rank()
dist.set_device(local)
comm = dist.Comm(rank, world)
case = make_case(seed=41, T=12, H=90, W=70, n_cand=60)
s = kb.StackSearch(case["sci"], case["var"], case["psfs"], case["times"], -1)
s.set_min_obs(4); s.enable_gpu_sigmag_filter([0.25, 0.75], 0.7413, 1.0)
s.search_all((case["vx"], case["vy"]), True)
full = s.get_results_array().copy()
n = dist.sharded_search(s, (case["vx"], case["vy"]), comm, top_n=0)
merged = s.get_results_array()
ok = (n == len(full)) and merged.tobytes() == full.tobytes()
n2 = dist.sharded_search(s, (case["vx"], case["vy"]), comm, top_n=100)
top = s.get_results_array()
ok2 = top.tobytes() == full[:100].tobytes()
print(f"rank {rank}/{world}: merged {n} results, identical to single-GPU: {ok}; top-100 identical: {ok2}", flush=True)
comm.finalize()
sys.exit(0 if (ok and ok2) else 1)
