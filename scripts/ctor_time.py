import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kbmod_b200 import fake_data as fd, search as kb
T, H, W = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
rng = np.random.default_rng(1)
base = rng.normal(0, 2, (H, W)).astype(np.float32)
sci = [base + np.float32(i) for i in range(T)]
var = [np.full((H, W), 4.0, np.float32) for _ in range(T)]
psf = fd.make_gaussian_kernel(1.0)
for nb in (-1, 2):
    t0 = time.perf_counter()
    s = kb.StackSearch(sci, var, [psf] * T, np.arange(T) * 0.25, nb)
    print(f"T={T} {H}x{W} num_bytes={nb}: ctor {time.perf_counter() - t0:.2f} s", flush=True)
    del s
