#!/usr/bin/env python
"""GHMC on the funnel (d and nleap from argv): one hmcb_ghmc_step per transition (cached V, grad V in and out) against
hmcb_ghmc_sample with K transitions per call; HMCB_LPC picks the lane shape."""
import json
import sys
import time

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from hmc_b200 import algorithms, models  # noqa: E402

d = int(sys.argv[1]) if len(sys.argv) > 1 else 50
L = int(sys.argv[2]) if len(sys.argv) > 2 else 20
C = 1 << 20
dev = torch.device("cuda:0")
gh = algorithms.GHMC(models.Funnel(d), device=dev, seed=1)
e = gh.engine
q = (torch.rand(C, d, dtype=torch.float64) * 2 - 1).to(dev)
p = torch.randn(C, d, dtype=torch.float64).to(dev)
o = e.ghmc_sample(q, p, L, 0.1, 0.5, 30)


def timed(fn, n):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


ms1 = timed(lambda: e.ghmc_step(o["q"], o["p"], L, 0.1, 0.5, V=o["V"], Vg=o["Vg"]), 5)
row = {"d": d, "nleap": L, "chains": C, "ms_single": round(ms1, 4), "single_TBps_6_arrays": round(6 * C * d * 8 / ms1 / 1e9, 2)}
flops = C * L * 7 * d           # L gradients per transition with the caches
row["single_fp64_tflops"] = round(flops / ms1 / 1e9, 2)
for K in (2, 4, 8, 16, 32):
    ms = timed(lambda: e.ghmc_sample(o["q"], o["p"], L, 0.1, 0.5, K, V=o["V"], Vg=o["Vg"]), 3) / K
    row["ms_K%d" % K] = round(ms, 4)
    row["tflops_K%d" % K] = round(flops / ms / 1e9, 2)
print(json.dumps(row), flush=True)
