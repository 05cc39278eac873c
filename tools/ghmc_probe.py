#!/usr/bin/env python
"""GHMC transition (cached V, grad V in and out) on the funnel, d and nleap from argv; HMCB_LPC picks the lane shape."""
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
q = (torch.rand(C, d, dtype=torch.float64) * 2 - 1).to(dev)
p = torch.randn(C, d, dtype=torch.float64).to(dev)
o = gh.engine.ghmc_step(q, p, L, 0.1, 0.5)
for _ in range(2):
    o2 = gh.engine.ghmc_step(o["q"], o["p"], L, 0.1, 0.5, V=o["V"], Vg=o["Vg"])
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    o2 = gh.engine.ghmc_step(o["q"], o["p"], L, 0.1, 0.5, V=o["V"], Vg=o["Vg"])
torch.cuda.synchronize()
ms = (time.perf_counter() - t0) / 5 * 1e3
print("GHMC d=%d nleap=%d: %.3f ms per transition (6 state arrays of %.0f MB move per transition = %.2f TB/s)" %
      (d, L, ms, C * d * 8 / 1e6, 6 * C * d * 8 / ms / 1e9))
