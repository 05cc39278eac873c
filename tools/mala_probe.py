#!/usr/bin/env python
"""One MALA launch of the cfg5 shape (Rosenbrock, 131,072 chains, 32 fused steps) for ncu."""
import sys
import time

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from hmc_b200 import mala, models  # noqa: E402

d = int(sys.argv[1]) if len(sys.argv) > 1 else 100
K = int(sys.argv[2]) if len(sys.argv) > 2 else 32
dev = torch.device("cuda:0")
C = 131072
m = mala.MALA(models.Rosenbrock(d), device=dev, seed=4)
q = (1.0 + 0.1 * torch.randn(C, d, dtype=torch.float64)).to(dev)
for _ in range(3):
    out = m.engine.mala_step(q, 0.02, n_steps=K)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    out = m.engine.mala_step(q, 0.02, n_steps=K)
torch.cuda.synchronize()
ms = (time.perf_counter() - t0) / 5 * 1e3
print("d=%d K=%d: %.3f ms per launch, %.3e chain-steps/s, %.3e normals/s" % (d, K, ms, C * K / ms * 1e3, C * K * d / ms * 1e3))
