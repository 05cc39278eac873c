#!/usr/bin/env python
"""One vanilla-HMC launch shape for ncu: funnel, d, L, C from argv."""
import sys
import time

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from hmc_b200 import _lib, models, pyhmc  # noqa: E402

d = int(sys.argv[1]) if len(sys.argv) > 1 else 50
L = int(sys.argv[2]) if len(sys.argv) > 2 else 20
C = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 20
dev = torch.device("cuda:0")
s = pyhmc.HMC(models.Funnel(d), device=dev, seed=1)
q = (torch.rand(C, d, dtype=torch.float64) * 2 - 1).to(dev)
qo = torch.empty_like(q)
for _ in range(3):
    s.engine.hmc_step(q, L, 0.1, _lib.RULE_PYHMC, want_p=False, q_out=qo)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10):
    s.engine.hmc_step(q, L, 0.1, _lib.RULE_PYHMC, want_p=False, q_out=qo)
torch.cuda.synchronize()
ms = (time.perf_counter() - t0) / 10 * 1e3
print("d=%d L=%d C=%d: %.3f ms per transition, %.3e grad evals/s" % (d, L, C, ms, C * (L + 1) / ms * 1e3))
