#!/usr/bin/env python
"""The bench line's timed region and nothing else, for the ncu launch list: 2 transitions of the headline workload through
one hmcb_drhmc_sample call (as bench.py times it), then the same 2 transitions as single hmcb_drhmc_step calls + moments."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import algorithms, models, utils  # noqa: E402

dev = torch.device("cuda:0")
d, C, K = 50, 1 << 20, 2
eng = algorithms.DRHMC(models.Funnel(d), device=dev, seed=2).engine
g = torch.Generator().manual_seed(2)
q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
nxt = torch.empty_like(q)
mom = utils.RunningMoments(eng, C, d)
counts = torch.zeros((C, 3), dtype=torch.int32, device=dev)
eng.drhmc_sample(q, 150, 0.2, 3, 2.0, K, msum=mom.sum, msumsq=mom.sumsq, q_out=nxt, want_avec=False, counts_out=counts)
torch.cuda.synchronize()
for k in range(K):
    eng.drhmc_step(nxt, 150, 0.2, 3, 2.0, want_p=False, q_out=q, want_avec=False, counts_out=counts)
    mom.add(q)
    q, nxt = nxt, q
torch.cuda.synchronize()
print("done")
