#!/usr/bin/env python
"""ncu target: cfg5 Rosenbrock d=100, 131,072 chains, 8 HMC transitions (L=20) in one launch (k_hmc_multi); the second
launch is bracketed by cudaProfilerStart/Stop (run ncu with --profile-from-start off)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import _lib, algorithms, models  # noqa: E402
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
d, C = 100, 131072
eng = algorithms.HMC(models.Rosenbrock(d), device=dev, seed=4).engine
q = (1.0 + 0.1 * torch.randn(C, d, generator=g, dtype=torch.float64)).to(dev)
nxt = torch.empty_like(q)
eng.hmc_sample(q, 20, 0.02, _lib.RULE_ALG, 8, q_out=nxt); torch.cuda.synchronize()
torch.cuda.profiler.start()
eng.hmc_sample(q, 20, 0.02, _lib.RULE_ALG, 8, q_out=nxt); torch.cuda.synchronize()
torch.cuda.profiler.stop()
