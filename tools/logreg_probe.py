#!/usr/bin/env python
"""One cfg4 transition (Bayesian logistic regression, N=100k, D=128, 16,384 chains, HMC L=10) through the
Python class -> C ABI, for ncu and for a launch count."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from hmc_b200 import _lib, algorithms, models  # noqa: E402

dev = torch.device("cuda:0")
N, D, C, L = 100000, 128, 16384, 10
rng = np.random.default_rng(3)
X = (rng.standard_normal((N, D)) / np.sqrt(D)).astype(np.float32)
ths = rng.standard_normal(D)
y = (rng.random(N) < 1 / (1 + np.exp(-(X.astype(np.float64) @ ths)))).astype(np.float32)
lm = algorithms.HMC(models.LogisticRegression(X, y, 1.0), device=dev, seed=3)
q = (0.1 * torch.randn(C, D, generator=torch.Generator().manual_seed(1))).to(dev)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
lm.engine.hmc_step(q, L, 0.01, _lib.RULE_ALG, want_p=False)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(reps):
    out = lm.engine.hmc_step(q, L, 0.01, _lib.RULE_ALG, want_p=False)
torch.cuda.synchronize()
ms = (time.perf_counter() - t0) / reps * 1e3
print("cfg4: %.2f ms per transition, accept %.3f" % (ms, float((out["code"] == 1).double().mean())))
