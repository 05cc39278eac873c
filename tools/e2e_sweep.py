#!/usr/bin/env python
"""End-to-end DR step with host buffers (Engine.drhmc_step_host): chunk count and compute-stream count sweep."""
import json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import algorithms, models  # noqa: E402
dev = torch.device("cuda:0")
d, C = 50, 1 << 20
dr = algorithms.DRHMC(models.Funnel(d), device=dev, seed=3)
e = dr.engine
e._ensure()
g = torch.Generator().manual_seed(0)
q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
q = e.drhmc_sample(q, 150, 0.2, 3, 2.0, 100, want_avec=False)["q"]
qh = torch.empty((C, d), dtype=torch.float64).pin_memory()
qh.copy_(q)
for n_chunks, rs in ((8, 2), (4, 2), (6, 2), (12, 2), (16, 2), (8, 1), (8, 3), (12, 3), (16, 3), (24, 3)):
    h = e.drhmc_step_host(qh, 150, 0.2, 3, 2.0, want_p=False, n_chunks=n_chunks, run_streams=rs)
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(5):
        h = e.drhmc_step_host(h["q"], 150, 0.2, 3, 2.0, want_p=False, n_chunks=n_chunks, run_streams=rs)
    torch.cuda.synchronize()
    print(json.dumps({"n_chunks": n_chunks, "run_streams": rs, "ms": (time.time() - t0) / 5 * 1e3}), flush=True)
