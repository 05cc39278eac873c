#!/usr/bin/env python
"""DRHMC nleap=20 (the light variant of config 3), one call per transition, under different kernel choices (env)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import algorithms, models  # noqa: E402
dev = torch.device("cuda:0")
d, C = 50, 1 << 20
g = torch.Generator().manual_seed(0)
q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
nxt = torch.empty_like(q)
counts = torch.zeros((C, 3), dtype=torch.int32, device=dev)
for env in ({}, {"HMCB_WS_Q0S": "2"}, {"HMCB_NO_WS": "1"}, {"HMCB_DR_SPLIT": "1"}, {"HMCB_DR_SPLIT": "1", "HMCB_WS_Q0S": "2"}):
    for k in ("HMCB_WS_Q0S", "HMCB_NO_WS", "HMCB_DR_SPLIT"):
        os.environ.pop(k, None)
    os.environ.update(env)
    dr = algorithms.DRHMC(models.Funnel(d), device=dev, seed=1).engine
    def one():
        dr.drhmc_step(q, 20, 0.1, 3, 2.0, want_p=False, q_out=nxt, counts_out=counts, want_avec=False)
    one(); torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(8): one()
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 8)
    print(json.dumps({"env": env, "ms": best, "checksum": float(nxt.sum())}), flush=True)
