#!/usr/bin/env python
"""What the running moments cost inside hmcb_drhmc_sample at the headline parameters (with / without msum)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import algorithms, models  # noqa: E402
dev = torch.device("cuda:0")
d, C, K = 50, 1 << 20, 20
eng = algorithms.DRHMC(models.Funnel(d), device=dev, seed=2).engine
g = torch.Generator().manual_seed(2)
q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
q = eng.drhmc_sample(q, 150, 0.2, 3, 2.0, 200, want_avec=False)["q"]
nxt = torch.empty_like(q)
s1, s2 = torch.zeros_like(q), torch.zeros_like(q)
counts = torch.zeros((C, 3), dtype=torch.int32, device=dev)
for name, kw in (("with moments", dict(msum=s1, msumsq=s2)), ("without", {}),                  ("with moments, no fold (HMCB_NO_TILE)", dict(msum=s1, msumsq=s2)), ("without, HMCB_NO_TILE", {})):
    os.environ.pop("HMCB_DEBUG_SKIP_LIST_MOMENTS", None); os.environ.pop("HMCB_NO_TILE", None)
    if "NO_TILE" in name: os.environ["HMCB_NO_TILE"] = "1"
    def run():
        eng.drhmc_sample(q, 150, 0.2, 3, 2.0, K, q_out=nxt, want_avec=False, counts_out=counts, **kw)
    run(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(); e1.record(); torch.cuda.synchronize()
    print(json.dumps({"form": name, "ms_per_transition": e0.elapsed_time(e1) / K}), flush=True)
