#!/usr/bin/env python
"""MALA on Rosenbrock d=100 (config 5) under different lanes-per-chain shapes (HMCB_LPC) and fused step counts."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import mala, models  # noqa: E402
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
for d in (100, 20):
    C = 131072
    q = (1.0 + 0.1 * torch.randn(C, d, generator=g, dtype=torch.float64)).to(dev)
    for lpc in ("", "2", "4", "8", "32"):
        if lpc: os.environ["HMCB_LPC"] = lpc
        else: os.environ.pop("HMCB_LPC", None)
        m = mala.MALA(models.Rosenbrock(d), device=dev, seed=4).engine
        for K in (32,):
            try:
                m.mala_step(q, 0.02, n_steps=K); torch.cuda.synchronize()
            except Exception as e:
                print(json.dumps({"d": d, "lpc": lpc, "error": str(e)[:80]})); continue
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(4): o = m.mala_step(q, 0.02, n_steps=K)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 4 / K
            print(json.dumps({"d": d, "lpc": lpc or "auto", "K": K, "ms_per_step": ms, "normals_per_s": C * d / (ms * 1e-3), "checksum": float(o["q"].sum())}), flush=True)
