#!/usr/bin/env python
"""Timeline of Engine.drhmc_step_host's chunk pipeline at the headline shape: per chunk, when its copy-in ended, when its
stage kernels started and ended, and when its copy-out ended (ms from the start of the call)."""
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
for ramp in ("", "0.25,0.5,0.75;0.5", "0.25,0.5;0.25", "0.3,0.7;0.5,0.5", "0.25,0.75;1.0", "0.5;1.0"):
    nc, rs = 8, 2
    os.environ["HMCB_E2E_RAMP"] = ramp
    h = e.drhmc_step_host(qh, 150, 0.2, 3, 2.0, want_p=False, n_chunks=nc, run_streams=rs)
    for _ in range(3):
        h = e.drhmc_step_host(h["q"], 150, 0.2, 3, 2.0, want_p=False, n_chunks=nc, run_streams=rs)
    torch.cuda.synchronize()
    e.host_timeline = []
    t0 = time.perf_counter()
    h = e.drhmc_step_host(h["q"], 150, 0.2, 3, 2.0, want_p=False, n_chunks=nc, run_streams=rs)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    tl, e.host_timeline = e.host_timeline, None
    z = tl[0]
    rows = [[r[0], r[1]] + [round(z.elapsed_time(x), 3) for x in r[2:]] for r in tl[1:]]
    print(json.dumps({"ramp": ramp, "n_chunks": nc, "run_streams": rs, "wall_ms": round(wall, 3), "rows[idx,chains,in_done,run_start,run_done,out_done]": rows}), flush=True)
