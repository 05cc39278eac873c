#!/usr/bin/env python
"""single-call ws kernel with and without the shared-memory q0 copy (HMCB_WS_Q0S)"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import _lib, pyhmc, algorithms, models  # noqa: E402
dev = torch.device("cuda:0")
d, C = 50, 1 << 20
g = torch.Generator().manual_seed(0)
q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
nxt = torch.empty_like(q)
for env in ("0", "1"):
    os.environ["HMCB_WS_Q0S"] = env
    eng = pyhmc.HMC(models.Funnel(d), device=dev, seed=1).engine
    dr = algorithms.DRHMC(models.Funnel(d), device=dev, seed=1).engine
    res = {}
    for L, eps in ((20, 0.1), (150, 0.2)):
        def single():
            eng.hmc_step(q, L, eps, _lib.RULE_PYHMC, want_p=False, q_out=nxt)
        single(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(8): single()
        e1.record(); torch.cuda.synchronize()
        res["hmc L=%d" % L] = e0.elapsed_time(e1) / 8
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    for ev in evs: ev.record()
    tot = [0, 0, 0]
    for it in range(5):
        dr.drhmc_step(q, 150, 0.2, 3, 2.0, want_p=False, q_out=nxt, stage_events=evs, want_avec=False)
        torch.cuda.synchronize()
        if it: tot = [t + evs[j].elapsed_time(evs[j + 1]) for j, t in enumerate(tot)]
    res["dr stages"] = [t / 4 for t in tot]
    o1 = eng.hmc_step(q, 20, 0.1, _lib.RULE_PYHMC)
    res["checksum"] = float(o1["q"].sum()) 
    print("HMCB_WS_Q0S=%s" % env, json.dumps(res), flush=True)
