#!/usr/bin/env python
"""funnel d=50 vanilla HMC: one hmcb_hmc_step call per transition against hmcb_hmc_sample (K transitions per launch)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import _lib, pyhmc, models  # noqa: E402
from tools.dr_sample_bench import fp64_peak  # noqa: E402

dev = torch.device("cuda:0")
d, C = 50, 1 << 20
eng = pyhmc.HMC(models.Funnel(d), device=dev, seed=1).engine
eng._ensure()
peak = fp64_peak(eng, dev)
g = torch.Generator().manual_seed(0)
q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
nxt = torch.empty_like(q)
for L, eps in ((20, 0.1), (150, 0.2), (5, 0.1)):
    for K in (1, 4, 16, 64):
        def run():
            eng.hmc_sample(q, L, eps, _lib.RULE_PYHMC, K, q_out=nxt)
        run(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        tf = C * (L + 1) * 7 * d / (ms * 1e-3) / 1e12
        print(json.dumps({"L": L, "eps": eps, "K": K, "ms_per_transition": ms, "frac_of_fp64_peak": tf / peak}), flush=True)
    def single():
        eng.hmc_step(q, L, eps, _lib.RULE_PYHMC, want_p=False, q_out=nxt)
    single(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(8): single()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 8
    print(json.dumps({"L": L, "single_call_ms": ms, "frac": C * (L + 1) * 7 * d / (ms * 1e-3) / 1e12 / peak}), flush=True)
