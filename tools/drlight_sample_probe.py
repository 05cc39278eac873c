#!/usr/bin/env python
"""DRHMC nleap=20 eps=0.1 nprops=3 (config 3, light): hmcb_drhmc_sample's short-trajectory mode against the pipelined
stage kernels (HMCB_NO_DR_SHORT=1) and single calls, no running moments."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import algorithms, models  # noqa: E402
from tools.dr_sample_bench import fp64_peak  # noqa: E402
dev = torch.device("cuda:0")
d = 50
for C in (1 << 20, 1 << 17):
    g = torch.Generator().manual_seed(0)
    q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
    eng = algorithms.DRHMC(models.Funnel(d), device=dev, seed=1).engine
    eng._ensure()
    peak = fp64_peak(eng, dev)
    q = eng.drhmc_sample(q, 20, 0.1, 3, 2.0, 64, want_avec=False)["q"]
    nxt = torch.empty_like(q)
    counts = torch.zeros((C, 3), dtype=torch.int32, device=dev)
    for K in (16,):
        for env in ("0", "1", "c2", "c3", "c4", "c5", "c6"):
            os.environ.pop("HMCB_NO_DR_SHORT", None); os.environ.pop("HMCB_DR_SHORT_CHUNK", None)
            if env == "1": os.environ["HMCB_NO_DR_SHORT"] = "1"
            elif env.startswith("c"): os.environ["HMCB_DR_SHORT_CHUNK"] = env[1:]
            def run():
                eng.drhmc_sample(q, 20, 0.1, 3, 2.0, K, q_out=nxt, want_avec=False, counts_out=counts)
            run(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); run(); e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / K
            grads = float(counts[:, 1].sum(dtype=torch.float64)) / K
            print(json.dumps({"chains": C, "K": K, "mode": {"0": "short, chunk 8", "1": "pipelined"}.get(env, "short, chunk " + env[1:]), "ms_per_transition": ms, "grad_evals_per_chain": grads / C,
                              "frac_of_fp64_peak": grads * 350 / (ms * 1e-3) / 1e12 / peak}), flush=True)
    os.environ.pop("HMCB_NO_DR_SHORT", None)
