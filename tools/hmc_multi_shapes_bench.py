#!/usr/bin/env python
"""K-fused HMC (hmcb_hmc_sample) against single calls on the shapes of BASELINE config 5 and the funnel."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import _lib, pyhmc, algorithms, models  # noqa: E402
from tools.dr_sample_bench import fp64_peak  # noqa: E402
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
peak = None
for name, mk, d, C, L, eps, fpg in (("rosenbrock", models.Rosenbrock, 100, 131072, 20, 0.02, 8.5), ("rosenbrock", models.Rosenbrock, 10, 131072, 20, 0.02, 8.5),
                                    ("rosenbrock", models.Rosenbrock, 2, 131072, 20, 0.02, 8.5),
                                    ("funnel", models.Funnel, 10, 1 << 20, 20, 0.1, 7), ("funnel", models.Funnel, 100, 1 << 19, 20, 0.1, 7),
                                    ("funnel", models.Funnel, 50, 1 << 20, 20, 0.1, 7)):
    eng = algorithms.HMC(mk(d), device=dev, seed=4).engine
    eng._ensure()
    if peak is None: peak = fp64_peak(eng, dev)
    q = torch.randn(C, d, generator=g, dtype=torch.float64)
    q = ((1.0 + 0.1 * q) if name == "rosenbrock" else 0.5 * q).to(dev)
    nxt = torch.empty_like(q)
    def t(fn, n):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    ms1 = t(lambda: eng.hmc_step(q, L, eps, _lib.RULE_ALG, want_p=False, q_out=nxt), 8)
    K = 16
    msK = t(lambda: eng.hmc_sample(q, L, eps, _lib.RULE_ALG, K, q_out=nxt), 2) / K
    f = lambda ms: C * (L + 1) * fpg * d / (ms * 1e-3) / 1e12 / peak
    print(json.dumps({"model": name, "d": d, "chains": C, "L": L, "single_ms": ms1, "single_frac": f(ms1), "fused_K16_ms": msK, "fused_frac": f(msK)}), flush=True)
