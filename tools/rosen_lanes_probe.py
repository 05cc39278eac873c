#!/usr/bin/env python
"""cfg5 Rosenbrock d=100 (131,072 chains, HMC L=20): single launches and 16 transitions per launch under the compiled
lanes-per-chain shapes (HMCB_LPC = 2 / 4 / 8 / 32; the pairwise gradient needs no cross-lane sum inside the trajectory)."""
import json, os, subprocess, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) == 1:
    for lpc in ("", "2", "4", "8", "32"):
        env = dict(os.environ)
        env.pop("HMCB_LPC", None)
        if lpc:
            env["HMCB_LPC"] = lpc
        subprocess.run([sys.executable, __file__, lpc or "default"], env=env)
    sys.exit(0)
from hmc_b200 import _lib, algorithms, models  # noqa: E402
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
for d in (100, 20):
    C, L, eps = 131072, 20, 0.02
    eng = algorithms.HMC(models.Rosenbrock(d), device=dev, seed=4).engine
    eng._ensure()
    q = (1.0 + 0.1 * torch.randn(C, d, generator=g, dtype=torch.float64)).to(dev)
    nxt = torch.empty_like(q)

    def t(fn, n):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    try:
        ms1 = t(lambda: eng.hmc_step(q, L, eps, _lib.RULE_ALG, want_p=False, q_out=nxt), 20)
        msK = t(lambda: eng.hmc_sample(q, L, eps, _lib.RULE_ALG, 16, q_out=nxt), 4) / 16
        fl = C * (L + 1) * 8.5 * d / 1e9
        print(json.dumps({"lpc": sys.argv[1], "d": d, "single_ms": round(ms1, 4), "fused_ms": round(msK, 4),
                          "single_tflops": round(fl / ms1, 2), "fused_tflops": round(fl / msK, 2)}), flush=True)
    except Exception as ex:
        print(json.dumps({"lpc": sys.argv[1], "d": d, "error": str(ex)[:100]}), flush=True)
