#!/usr/bin/env python
"""Secondary measurements (not the bench.py contract line): the other BASELINE.json
configs that fit one GPU, each as grad evals/s and as a fraction of the measured FP64
DFMA peak.  Writes a markdown table to stdout.   python tools/bench_configs.py [--quick]"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hmc_b200 import _lib, adaptivehmc, algorithms, mala, models, pyhmc  # noqa: E402


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best, r


def fp64_peak(eng):
    eng._ensure()
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    sink = torch.empty(sms * 8 * 256, dtype=torch.float64, device="cuda")
    cnt = _lib.C.c_double(0)

    def go():
        _lib.check(eng.lib.hmcb_probe_fma(_lib.F64, sms * 8, 256, 4000, _lib.ptr(sink), _lib.C.byref(cnt), eng._stream()))
    ms, _ = timed(go)
    return 2 * cnt.value / (ms * 1e-3) / 1e12


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    rows = []
    g = torch.Generator().manual_seed(0)

    # config 2: correlated Gaussian d=100, 65,536 chains, fixed step / L
    d, C, L = 100, 65536, 32
    gm = models.CorrGaussian.ar1(d)
    s = pyhmc.HMC(gm, device=dev, seed=1)
    peak = fp64_peak(s.engine)
    q = torch.randn(C, d, generator=g, dtype=torch.float64).to(dev)
    eps = 0.1 * float(np.sqrt(np.linalg.eigvalsh(gm.Sigma).min()))
    ms, out = timed(lambda: s.engine.hmc_step(q, L, eps, _lib.RULE_PYHMC, want_p=False))
    grads = C * (L + 1)
    flops = grads * (2 * d * d + 4 * d)
    rows.append(("cfg2 gaussian d=100 HMC L=32, 65,536 chains (eps=%.4f)" % eps, ms, grads / ms * 1e3, flops / ms * 1e-9, flops / ms * 1e-9 / peak,
                 float((out["code"] == 1).double().mean())))

    # config 1 shape on the GPU: funnel d=10 HMC L=20 (many chains)
    d, C, L = 10, 1 << 20, 20
    s = pyhmc.HMC(models.Funnel(d), device=dev, seed=1)
    q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
    ms, out = timed(lambda: s.engine.hmc_step(q, L, 0.1, _lib.RULE_PYHMC, want_p=False))
    grads = C * (L + 1)
    rows.append(("funnel d=10 HMC L=20, 2^20 chains", ms, grads / ms * 1e3, grads * 7 * d / ms * 1e-9, grads * 7 * d / ms * 1e-9 / peak,
                 float((out["code"] == 1).double().mean())))

    # funnel d=50 vanilla HMC L=20 (light variant) and L=150
    for L in (20, 150):
        d, C = 50, 1 << 20
        s = pyhmc.HMC(models.Funnel(d), device=dev, seed=1)
        q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
        ms, out = timed(lambda: s.engine.hmc_step(q, L, 0.1, _lib.RULE_PYHMC, want_p=False))
        grads = C * (L + 1)
        rows.append(("funnel d=50 HMC L=%d, 2^20 chains" % L, ms, grads / ms * 1e3, grads * 7 * d / ms * 1e-9, grads * 7 * d / ms * 1e-9 / peak,
                     float((out["code"] == 1).double().mean())))

    # config 5: Rosenbrock sweep, 131,072 chains: HMC L=20, MALA (32 fused steps), AdHMC_eps0
    for d in ((2, 10, 100) if a.quick else (2, 4, 10, 20, 50, 100)):
        C = 131072
        rm = models.Rosenbrock(d)
        q = (1.0 + 0.1 * torch.randn(C, d, generator=g, dtype=torch.float64)).to(dev)
        fpg = 8.5 * d
        h = algorithms.HMC(rm, device=dev, seed=4)
        ms, out = timed(lambda: h.engine.hmc_step(q, 20, 0.02, _lib.RULE_ALG, want_p=False))
        grads = C * 21
        rows.append(("cfg5 rosenbrock d=%d HMC L=20" % d, ms, grads / ms * 1e3, grads * fpg / ms * 1e-9, grads * fpg / ms * 1e-9 / peak,
                     float((out["code"] == 1).double().mean())))
        m = mala.MALA(rm, device=dev, seed=4)
        K = 32
        ms, out = timed(lambda: m.engine.mala_step(q, 0.02, n_steps=K))
        grads = C * (K + 1)
        rows.append(("cfg5 rosenbrock d=%d MALA %d fused steps" % (d, K), ms, grads / ms * 1e3, grads * fpg / ms * 1e-9,
                     grads * fpg / ms * 1e-9 / peak, float(out["n_accept"].double().mean()) / K))
        ad = adaptivehmc.AdHMC_eps0(rm, device=dev, seed=4)
        steps = ad.get_steps(0.01, 1.0, 10)
        ms, out = timed(lambda: ad.engine.adhmc_eps0_step(q, 20, steps, 1))
        grads = float(out["counts"][:, 1].double().sum())
        rows.append(("cfg5 rosenbrock d=%d AdHMC_eps0 ntry=10 Nleap=20" % d, ms, grads / ms * 1e3, grads * fpg / ms * 1e-9,
                     grads * fpg / ms * 1e-9 / peak, float((out["code"] == 1).double().mean())))

    print("Measured FP64 DFMA peak: %.2f TFLOP/s\n" % peak)
    print("| config | ms / transition | grad evals/s | algorithmic TFLOP/s | of FP64 peak | accept |")
    print("|---|---:|---:|---:|---:|---:|")
    for r in rows:
        print("| %s | %.3f | %.3e | %.2f | %.3f | %.2f |" % r)


if __name__ == "__main__":
    main()
