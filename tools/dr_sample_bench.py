#!/usr/bin/env python
"""One hmcb_drhmc_step call per transition against hmcb_drhmc_sample (K transitions per call, chain ranges pipelined over
streams) on the headline workload:
funnel d=50, nleap=150, eps=0.2, nprops=3, f=2.  Prints ms per transition and the fraction of the FP64 peak
(gradient evaluations x 350 flop / time / measured DFMA peak) for each form and chain count."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import _lib, algorithms, models  # noqa: E402


def fp64_peak(eng, dev):
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    sink = torch.empty(sms * 8 * 256, dtype=torch.float64, device=dev)
    cnt = _lib.C.c_double(0)
    best = 0.0
    for it in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(eng.lib.hmcb_probe_fma(_lib.F64, sms * 8, 256, 4000, _lib.ptr(sink), _lib.C.byref(cnt), eng._stream()))
        e1.record()
        torch.cuda.synchronize()
        if it:
            best = max(best, 2.0 * cnt.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chains", type=int, nargs="+", default=[1 << 20, 1 << 17])
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--nleap", type=int, default=150)
    ap.add_argument("--eps", type=float, default=0.2)
    ap.add_argument("--burnin", type=int, default=200)
    ap.add_argument("--d", type=int, default=50)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    d, K = a.d, a.steps
    rows = []
    for C in a.chains:
        eng = algorithms.DRHMC(models.Funnel(d), device=dev, seed=2).engine
        eng._ensure()
        peak = fp64_peak(eng, dev)
        g = torch.Generator().manual_seed(2)
        q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
        q = eng.drhmc_sample(q, a.nleap, a.eps, 3, 2.0, a.burnin, want_avec=False)["q"]   # burn-in in one launch
        torch.cuda.synchronize()
        nxt = torch.empty_like(q)
        s1, s2 = torch.zeros_like(q), torch.zeros_like(q)
        code = torch.empty(C, dtype=torch.int32, device=dev)
        counts = torch.empty((K, C, 3), dtype=torch.int32, device=dev)

        def timed(fn, reps=1):
            fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps

        state = {"q": q.clone()}

        def staged():
            qi = state["q"]
            for k in range(K):
                eng.drhmc_step(qi, a.nleap, a.eps, 3, 2.0, want_p=False, q_out=nxt, want_avec=False, code_out=code,
                               counts_out=counts[k])
                eng.moments_accumulate(nxt, s1, s2)
                qi, _ = nxt, qi

        def tasks_k1():
            qi = state["q"]
            for k in range(K):
                eng.drhmc_sample(qi, a.nleap, a.eps, 3, 2.0, 1, msum=s1, msumsq=s2, q_out=nxt, want_avec=False,
                                 code_out=code, counts_out=counts[k])
                qi = nxt

        def tasks_fused():
            eng.drhmc_sample(state["q"], a.nleap, a.eps, 3, 2.0, K, msum=s1, msumsq=s2, q_out=nxt, want_avec=False,
                             code_out=code, counts_out=counts[0])

        for name, fn in (("step call per transition (+ moments kernel)", staged),
                         ("sample call, 1 transition per call (moments folded)", tasks_k1),
                         ("sample call, %d transitions per call (moments folded)" % K, tasks_fused)):
            ms = timed(fn)
            if fn is tasks_fused:
                grads = float(counts[0, :, 1].sum(dtype=torch.float64))
            else:
                grads = float(counts[:, :, 1].sum(dtype=torch.float64))
            tf = grads * 7 * d / (ms * 1e-3) / 1e12
            rows.append({"chains": C, "form": name, "ms_per_transition": ms / K, "grad_evals_per_transition_per_chain":
                         grads / K / C, "grad_evals_per_s": grads / (ms * 1e-3), "tflops": tf, "frac_of_fp64_peak": tf / peak,
                         "fp64_peak_tflops": peak})
            print(json.dumps(rows[-1]), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/dr_sample_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
