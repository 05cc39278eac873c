#!/usr/bin/env python
"""One launch of every kernel the bench line and its secondary block rest on, at the bench's shapes, for
`ncu --set full` (see profiles/README.md for the command).  Each kernel is launched twice: the first launch warms
caches and attributes, ncu (run with --profile-from-start off) captures the second, bracketed by cudaProfilerStart/Stop."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import _lib, adaptivehmc, algorithms, mala, models, pyhmc  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
which = set(sys.argv[1:]) or {"dr", "hmc", "gauss", "mala", "ghmc", "adhmc", "walk", "logreg", "moments"}


def twice(fn):
    fn()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()     # ncu --profile-from-start off: only the second launch of each target is captured
    fn()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()


d, C = 50, 1 << 20
fm = models.Funnel(d)
if "dr" in which or "hmc" in which or "ghmc" in which or "moments" in which:
    q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
if "dr" in which:
    dr = algorithms.DRHMC(fm, device=dev, seed=2).engine
    qb = dr.drhmc_sample(q, 150, 0.2, 3, 2.0, 40, want_avec=False)["q"]      # some burn-in so the stage lists look like the bench's
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    for e in evs:
        e.record()
    twice(lambda: dr.drhmc_step(qb, 150, 0.2, 3, 2.0, want_p=False, stage_events=evs, want_avec=False))   # staged, serialised
if "hmc" in which:
    h = pyhmc.HMC(fm, device=dev, seed=1).engine
    twice(lambda: h.hmc_step(q, 20, 0.1, _lib.RULE_PYHMC, want_p=False))
    twice(lambda: h.hmc_sample(q, 20, 0.1, _lib.RULE_PYHMC, 8))
    twice(lambda: h.hmc_sample(q, 150, 0.2, _lib.RULE_PYHMC, 4))
if "moments" in which:
    s1, s2 = torch.zeros_like(q), torch.zeros_like(q)
    e = pyhmc.HMC(fm, device=dev).engine
    twice(lambda: e.moments_accumulate(q, s1, s2))
if "ghmc" in which:
    gh = algorithms.GHMC(fm, device=dev, seed=1).engine
    pp = torch.randn(C, d, generator=g, dtype=torch.float64).to(dev)
    o0 = gh.ghmc_step(q, pp, 20, 0.1, 0.5)
    twice(lambda: gh.ghmc_step(o0["q"], o0["p"], 20, 0.1, 0.5, V=o0["V"], Vg=o0["Vg"]))
    twice(lambda: gh.ghmc_sample(o0["q"], o0["p"], 20, 0.1, 0.5, 8, V=o0["V"], Vg=o0["Vg"]))     # k_ghmc_multi, 8 transitions
    twice(lambda: gh.ghmc_step(o0["q"], o0["p"], 1, 0.1, 0.5, V=o0["V"], Vg=o0["Vg"]))
    twice(lambda: gh.ghmc_sample(o0["q"], o0["p"], 1, 0.1, 0.5, 16, V=o0["V"], Vg=o0["Vg"]))
if "gauss" in which:
    gm = models.CorrGaussian.ar1(100)
    s = pyhmc.HMC(gm, device=dev, seed=1).engine
    qg = torch.randn(65536, 100, generator=g, dtype=torch.float64).to(dev)
    eps = 0.1 * float(np.sqrt(np.linalg.eigvalsh(gm.Sigma).min()))
    twice(lambda: s.hmc_step(qg, 32, eps, _lib.RULE_PYHMC, want_p=False))
if which & {"mala", "adhmc", "walk"}:
    rm = models.Rosenbrock(100)
    qr = (1.0 + 0.1 * torch.randn(131072, 100, generator=g, dtype=torch.float64)).to(dev)
if "mala" in which:
    m = mala.MALA(rm, device=dev, seed=4).engine
    twice(lambda: m.mala_step(qr, 0.02, n_steps=32))
if "adhmc" in which:
    ad = adaptivehmc.AdHMC_eps0(rm, device=dev, seed=4)
    steps = ad.get_steps(0.01, 1.0, 10)
    twice(lambda: ad.engine.adhmc_eps0_step(qr, 20, steps, 1))
if "walk" in which:
    aw = adaptivehmc.AdHMC(rm, device=dev, seed=4)
    steps = aw.get_steps(0.005, 0.2, 10)
    twice(lambda: aw.engine.adhmc_walk_step(qr, 4, 5, 0.0, [2 ** -0.5, 2 ** 0.5], [1.0, 1.0], 0.005, 0.2, steps, 1))
if "logreg" in which:
    N, D, Cl = 100000, 128, 16384
    rng = np.random.default_rng(3)
    X = (rng.standard_normal((N, D)) / np.sqrt(D)).astype(np.float32)
    y = (rng.random(N) < 0.5).astype(np.float32)
    lm = algorithms.HMC(models.LogisticRegression(X, y, 1.0), device=dev, seed=3).engine
    ql = (0.1 * torch.randn(Cl, D, generator=g)).to(dev)
    twice(lambda: lm.logp_grad(ql))
print("done")
