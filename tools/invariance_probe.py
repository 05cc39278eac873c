#!/usr/bin/env python
"""Does one transition leave the funnel invariant?  Start 2^20 chains from EXACT funnel draws, apply T transitions, and
compare the pooled moments of v afterwards with (0, 3) in Monte Carlo errors.  Variants separate the sampler's rule from
the draws: plain HMC, delayed rejection, the exact-f64 normal generator, smaller steps."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmc_b200 import _lib, algorithms, models, pyhmc  # noqa: E402

dev = torch.device("cuda:0")
d, C = 50, 1 << 20


def exact(seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    v = 3.0 * torch.randn(C, 1, generator=g, dtype=torch.float64, device=dev)
    return torch.cat([v, torch.exp(v / 2) * torch.randn(C, d - 1, generator=g, dtype=torch.float64, device=dev)], dim=1).contiguous()


def report(name, q, extra=""):
    v = q[:, 0]
    m, s = v.mean().item(), v.std().item()
    se_m = s / C ** 0.5
    se_s = s / (2 * C) ** 0.5
    print("%-62s v mean %+.4f (%+.1f se)  std %.4f (%+.1f se) %s" % (name, m, m / se_m, s, (s - 3) / se_s, extra), flush=True)


q0 = exact(1)
report("exact draws", q0)
for name, kw, T in (("DRHMC nleap=150 eps=0.2 nprops=3, 1 transition", dict(), 1),
                    ("DRHMC nleap=150 eps=0.2 nprops=3, 8 transitions", dict(), 8),
                    ("DRHMC, exact f64 normals, 8 transitions", dict(rng_f64=True), 8)):
    e = algorithms.DRHMC(models.Funnel(d), device=dev, seed=5, **kw).engine
    q = q0
    for t in range(T):
        o = e.drhmc_step(q, 150, 0.2, 3, 2.0, want_p=False)
        q = o["q"]
    report(name, q, "codes %s" % (torch.bincount(o["code"].long(), minlength=4).double() / C).tolist())
for name, L, eps, T in (("HMC L=150 eps=0.2, 8 transitions", 150, 0.2, 8), ("HMC L=150 eps=0.05, 8 transitions", 150, 0.05, 8),
                        ("HMC L=20 eps=0.1, 8 transitions", 20, 0.1, 8), ("HMC L=600 eps=0.05, 8 transitions", 600, 0.05, 8)):
    e = pyhmc.HMC(models.Funnel(d), device=dev, seed=5).engine
    q = q0
    for t in range(T):
        o = e.hmc_step(q, L, eps, _lib.RULE_PYHMC, want_p=False)
        q = o["q"]
    report(name, q, "accept %.3f code2 %.4f" % ((o["code"] == 1).double().mean().item(), (o["code"] == 2).double().mean().item()))
# stage-wise: DR with nprops 1 and 2
for nprops in (1, 2):
    e = algorithms.DRHMC(models.Funnel(d), device=dev, seed=5).engine
    q = q0
    for t in range(8):
        q = e.drhmc_step(q, 150, 0.2, nprops, 2.0, want_p=False)["q"]
    report("DRHMC nprops=%d, 8 transitions" % nprops, q)
