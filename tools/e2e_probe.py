#!/usr/bin/env python
"""Where the end-to-end DRHMC step (host buffers in and out) spends its time: raw pinned copy
rates in each direction, both at once, the device-resident step, and the chunked pipeline at
several chunk counts."""
import sys
import time

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from hmc_b200 import algorithms, models  # noqa: E402

dev = torch.device("cuda:0")
C, d = 1 << 20, 50
q = (torch.rand(C, d, dtype=torch.float64) * 2 - 1).pin_memory()
out = torch.empty_like(q).pin_memory()
qd = torch.empty((C, d), dtype=torch.float64, device=dev)
qd2 = torch.empty_like(qd)


def timed(fn, n=5):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


gb = q.numel() * 8 / 1e9
t = timed(lambda: qd.copy_(q, non_blocking=True))
print("H2D %.1f MB: %.2f ms = %.1f GB/s" % (gb * 1e3, t, gb / t * 1e3))
t = timed(lambda: out.copy_(qd2, non_blocking=True))
print("D2H: %.2f ms = %.1f GB/s" % (t, gb / t * 1e3))
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def both():
    with torch.cuda.stream(s1):
        qd.copy_(q, non_blocking=True)
    with torch.cuda.stream(s2):
        out.copy_(qd2, non_blocking=True)


t = timed(both)
print("H2D + D2H at once: %.2f ms" % t)

dr = algorithms.DRHMC(models.Funnel(d), device=dev, seed=3)
qdev = q.to(dev)
for _ in range(30):
    qdev = dr.step(qdev, 150, 0.2, 3, 2.0, want_p=False)[0]
t = timed(lambda: dr.step(qdev, 150, 0.2, 3, 2.0, want_p=False))
print("device-resident step: %.2f ms" % t)
qh = qdev.cpu().pin_memory()
for nch, rs in ((8, 1), (8, 2), (12, 2), (16, 2), (24, 2), (32, 2)):
    e = dr.engine
    t = timed(lambda: e.drhmc_step_host(qh, 150, 0.2, 3, 2.0, want_p=False, n_chunks=nch, run_streams=rs), n=4)
    print("pipeline n_chunks=%d run_streams=%d: %.2f ms" % (nch, rs, t))
t = timed(lambda: dr.step(qh, 150, 0.2, 3, 2.0, want_p=False), n=4)
print("DRHMC.step(host): %.2f ms" % t)
t0 = time.perf_counter()
for _ in range(4):
    x = torch.empty((C, d), dtype=torch.float64, pin_memory=True)
    del x
print("pinned alloc+free of 419 MB: %.2f ms" % ((time.perf_counter() - t0) / 4 * 1e3))
