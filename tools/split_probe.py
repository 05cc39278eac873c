#!/usr/bin/env python
"""Does running the DR step of disjoint chain ranges on several streams fill the tails of the
stage kernels?  Device-resident funnel d=50, 2^20 chains, nleap=150, nprops=3."""
import sys
import time

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from hmc_b200 import algorithms, models  # noqa: E402

dev = torch.device("cuda:0")
C, d = 1 << 20, 50
dr = algorithms.DRHMC(models.Funnel(d), device=dev, seed=3)
e = dr.engine
q = (torch.rand(C, d, dtype=torch.float64, device=dev) * 2 - 1)
for _ in range(100):
    q = e.drhmc_step(q, 150, 0.2, 3, 2.0, want_p=False)["q"]
qo = torch.empty_like(q)
code = torch.empty(C, dtype=torch.int32, device=dev)
counts = torch.empty((C, 3), dtype=torch.int32, device=dev)
avec = torch.empty((C, 3), dtype=torch.float64, device=dev)
streams = [torch.cuda.Stream() for _ in range(8)]


def run(nsplit, order="seq"):
    cur = torch.cuda.current_stream()
    if nsplit == 1:
        e.drhmc_step(q, 150, 0.2, 3, 2.0, want_p=False, q_out=qo, code_out=code, counts_out=counts, avec_out=avec,
                     advance=False)
        return
    bounds = [(i * C) // nsplit for i in range(nsplit + 1)]
    for i in range(nsplit):
        st = streams[i]
        st.wait_stream(cur)
        lo, hi = bounds[i], bounds[i + 1]
        with torch.cuda.stream(st):
            e.drhmc_step(q[lo:hi], 150, 0.2, 3, 2.0, want_p=False, q_out=qo[lo:hi], code_out=code[lo:hi],
                         counts_out=counts[lo:hi], avec_out=avec[lo:hi], extra_offset=lo, ws_slot=i, advance=False)
    for i in range(nsplit):
        cur.wait_stream(streams[i])


ref = None
for ns in (1, 2, 3, 4, 6, 8, 1):
    run(ns)
    torch.cuda.synchronize()
    if ref is None:
        ref = (qo.clone(), code.clone(), counts.clone())
    else:
        assert torch.equal(ref[0], qo) and torch.equal(ref[1], code) and torch.equal(ref[2], counts)
    t0 = time.perf_counter()
    for _ in range(10):
        run(ns)
    torch.cuda.synchronize()
    print("nsplit=%d: %.3f ms per step (bit-identical)" % (ns, (time.perf_counter() - t0) / 10 * 1e3))
