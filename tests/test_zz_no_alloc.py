"""hmcb_drhmc_step_static -- the delayed-rejection step for callers that cannot hand
the library a workspace (no device allocator of their own, device memory already
reserved by the host application, or a loop where cudaMalloc's device sync is not
acceptable).  Its scratch is an array built into libhmcb.so, so the call allocates
nothing.

Checked here, through the C ABI (ctypes, raw pointers):
  * two calls in a row give exactly what hmcb_drhmc_step gives with a caller-owned
    workspace -- the second call starts from the first call's leftovers in the shared
    array and uses a different batch size, so stale contents must not matter;
  * a batch whose scratch would not fit in HMCB_STATIC_SCRATCH_BYTES is refused with a
    message that names both byte counts.

This file is named so that pytest collects it last.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from hmc_b200 import _lib

pytestmark = pytest.mark.gpu

D, NPROPS, NLEAP, EPS, FACTOR = 50, 3, 10, 0.35, 2.0


def _model():
    m = _lib.Model()
    m.kind, m.d, m.dtype = _lib.MODEL_FUNNEL, D, _lib.F64
    m.par[0] = 3.0
    return m


def _call(lib, static, q, iteration, dev):
    """One DR transition of every row of q via the chosen entry point; returns device tensors."""
    n = q.shape[0]
    out = {"q": torch.empty_like(q), "p": torch.empty_like(q),
           "code": torch.empty(n, dtype=torch.int32, device=dev),
           "avec": torch.empty((n, NPROPS), dtype=torch.float64, device=dev),
           "counts": torch.empty((n, 3), dtype=torch.int32, device=dev)}
    m = _model()
    ch = _lib.Chains(n, 0, q.data_ptr(), out["q"].data_ptr(), q.stride(0), q.stride(1))
    r = _lib.Rng()
    r.seed, r.iteration = 1234, iteration
    so = _lib.StepOut()
    so.p_out, so.code, so.avec, so.counts = (out["p"].data_ptr(), out["code"].data_ptr(), out["avec"].data_ptr(),
                                             out["counts"].data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    if static:
        rc = lib.hmcb_drhmc_step_static(C.byref(m), C.byref(ch), C.byref(r), NLEAP, EPS, NPROPS, FACTOR, 0, C.byref(so),
                                        stream)
    else:
        need = lib.hmcb_drhmc_workspace_bytes(C.byref(m), n, NPROPS)
        assert 0 < need
        ws = torch.empty(need, dtype=torch.uint8, device=dev)
        rc = lib.hmcb_drhmc_step(C.byref(m), C.byref(ch), C.byref(r), NLEAP, EPS, NPROPS, FACTOR, 0, C.byref(so),
                                 C.c_void_p(ws.data_ptr()), need, None, stream)
    assert rc == 0, lib.hmcb_last_error()
    torch.cuda.synchronize(dev)
    return out


def _same(a, b):
    assert torch.equal(a["code"], b["code"])
    assert torch.equal(a["counts"], b["counts"])
    for k in ("q", "p", "avec"):
        assert torch.allclose(a[k], b[k], rtol=1e-10, atol=1e-11), k


def test_static_scratch_matches_workspace_entry_point_twice():
    assert torch.cuda.is_available()
    dev = torch.device("cuda:0")
    lib = _lib.load()
    g = torch.Generator().manual_seed(7)
    q0 = (torch.rand(6000, D, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
    with torch.cuda.device(dev):
        # call 1: 6000 chains
        ref1 = _call(lib, False, q0, 0, dev)
        got1 = _call(lib, True, q0, 0, dev)
        _same(got1, ref1)
        assert len(torch.unique(ref1["code"])) >= 3          # several stages really ran (lists, frames used)
        # call 2, straight after: fewer chains, next iteration, starting from call 1's result;
        # the shared array still holds call 1's lists and frames
        q1 = ref1["q"][:4321].contiguous()
        ref2 = _call(lib, False, q1, 1, dev)
        got2 = _call(lib, True, q1, 1, dev)
        _same(got2, ref2)
        assert not torch.equal(got2["q"], q1)


def test_static_scratch_refuses_oversized_batches():
    dev = torch.device("cuda:0")
    lib = _lib.load()
    n = 200_000                                               # needs far more than 32 MiB of scratch
    with torch.cuda.device(dev):
        q = torch.zeros((n, D), dtype=torch.float64, device=dev)
        qo = torch.empty_like(q)
        m = _model()
        need = lib.hmcb_drhmc_workspace_bytes(C.byref(m), n, NPROPS)
        assert need > _lib.STATIC_SCRATCH_BYTES
        ch = _lib.Chains(n, 0, q.data_ptr(), qo.data_ptr(), q.stride(0), q.stride(1))
        r, so = _lib.Rng(), _lib.StepOut()
        rc = lib.hmcb_drhmc_step_static(C.byref(m), C.byref(ch), C.byref(r), NLEAP, EPS, NPROPS, FACTOR, 0, C.byref(so),
                                        None)
    assert rc == _lib.E_WORKSPACE
    msg = lib.hmcb_last_error().decode()
    # the message names both byte counts: what this batch needs as one range (hmcb_drhmc_workspace_bytes may
    # be a little larger: it sizes the multi-stream form of hmcb_drhmc_step) and what the library holds
    import re
    stated = [int(x) for x in re.findall(r"\d{6,}", msg)]
    assert _lib.STATIC_SCRATCH_BYTES in stated
    assert any(_lib.STATIC_SCRATCH_BYTES < x <= need for x in stated), msg
