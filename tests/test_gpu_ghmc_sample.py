"""hmcb_ghmc_sample (K GHMC / Persistent_GHMC transitions per call, the loops of src/algorithms.py:284-303 and :336-357
without the stored history) against K calls of hmcb_ghmc_step chained through their outputs: bit-identical state, momentum,
caches (None = NaN for chains that have not accepted yet), persistent uniform, last code and energies, summed work
counters and accept counts.  The single step is pinned to the oracle in tests/test_gpu_ghmc.py."""
import numpy as np
import pytest
import torch

from hmc_b200 import algorithms, models

pytestmark = pytest.mark.gpu


def _dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _model(kind, d):
    if kind == "funnel":
        return models.Funnel(d)
    if kind == "rosenbrock":
        return models.Rosenbrock(d)
    return models.CorrGaussian.ar1(d)


def _start(kind, C, d, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    q = torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1
    if kind == "rosenbrock":
        q = 1.0 + 0.1 * q
    p = torch.randn(C, d, generator=g, dtype=torch.float64)
    u = torch.randn(C, generator=g, dtype=torch.float64)
    return q.to(_dev(), dtype), p.to(_dev(), dtype), u.to(_dev())


def _chained_single(e, q, p, nleap, eps, alpha, K, pu=None, delta=0.0, V=None, Vg=None):
    counts = torch.zeros((q.shape[0], 3), dtype=torch.int32, device=q.device)
    nacc = torch.zeros(q.shape[0], dtype=torch.int32, device=q.device)
    o = None
    for _ in range(K):
        o = e.ghmc_step(q, p, nleap, eps, alpha, V=V, Vg=Vg, pu=pu, delta=delta)
        q, p, V, Vg, pu = o["q"], o["p"], o["V"], o["Vg"], o["u"]
        counts += o["counts"]
        nacc += (o["code"] == 1).int()
    o = dict(o)
    o["counts"], o["n_accept"] = counts, nacc
    return o


def _same(a, b, keys):
    for k in keys:
        x, y = a[k], b[k]
        assert x.shape == y.shape, k
        assert torch.equal(torch.nan_to_num(x.double(), nan=-7.25), torch.nan_to_num(y.double(), nan=-7.25)), k
        assert torch.equal(torch.isnan(x.double()), torch.isnan(y.double())), k


KEYS = ("q", "p", "V", "Vg", "code", "Hs", "counts", "n_accept")


@pytest.mark.parametrize("kind,d,eps,nleap", [("funnel", 50, 0.25, 4), ("funnel", 50, 0.1, 1), ("funnel", 10, 0.3, 6),
                                              ("funnel", 100, 0.2, 3), ("rosenbrock", 20, 0.05, 5),
                                              ("rosenbrock", 100, 0.04, 3), ("gaussian", 16, 0.4, 4)])
@pytest.mark.parametrize("persistent", [False, True])
def test_fused_equals_chained_single_calls(kind, d, eps, nleap, persistent, monkeypatch):
    C, K, alpha = 2500, 7, 0.6
    if d >= 32:
        monkeypatch.setenv("HMCB_LPC", "4")       # the step's lane shape for the fused launch too: same in-chain sums
    q, p, u = _start(kind, C, d, torch.float64, 40 + d)
    a = algorithms.GHMC(_model(kind, d), device=_dev(), seed=5)
    b = algorithms.GHMC(_model(kind, d), device=_dev(), seed=5)
    pu = u if persistent else None
    ref = _chained_single(a.engine, q, p, nleap, eps, alpha, K, pu=pu, delta=0.37)
    out = b.engine.ghmc_sample(q, p, nleap, eps, alpha, K, pu=pu, delta=0.37)
    _same(out, ref, KEYS + (("u",) if persistent else ()))
    assert a.engine.iteration == b.engine.iteration == K
    na = out["n_accept"].cpu().numpy()
    assert 0 < na.sum() < C * K                                   # both outcomes occur
    assert torch.equal(torch.isnan(out["V"]).cpu(), torch.as_tensor(na == 0))   # V stays None until a chain first accepts


def test_fused_continues_from_caches_and_accumulates_n_accept(monkeypatch):
    monkeypatch.setenv("HMCB_LPC", "4")
    d, C, nleap, eps, alpha = 50, 3000, 3, 0.2, 0.3
    q, p, _ = _start("funnel", C, d, torch.float64, 3)
    a = algorithms.GHMC(models.Funnel(d), device=_dev(), seed=9)
    b = algorithms.GHMC(models.Funnel(d), device=_dev(), seed=9)
    ref = _chained_single(a.engine, q, p, nleap, eps, alpha, 9)
    o1 = b.engine.ghmc_sample(q, p, nleap, eps, alpha, 4)
    o2 = b.engine.ghmc_sample(o1["q"], o1["p"], nleap, eps, alpha, 5, V=o1["V"], Vg=o1["Vg"], n_accept=o1["n_accept"])
    _same(o2, ref, ("q", "p", "V", "Vg", "code", "Hs", "n_accept"))
    assert torch.equal(o1["counts"] + o2["counts"], ref["counts"])


def test_one_transition_is_the_step():
    d, C = 50, 1500
    q, p, u = _start("funnel", C, d, torch.float64, 4)
    a = algorithms.GHMC(models.Funnel(d), device=_dev(), seed=2)
    b = algorithms.GHMC(models.Funnel(d), device=_dev(), seed=2)
    ref = _chained_single(a.engine, q, p, 5, 0.2, 0.5, 1, pu=u, delta=0.1)
    out = b.engine.ghmc_sample(q, p, 5, 0.2, 0.5, 1, pu=u, delta=0.1)
    _same(out, ref, KEYS + ("u",))


def test_float32_and_diagonal_metric(monkeypatch):
    monkeypatch.setenv("HMCB_LPC", "4")
    d, C, K = 50, 2000, 5
    q, p, _ = _start("funnel", C, d, torch.float32, 6)
    a = algorithms.GHMC(models.Funnel(d), device=_dev(), dtype=torch.float32, seed=3)
    b = algorithms.GHMC(models.Funnel(d), device=_dev(), dtype=torch.float32, seed=3)
    _same(b.engine.ghmc_sample(q, p, 4, 0.2, 0.5, K), _chained_single(a.engine, q, p, 4, 0.2, 0.5, K), KEYS)
    minv = np.linspace(0.5, 2.0, d)
    q, p, _ = _start("funnel", C, d, torch.float64, 7)
    a = algorithms.GHMC(models.Funnel(d), invmetric_diag=minv, device=_dev(), seed=3)
    b = algorithms.GHMC(models.Funnel(d), invmetric_diag=minv, device=_dev(), seed=3)
    _same(b.engine.ghmc_sample(q, p, 4, 0.2, 0.5, K), _chained_single(a.engine, q, p, 4, 0.2, 0.5, K), KEYS)


@pytest.mark.parametrize("cls", [algorithms.GHMC, algorithms.Persistent_GHMC])
def test_sample_without_history_equals_the_stored_run(cls, monkeypatch):
    monkeypatch.setenv("HMCB_LPC", "4")
    d, C, ns, nb = 50, 1200, 5, 3
    q, p, u = _start("funnel", C, d, torch.float64, 8)
    kw = dict(nsamples=ns, burnin=nb, step_size=0.2, nleap=4, alpha=0.7)
    if cls is algorithms.Persistent_GHMC:
        kw.update(delta=0.21, u=u)
    a = cls(models.Funnel(d), device=_dev(), seed=11)
    b = cls(models.Funnel(d), device=_dev(), seed=11)
    st = a.sample(q.cpu().numpy(), p=p.cpu().numpy(), **kw)
    f = b.sample(q.cpu().numpy(), p=p.cpu().numpy(), store=False, **kw)
    assert f.samples is None and f.i == st.i == ns + nb
    assert np.array_equal(f.q, st.samples[-1])
    assert np.array_equal(f.n_accept, (st.accepts[nb:] == 1).sum(axis=0))
    assert np.array_equal(f.counts, st.counts[nb:].sum(axis=0))
    if cls is algorithms.Persistent_GHMC:
        assert np.array_equal(f.u, st.u[-1])


def test_default_shapes_agree_to_rounding():
    """d >= 32 without HMCB_LPC: the fused launch runs 2 lanes per chain, the step 4 -- same draws, same decisions, state
    equal to the rounding of the in-chain sums."""
    d, C, K = 50, 4000, 6
    q, p, _ = _start("funnel", C, d, torch.float64, 12)
    a = algorithms.GHMC(models.Funnel(d), device=_dev(), seed=21)
    b = algorithms.GHMC(models.Funnel(d), device=_dev(), seed=21)
    ref = _chained_single(a.engine, q, p, 4, 0.2, 0.5, K)
    out = b.engine.ghmc_sample(q, p, 4, 0.2, 0.5, K)
    assert torch.equal(out["n_accept"], ref["n_accept"]) and torch.equal(out["counts"], ref["counts"])
    assert torch.equal(out["code"], ref["code"])
    for k in ("q", "p", "Vg", "V", "Hs"):
        assert torch.allclose(out[k], ref[k], rtol=1e-10, atol=1e-11, equal_nan=True), k


def test_argument_errors():
    d, C = 10, 64
    q, p, _ = _start("funnel", C, d, torch.float64, 9)
    e = algorithms.GHMC(models.Funnel(d), device=_dev()).engine
    with pytest.raises(Exception):
        e.ghmc_sample(q, p, 3, 0.1, 0.5, 0)
    with pytest.raises(Exception):
        e.ghmc_sample(q, p, 3, 0.1, 1.5, 2)


@pytest.mark.parametrize("d,K", [(50, 5), (10, 4), (100, 3)])
def test_outputs_stay_inside_their_buffers(d, K):
    """hmcb_ghmc_sample through the C ABI with every output placed inside a larger, sentinel-filled allocation and a
    ragged chain count (the last warp is partial, and at d=50 / 100 the lanes carry padding slots): nothing outside the
    [C, ...] windows may change, and the windows must hold what the engine call returns."""
    import ctypes as C_
    from hmc_b200 import _lib
    Cn, G = 1237, 97
    q, p, u = _start("funnel", Cn, d, torch.float64, 60 + d)
    a = algorithms.GHMC(models.Funnel(d), device=_dev(), seed=31)
    b = algorithms.GHMC(models.Funnel(d), device=_dev(), seed=31)
    ref = a.engine.ghmc_sample(q, p, 3, 0.2, 0.5, K, pu=u, delta=0.2)
    e = b.engine
    m = e._ensure()
    S = -12345.0

    def padded(shape, dtype):
        n = int(np.prod(shape))
        buf = torch.full((n + 2 * G,), S if dtype.is_floating_point else -12345, dtype=dtype, device=_dev())
        return buf, buf[G:G + n].view(shape)
    bufs = {k: padded(s, t) for k, (s, t) in dict(q=((Cn, d), torch.float64), p=((Cn, d), torch.float64), V=((Cn,), torch.float64),
                                                  Vg=((Cn, d), torch.float64), u=((Cn,), torch.float64), code=((Cn,), torch.int32),
                                                  Hs=((Cn, 2), torch.float64), counts=((Cn, 3), torch.int32),
                                                  n_accept=((Cn,), torch.int32)).items()}
    bufs["n_accept"][1].zero_()
    io = _lib.GhmcIO()
    io.p_out, io.V_in, io.Vg_in = _lib.ptr(bufs["p"][1]), None, None
    io.V_out, io.Vg_out = _lib.ptr(bufs["V"][1]), _lib.ptr(bufs["Vg"][1])
    io.u_in, io.u_out, io.delta = _lib.ptr(u), _lib.ptr(bufs["u"][1]), 0.2
    io.code, io.energies, io.counts = _lib.ptr(bufs["code"][1]), _lib.ptr(bufs["Hs"][1]), _lib.ptr(bufs["counts"][1])
    ch = e._chains(q, bufs["q"][1])
    r, keep = e._rng(None, None, 1, Cn, p, 0.5)
    with torch.cuda.device(_dev()):
        _lib.check(e.lib.hmcb_ghmc_sample(C_.byref(m), C_.byref(ch), C_.byref(r), 3, 0.2, C_.byref(io), K,
                                          C_.c_void_p(bufs["n_accept"][1].data_ptr()), e._stream()))
    torch.cuda.synchronize()
    for k, (buf, win) in bufs.items():
        sent = S if buf.dtype.is_floating_point else -12345
        assert bool((buf[:G] == sent).all()) and bool((buf[-G:] == sent).all()), "wrote outside " + k
    got = {k: w for k, (_, w) in bufs.items()}
    _same(got, ref, KEYS + ("u",))
