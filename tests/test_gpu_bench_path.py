"""The path bench.py times, checked against the oracle.

Every other parity test injects the momentum (``p=``), and injection switches the
warp-specialised kernel off; what the benchmark runs is ``k_transition_ws`` (momenta drawn in
the kernel from Philox) followed by the delayed-rejection stages at nleap=150, eps=0.2,
nprops=3, stepfactor=2 on the d=50 funnel.  Here that path runs WITHOUT injection: the
momentum each chain drew is read back (``hmcb_step_out.p0_out``), the uniforms it consumed
are recomputed on the host from the documented Philox counters (oracle/philox.py, pinned
to the library's generator and the Random123 vectors in tests/test_rng_cpu.py), and both
are replayed through ``oracle.samplers.drhmc_step`` (bit-pinned to
/root/reference/src/algorithms.py:474-530).

Also here: the error-vs-trajectory-length curve and the check
that the exact-f64 normal generator gives the same draws in every kernel that serves a shape.

Measured values are written to gpurun_out/ (copied under profiles/ when committed)."""
import json
import os

import numpy as np
import pytest
import torch

from hmc_b200 import _lib, algorithms, models, pyhmc
from oracle import models as omodels
from oracle import philox as ophilox
from oracle import replay, samplers

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _exact_funnel(C, d, seed):
    rng = np.random.default_rng(seed)
    v = 3.0 * rng.standard_normal((C, 1))
    return np.concatenate([v, np.exp(v / 2) * rng.standard_normal((C, d - 1))], axis=1)


def _relerr(a, b):
    """|a-b| / (|b| + 0.1): the denominator of the 1e-10 / 1e-11 (rtol / atol) parity bound."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / (np.abs(b) + 0.1)


def _dump(name, obj):
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, name), "w") as f:
        json.dump(obj, f, indent=1)


# ------------------------------------------------------------------ which draws a chain gets
@pytest.mark.parametrize("C", [100, 4096 + 37])        # plain kernel / warp-specialised kernel (>= 4096 chains)
@pytest.mark.parametrize("lpc", [0, 4])
def test_drawn_momenta_are_the_documented_philox_normals(C, lpc, monkeypatch):
    if lpc:
        monkeypatch.setenv("HMCB_LPC", str(lpc))
    else:
        monkeypatch.delenv("HMCB_LPC", raising=False)
    d, seed, off = 50, 987654321, 12345
    q0 = torch.as_tensor(_exact_funnel(C, d, 1), device=_dev())
    chains = np.arange(off, off + C, dtype=np.uint64)
    for f64 in (False, True):
        s = pyhmc.HMC(models.Funnel(d), device=_dev(), seed=seed, chain_offset=off, rng_f64=f64)
        s.engine.iteration = 5
        out = s.engine.hmc_step(q0, 3, 0.01, _lib.RULE_PYHMC, want_p0=True)
        p0 = out["p0"].cpu().numpy()
        if f64:
            want = ophilox.momentum_normals_f64(seed, chains, 5, d)
            assert np.abs(p0 - want).max() < 1e-12
        else:
            want = ophilox.momentum_normals_f32(seed, chains, 5, d)
            # MUFU lg2 / sin / cos against correctly rounded float32: ~1e-6 absolute
            assert np.abs(p0 - want).max() < 2e-5
            assert np.array_equal(p0, p0.astype(np.float32).astype(np.float64))   # the draw is a float32 value


@pytest.mark.parametrize("f64", [False, True])
def test_warp_specialised_kernel_draws_what_the_plain_kernel_draws(f64, monkeypatch):
    """ws == plain bit for bit, for HMC and for DR with nprops >= 2, including the momentum handed back for
    rejected chains (re-drawn in later kernels) -- with either normal generator."""
    d, C = 50, 4096 + 64
    q0 = torch.as_tensor(_exact_funnel(C, d, 2), device=_dev())
    res = []
    for no_ws in ("", "1"):
        if no_ws:
            monkeypatch.setenv("HMCB_NO_WS", "1")
        else:
            monkeypatch.delenv("HMCB_NO_WS", raising=False)
        h = pyhmc.HMC(models.Funnel(d), device=_dev(), seed=4, rng_f64=f64)
        o1 = h.engine.hmc_step(q0, 7, 0.2, _lib.RULE_PYHMC, want_p0=True)
        dr = algorithms.DRHMC(models.Funnel(d), device=_dev(), seed=4, rng_f64=f64)
        o2 = dr.engine.drhmc_step(q0, 7, 0.5, 3, 2.0, want_p0=True)
        res.append((o1, o2))
    for k in ("q", "p", "p0", "code", "Hs", "counts"):
        assert torch.equal(res[0][0][k], res[1][0][k]), "HMC %s differs between the two kernels" % k
    for k in ("q", "p", "p0", "code", "avec", "counts"):
        assert torch.equal(res[0][1][k], res[1][1][k]), "DRHMC %s differs between the two kernels" % k
    code = res[0][1]["code"]
    assert (code == 0).any() and (code >= 2).any()            # rejected chains and later-stage acceptances occur
    rej = code == 0
    assert torch.equal(res[0][1]["p"][rej], res[0][1]["p0"][rej])   # a rejected chain gets the momentum it started from


# ------------------------------------------------------------------ the benchmarked transition against the oracle
def test_headline_transition_without_injection_vs_oracle():
    d, C = 50, 4096
    nleap, eps, nprops, f = 150, 0.2, 3, 2.0
    seed, off, it = 2, 1 << 33, 11          # a global chain id above 2^32 exercises the counter's high word
    q0 = _exact_funnel(C, d, 7)
    dr = algorithms.DRHMC(models.Funnel(d), device=_dev(), seed=seed, chain_offset=off)
    dr.engine.iteration = it
    out = dr.engine.drhmc_step(torch.as_tensor(q0, device=_dev()), nleap, eps, nprops, f, want_p0=True)
    q, p, p0 = out["q"].cpu().numpy(), out["p"].cpu().numpy(), out["p0"].cpu().numpy()
    code, counts, avec = out["code"].cpu().numpy(), out["counts"].cpu().numpy(), out["avec"].cpu().numpy()
    chains = np.arange(off, off + C, dtype=np.uint64)
    U = ophilox.accept_uniforms(seed, chains, it, nprops)
    assert np.abs(p0 - ophilox.momentum_normals_f32(seed, chains, it, d)).max() < 2e-5

    om = omodels.Funnel(d)
    ref = [samplers.drhmc_step(om, q0[c], nleap, eps, nprops, f, replay.Draws([p0[c]], U[c])) for c in range(C)]
    rcode = np.array([r[2] for r in ref])
    rcounts = np.array([r[4] for r in ref])
    ravec = np.array([r[3] for r in ref])
    rq = np.array([r[0] for r in ref])
    rp = np.array([r[1] for r in ref])

    differ = np.nonzero(code != rcode)[0]
    # a decision may only differ where the uniform sits within rounding of the acceptance probability
    for c in differ:
        gap = np.abs(U[c, :nprops] - ravec[c]).min()
        assert gap < 1e-7, "chain %d: codes %d vs %d with |u - alpha| = %.3e" % (c, code[c], rcode[c], gap)
    same = code == rcode
    assert (~same).sum() <= 2, "%d decisions differ" % (~same).sum()
    assert np.array_equal(counts[same], rcounts[same]), "[Hcount, Vgcount, leapcount] differ"

    stats = {"config": "funnel d=50, DRHMC nleap=150 step_size=0.2 nprops=3 stepfactor=2, %d chains from exact funnel "
                       "draws, Philox momenta (k_transition_ws) read back through p0_out, host-computed uniforms" % C,
             "decisions_differing": int((~same).sum()), "by_code": {}}
    for k in range(nprops + 1):
        m = same & (rcode == k)
        if not m.any():
            continue
        eq, ep = _relerr(q[m], rq[m]).max(axis=1), _relerr(p[m], rp[m]).max(axis=1)
        stats["by_code"][str(k)] = {
            "chains": int(m.sum()), "grad_evals_mean": float(rcounts[m, 1].mean()),
            "q_relerr_max": float(eq.max()), "q_relerr_p99": float(np.quantile(eq, 0.99)), "q_relerr_median": float(np.median(eq)),
            "p_relerr_max": float(ep.max()), "p_relerr_median": float(np.median(ep))}
    ea = _relerr(avec[same], ravec[same])
    stats["avec_relerr_max"] = float(ea.max())
    stats["avec_relerr_p99"] = float(np.quantile(ea, 0.99))
    stats["avec_note"] = ("alpha = exp(H0 - Hj) * num / den carries the ABSOLUTE error of energies of size up to 1e5 in the "
                          "neck of the funnel; |u - alpha| of the worst chain is far above it (no decision differs)")
    stats["code_fractions"] = [float((rcode == k).mean()) for k in range(nprops + 1)]
    _dump("headline_parity.json", stats)
    print(json.dumps(stats))
    assert len(stats["by_code"]) == nprops + 1, "the case must exercise every stage: %s" % stats["code_fractions"]
    # Rejected chains return q0 unchanged: exact.  Accepted chains carry the rounding differences of a 150- to
    # 600-step trajectory (fused multiply-adds on the GPU, none in numpy) amplified by the funnel's dynamics.
    assert stats["by_code"]["0"]["q_relerr_max"] == 0.0
    for k in range(1, nprops + 1):
        s = stats["by_code"][str(k)]
        assert s["q_relerr_max"] < 1e-10, (k, s)       # measured: 6e-12 (stage 0), 2e-12, 7e-13
        assert s["p_relerr_max"] < 1e-9, (k, s)        # measured: 1.5e-10 on one chain of 3053, median 3e-14


# ------------------------------------------------------------------ error against trajectory length
def test_error_vs_trajectory_length_curve():
    """max / median relative error of (q, p) after L leapfrog steps against the numpy oracle, funnel d=50, on
    stationary draws: once at fixed trajectory length 30 (eps = 30/L; L = 150, 300, 600 are the three stages
    of the headline), once at fixed eps = 0.05."""
    d, C = 50, 512
    om = omodels.Funnel(d)
    s = pyhmc.HMC(models.Funnel(d), device=_dev())
    q0 = _exact_funnel(C, d, 3)
    p0 = np.random.default_rng(4).standard_normal((C, d))
    kin = samplers.Kinetics()
    rows = []
    for mode in ("length30", "eps0.05"):
        for L in ((150, 300, 600, 1200) if mode == "length30" else (10, 20, 50, 150, 300, 600)):
            eps = 30.0 / L if mode == "length30" else 0.05
            q1, p1 = s.leapfrog(q0, p0, L, eps)
            ref = [samplers.leapfrog(om, kin, q0[c], p0[c], L, eps, samplers.Counters()) for c in range(C)]
            rq, rp = np.array([r[0] for r in ref]), np.array([r[1] for r in ref])
            ok = np.isfinite(rq).all(axis=1) & np.isfinite(rp).all(axis=1) & (np.abs(rq).max(axis=1) < 1e6)
            assert ok.sum() >= C // 4, "only %d of %d reference trajectories stay finite at L=%d eps=%g" % (ok.sum(), C, L, eps)
            e = np.maximum(_relerr(q1[ok], rq[ok]).max(axis=1), _relerr(p1[ok], rp[ok]).max(axis=1))
            rows.append({"mode": mode, "L": L, "eps": eps, "chains": int(ok.sum()),
                         "diverged_in_both": int(C - ok.sum()), "relerr_median": float(np.median(e)),
                         "relerr_p90": float(np.quantile(e, 0.9)), "relerr_p99": float(np.quantile(e, 0.99)),
                         "relerr_max": float(e.max())})
            # divergent trajectories (eps too large for the neck of the funnel) must diverge on both sides
            bad_gpu = ~(np.isfinite(q1).all(axis=1) & (np.abs(q1).max(axis=1) < 1e6))
            assert (bad_gpu & ok).sum() <= 1
    _dump("err_vs_L.json", rows)
    for r in rows:
        print(json.dumps(r))
    short = [r for r in rows if r["L"] <= 50 and r["mode"] == "eps0.05"]
    assert all(r["relerr_median"] < 1e-12 for r in short), short
    assert all(r["relerr_median"] < 1e-10 for r in rows), rows
