"""Generate tests/golden/*.npz by running the UNMODIFIED reference classes
(/root/reference/src, imported via oracle/ref_loader.py) on replayed draws.

Run in the build container only:  python -m oracle.make_goldens
Each file stores the inputs (model parameters, q0, normal draws P[T,C,d], uniform
draws U[T,C,nu], sampler arguments) and what the reference returned for every
(transition, chain).  The GPU box checks against these without the reference tree.
"""
import contextlib
import io
import json
import os

import numpy as np

from . import models, ref_loader, replay

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")


def build_model(spec):
    kind = spec["kind"]
    if kind == "funnel":
        return models.Funnel(spec["d"], spec.get("sigma", 3.0))
    if kind == "rosenbrock":
        return models.Rosenbrock(spec["d"], spec.get("a", 1.0), spec.get("b", 100.0), spec.get("T", 20.0))
    if kind == "gaussian":
        return models.CorrGaussian.ar1(spec["d"], spec.get("rho", 0.9))
    raise ValueError(kind)


CASES = [
    dict(name="hmc_funnel10", cls="pyhmc.HMC", model=dict(kind="funnel", d=10), C=8, T=5, seed=101,
         args=dict(N=20, step_size=0.1), nu=1),
    dict(name="hmc_funnel50_metric", cls="pyhmc.PyHMC", model=dict(kind="funnel", d=50), C=6, T=4, seed=102,
         args=dict(N=15, step_size=0.08), nu=1, metric=True),
    dict(name="alghmc_rosen8", cls="algorithms.HMC", model=dict(kind="rosenbrock", d=8), C=8, T=5, seed=103,
         args=dict(N=12, step_size=0.03), nu=1),
    dict(name="alghmc_gauss12", cls="algorithms.HMC", model=dict(kind="gaussian", d=12), C=8, T=4, seed=104,
         args=dict(N=10, step_size=0.15), nu=1),
    dict(name="drhmc_funnel50", cls="algorithms.DRHMC", model=dict(kind="funnel", d=50), C=24, T=4, seed=105,
         args=dict(nleap=10, step_size=0.4, nprops=3, stepfactor=2), nu=3),
    dict(name="drhmc_funnel10_np4", cls="algorithms.DRHMC", model=dict(kind="funnel", d=10), C=16, T=4, seed=106,
         args=dict(nleap=6, step_size=0.7, nprops=4, stepfactor=1.5), nu=4),
    dict(name="drhmc_rosen6", cls="algorithms.DRHMC", model=dict(kind="rosenbrock", d=6), C=16, T=4, seed=107,
         args=dict(nleap=8, step_size=0.12, nprops=3, stepfactor=2), nu=3),
]


def run_case(case):
    modname, clsname = case["cls"].split(".")
    mod = ref_loader.load(modname)
    model = build_model(case["model"])
    d, C, T = model.d, case["C"], case["T"]
    inv = np.linspace(0.5, 2.0, d) if case.get("metric") else None
    P, U = replay.make_draws(case["seed"], T, C, d, case["nu"])
    q0 = np.random.default_rng(case["seed"] + 1000).uniform(-1, 1, (C, d))
    if case["model"]["kind"] == "rosenbrock":
        q0 = 1.0 + 0.1 * q0
    cls = getattr(mod, clsname)
    if modname == "algorithms":
        ref = cls(model.log_prob, model.grad_log_prob, invmetric_diag=inv)
    else:
        ref = cls(model.log_prob, model.grad_log_prob, invmetric_diag=inv)
    a = case["args"]
    is_dr = clsname == "DRHMC"
    q = np.zeros((T, C, d)); p = np.zeros((T, C, d)); code = np.zeros((T, C)); counts = np.zeros((T, C, 3), dtype=np.int64)
    aux = np.zeros((T, C, a["nprops"] if is_dr else 2))
    sink = io.StringIO()
    with np.errstate(all="ignore"), contextlib.redirect_stdout(sink):
        for c in range(C):
            qc = q0[c]
            for t in range(T):
                with replay.patched_numpy_random(replay.Draws([P[t, c]], U[t, c])):
                    if is_dr:
                        out = ref.step(qc, a["nleap"], a["step_size"], a["nprops"], a["stepfactor"])
                    elif clsname == "PyHMC":
                        out = ref.hmc_step(qc, a["N"], a["step_size"])
                    else:
                        out = ref.step(qc, a["N"], a["step_size"])
                qc = out[0]
                q[t, c], p[t, c], code[t, c], aux[t, c], counts[t, c] = out[0], out[1], out[2], out[3], out[4]
    np.savez_compressed(os.path.join(OUT, case["name"] + ".npz"), spec=json.dumps(case), q0=q0, P=P, U=U,
                        invmetric=np.zeros(0) if inv is None else inv, q=q, p=p, code=code, aux=aux, counts=counts)
    return code


MALA_CASES = [
    dict(name="mala_funnel10", model=dict(kind="funnel", d=10), C=8, T=6, seed=201, eps=0.15),
    dict(name="mala_rosen8", model=dict(kind="rosenbrock", d=8), C=8, T=6, seed=202, eps=0.04),
    dict(name="mala_gauss12", model=dict(kind="gaussian", d=12), C=8, T=6, seed=203, eps=0.2),
]

ADHMC_CASES = [
    dict(name="adhmc_funnel10", model=dict(kind="funnel", d=10), C=12, T=4, seed=301,
         args=dict(Nleap=8, smin=0.01, smax=0.5, ntry=6, nsteps=2)),
    dict(name="adhmc_rosen4", model=dict(kind="rosenbrock", d=4), C=12, T=4, seed=302,
         args=dict(Nleap=10, smin=0.005, smax=0.2, ntry=8, nsteps=1)),
]


ADHMC_CASES.append(dict(name="adhmc_funnel10_tint", model=dict(kind="funnel", d=10), C=12, T=4, seed=303,
                        args=dict(Nleap=0, Tint=1.5, smin=0.02, smax=0.5, ntry=6, nsteps=2)))

WALK_CASES = [
    dict(name="adwalk_funnel10", cls="AdHMC", model=dict(kind="funnel", d=10), C=16, T=4, seed=501,
         args=dict(Nleap=24, nleap=4, smin=0.02, smax=0.8, ntry_eps0=6, nsteps_eps0=2)),
    dict(name="adwalk_funnel50_tint", cls="AdHMC_tint", model=dict(kind="funnel", d=50), C=12, T=3, seed=502,
         args=dict(Tint=2.0, ntimes=5, smin=0.02, smax=0.8, ntry_eps0=6, nsteps_eps0=2)),
    dict(name="adwalk_rosen8_3ratios", cls="AdHMC", model=dict(kind="rosenbrock", d=8), C=16, T=4, seed=503,
         args=dict(Nleap=20, nleap=4, ratios=[0.5, 1.0, 2.0], pwts0=[1.0, 2.0, 1.0], smin=0.005, smax=0.2, ntry_eps0=6,
                   nsteps_eps0=2, logeps=False)),
    dict(name="adwalk_funnel10_wide", cls="AdHMC", model=dict(kind="funnel", d=10), C=48, T=4, seed=504,
         args=dict(Nleap=24, nleap=4, smin=0.05, smax=3.0, ntry_eps0=6, nsteps_eps0=1)),
]


def run_walk(case):
    """hmc_step of adaptivehmc.AdHMC / AdHMC_tint; N+2 replayed uniforms per transition (np.random.choice included).
    The reference's blow-up return spells NaN as np.NaN, which numpy 2 no longer has: it is given back for the run."""
    ad = ref_loader.load("adaptivehmc")
    model = build_model(case["model"])
    d, C, T, a = model.d, case["C"], case["T"], case["args"]
    N = a["ntimes"] if "ntimes" in a else a["Nleap"] // a["nleap"]
    P, U = replay.make_draws(case["seed"], T, C, d, N + 2)
    q0 = _q0(case, model)
    ref = getattr(ad, case["cls"])(model.log_prob, model.grad_log_prob)
    q = np.zeros((T, C, d)); p = np.zeros((T, C, d)); code = np.zeros((T, C))
    probs = np.zeros((T, C, 3)); checks = np.zeros((T, C, 3, N)); counts = np.zeros((T, C, 3), dtype=np.int64)
    had_nan = hasattr(np, "NaN")
    if not had_nan:
        np.NaN = np.nan
    try:
        with np.errstate(all="ignore"), contextlib.redirect_stdout(io.StringIO()):
            for c in range(C):
                qc = q0[c]
                for t in range(T):
                    with replay.patched_numpy_random(replay.Draws([P[t, c]], U[t, c])):
                        out = ref.hmc_step(qc, **a)
                    qc = out[0]
                    q[t, c], p[t, c], code[t, c], probs[t, c], checks[t, c], counts[t, c] = out
    finally:
        if not had_nan:
            del np.NaN
    np.savez_compressed(os.path.join(OUT, case["name"] + ".npz"), spec=json.dumps(case), q0=q0, P=P, U=U, q=q, p=p,
                        code=code, probs=probs, checks=checks, counts=counts)
    return code


def _q0(case, model):
    q0 = np.random.default_rng(case["seed"] + 1000).uniform(-1, 1, (case["C"], model.d))
    if case["model"]["kind"] == "rosenbrock":
        q0 = 1.0 + 0.1 * q0
    return q0


def run_mala(case):
    mala = ref_loader.load("mala")
    model = build_model(case["model"])
    d, C, T = model.d, case["C"], case["T"]
    P, U = replay.make_draws(case["seed"], T, C, d, 1)
    q0 = _q0(case, model)
    ref = mala.MALA(model.log_prob, model.grad_log_prob)
    q = np.zeros((T, C, d)); Vg = np.zeros((T, C, d)); V = np.zeros((T, C)); code = np.zeros((T, C))
    with np.errstate(all="ignore"), contextlib.redirect_stdout(io.StringIO()):
        for c in range(C):
            qc, Vc, Vgc = q0[c], None, None
            for t in range(T):
                with replay.patched_numpy_random(replay.Draws([P[t, c]], U[t, c])):
                    qc, acc, Vc, Vgc = ref.step(qc, case["eps"], Vc, Vgc)
                q[t, c], code[t, c], V[t, c], Vg[t, c] = qc, acc, Vc, Vgc
    np.savez_compressed(os.path.join(OUT, case["name"] + ".npz"), spec=json.dumps(case), q0=q0, P=P, U=U, q=q,
                        code=code, V=V, Vg=Vg)
    return code


def run_adhmc(case):
    ad = ref_loader.load("adaptivehmc")
    model = build_model(case["model"])
    d, C, T = model.d, case["C"], case["T"]
    P, U = replay.make_draws(case["seed"], T, C, d, 3)
    q0 = _q0(case, model)
    ref = ad.AdHMC_eps0(model.log_prob, model.grad_log_prob)
    a = case["args"]
    q = np.zeros((T, C, d)); p = np.zeros((T, C, d)); code = np.zeros((T, C))
    probs = np.zeros((T, C, 3)); aux = np.zeros((T, C, 3)); counts = np.zeros((T, C, 3), dtype=np.int64)
    with np.errstate(all="ignore"), contextlib.redirect_stdout(io.StringIO()):
        for c in range(C):
            qc = q0[c]
            for t in range(T):
                with replay.patched_numpy_random(replay.Draws([P[t, c]], U[t, c])):
                    out = ref.hmc_step(qc, a["Nleap"], smin=a["smin"], smax=a["smax"], Tint=a.get("Tint", 0), ntry=a["ntry"],
                                       nsteps=a["nsteps"])
                qc = out[0]
                q[t, c], p[t, c], code[t, c], probs[t, c], aux[t, c], counts[t, c] = out
    np.savez_compressed(os.path.join(OUT, case["name"] + ".npz"), spec=json.dumps(case), q0=q0, P=P, U=U, q=q, p=p,
                        code=code, probs=probs, aux=aux, counts=counts)
    return code


PROBDR_CASES = [
    dict(name="probdr_metrop_funnel10", cls="algorithms.Prob_DRHMC_Metrop", model=dict(kind="funnel", d=10), C=16, T=5,
         seed=401, args=dict(m=3, N=6, ss=0.6, fsub=2.0)),
    dict(name="probdr_hastings_funnel10", cls="algorithms.Prob_DRHMC_Hastings", model=dict(kind="funnel", d=10), C=16,
         T=5, seed=402, args=dict(m=3, N=6, ss=0.6, fsub=2.0)),
    dict(name="probdr_tries_funnel50", cls="pyhmc.PyHMC_multistep_tries", model=dict(kind="funnel", d=50), C=12, T=4,
         seed=403, args=dict(m=4, N=10, ss=0.4, fsub=1.5)),
    dict(name="probdr_tries3_rosen8", cls="pyhmc.PyHMC_multistep_tries3", model=dict(kind="rosenbrock", d=8), C=12, T=4,
         seed=404, args=dict(m=3, N=8, ss=0.12, fsub=2.0)),
]


def run_probdr(case):
    """multi_step of the probabilistic delayed-rejection classes, 2m-1 replayed uniforms per transition."""
    modname, clsname = case["cls"].split(".")
    model = build_model(case["model"])
    d, C, T, a = model.d, case["C"], case["T"], case["args"]
    m = a["m"]
    P, U = replay.make_draws(case["seed"], T, C, d, 2 * m - 1)
    q0 = _q0(case, model)
    ref = getattr(ref_loader.load(modname), clsname)(model.log_prob, model.grad_log_prob)
    q = np.zeros((T, C, d)); p = np.zeros((T, C, d)); code = np.zeros((T, C)); avec = np.zeros((T, C, m))
    ptries = np.zeros((T, C, m)); counts = np.zeros((T, C, 3))
    with np.errstate(all="ignore"), contextlib.redirect_stdout(io.StringIO()):
        for c in range(C):
            qc = q0[c]
            for t in range(T):
                with replay.patched_numpy_random(replay.Draws([P[t, c]], U[t, c])):
                    qc, pc, acc, av, cnt, pt = ref.multi_step(m, qc, a["N"], a["ss"], a["fsub"])
                q[t, c], p[t, c], code[t, c], avec[t, c], counts[t, c], ptries[t, c] = qc, pc, acc, av, cnt, pt
    np.savez_compressed(os.path.join(OUT, case["name"] + ".npz"), spec=json.dumps(case), q0=q0, P=P, U=U, q=q, p=p,
                        code=code, avec=avec, counts=counts, ptries=ptries)
    return code


def run_logistic():
    """pyhmc.HMC.step on the Bayesian logistic-regression model (float64 reference arithmetic on
    float32-representable data); the GPU path runs in float32 and is held to 1e-5."""
    pyhmc = ref_loader.load("pyhmc")
    rng = np.random.default_rng(401)
    N, d, C, T = 3000, 24, 12, 3
    X = (rng.standard_normal((N, d)) / np.sqrt(d)).astype(np.float32)
    theta_star = rng.standard_normal(d)
    y = (rng.random(N) < 1.0 / (1.0 + np.exp(-(X.astype(np.float64) @ theta_star)))).astype(np.float32)
    model = models.LogisticRegression(X.astype(np.float64), y.astype(np.float64), 1.0)
    ref = pyhmc.HMC(model.log_prob, model.grad_log_prob)
    P, U = replay.make_draws(402, T, C, d, 1)
    P = P.astype(np.float32).astype(np.float64)       # momenta exactly representable in float32
    q0 = (0.1 * rng.standard_normal((C, d))).astype(np.float32).astype(np.float64)
    args = dict(N=10, step_size=0.01)
    q = np.zeros((T, C, d)); p = np.zeros((T, C, d)); code = np.zeros((T, C)); Hs = np.zeros((T, C, 2))
    counts = np.zeros((T, C, 3), dtype=np.int64)
    for c in range(C):
        qc = q0[c]
        for t in range(T):
            with replay.patched_numpy_random(replay.Draws([P[t, c]], U[t, c])):
                out = ref.step(qc, args["N"], args["step_size"])
            qc = out[0]
            q[t, c], p[t, c], code[t, c], Hs[t, c], counts[t, c] = out
    np.savez_compressed(os.path.join(OUT, "logit_hmc.npz"), spec=json.dumps(dict(args=args, tau=1.0)), X=X, y=y, q0=q0, P=P,
                        U=U, q=q, p=p, code=code, Hs=Hs, counts=counts)
    return code


def main():
    assert ref_loader.available(), "the reference tree is needed to make goldens"
    os.makedirs(OUT, exist_ok=True)
    for case in CASES:
        code = run_case(case)
        vals, n = np.unique(code, return_counts=True)
        print("%-24s codes %s" % (case["name"], dict(zip(vals.tolist(), n.tolist()))))
    vals, n = np.unique(run_logistic(), return_counts=True)
    print("%-24s codes %s" % ("logit_hmc", dict(zip(vals.tolist(), n.tolist()))))
    for cases, fn in ((MALA_CASES, run_mala), (ADHMC_CASES, run_adhmc), (PROBDR_CASES, run_probdr), (WALK_CASES, run_walk)):
        for case in cases:
            vals, n = np.unique(fn(case), return_counts=True)
            print("%-24s codes %s" % (case["name"], dict(zip(vals.tolist(), n.tolist()))))


if __name__ == "__main__":
    main()
