#!/usr/bin/env python
"""Benchmark of the sampler hot path on the BASELINE.json headline workload:
Neal's funnel d=50, delayed-rejection HMC (nleap=150, step 0.2 -> lpath 30 as
scripts_old/jobscripts/run_funnel2.sh:21-22; nprops=3, stepfactor=2), 2^20 chains.
Chains are independent, so N GPUs shard them by global chain id with no data-path
collective.  Default: weak scaling -- every GPU runs the N=1 workload (2^20 chains per
GPU, N*2^20 in total); at N > 1 the line also carries a `strong` block: 2^20 chains in
total, as the BASELINE config words it (`--scaling strong` makes that the headline).

A "step" is one DRHMC transition of every chain with its running moments; the timed
region is ONE hmcb_drhmc_sample call of K steps.  Prints ONE JSON line on stdout.
  value      : gradient evaluations/s (the reference's Vgcount counter, src/pyhmc.py:21-23),
               chain state resident in HBM; per_call: the same by single hmcb_drhmc_step calls
  e2e        : same metric through the public API with HOST (pinned) buffers, copies timed,
               and link_floor_ms (the copies alone)
  roofline   : algorithmic rate of the timed call over the FP64 peak measured in the run,
               per-stage kernel figures beside it
  strong, secondary, stationary, autocorr, gates, cpu_baseline, config1_cpu, clocks:
               see DESIGN.md "Measurement"
`--impl reference` times the CPU path (the numpy oracle port of the reference's
DRHMC.step; the reference itself cannot travel to the GPU box) on all host cores.
"""
import argparse
import json
import multiprocessing as mp
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries exactly one JSON line: NCCL's own banner/debug lines go to stderr
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")

FLOPS_PER_GRAD = {"funnel": lambda d: 7 * d}  # 7d flops + 1 exp per gradient evaluation


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="hmcb", choices=["hmcb", "reference"])
    ap.add_argument("--chains", type=int, default=1 << 20, help="chains per GPU (weak) / in total (strong)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--ndim", type=int, default=50)
    ap.add_argument("--nleap", type=int, default=150)
    ap.add_argument("--step_size", type=float, default=0.2)
    ap.add_argument("--nprops", type=int, default=3)
    ap.add_argument("--stepfactor", type=float, default=2.0)
    ap.add_argument("--burnin", type=int, default=200)
    ap.add_argument("--seed", type=int, default=2)
    ap.add_argument("--cpu_seconds", type=float, default=12.0)
    ap.add_argument("--no_cpu_baseline", action="store_true")
    ap.add_argument("--no_e2e", action="store_true")
    ap.add_argument("--e2e_steps", type=int, default=0)
    ap.add_argument("--no_secondary", action="store_true")
    ap.add_argument("--ess_steps", type=int, default=64, help="transitions of the stationary ESS window (0: skip)")
    ap.add_argument("--acf_steps", type=int, default=256, help="stored transitions of the autocorrelation window (0: skip)")
    return ap.parse_args()


# ------------------------------------------------------------------ CPU arm (oracle port on the host cores)
_W = {}


def _cpu_init(d, seed, chains_per_proc, nleap, eps, nprops, f):
    os.environ["OMP_NUM_THREADS"] = "1"
    from oracle import models as om
    ident = mp.current_process()._identity
    wid = ident[0] if ident else 0
    rng = np.random.Generator(np.random.PCG64(seed * 1000 + wid))
    _W.update(model=om.Funnel(d), rng=rng, q=rng.uniform(-1, 1, (chains_per_proc, d)),
              args=(nleap, eps, nprops, f))


def _cpu_step(n_chains):
    """Advance this worker's first n_chains chains by one DRHMC transition."""
    from oracle import samplers

    class R:  # the reference draws from numpy's global generator; here a per-worker one
        normal = staticmethod(lambda size=None: _W["rng"].standard_normal(size))
        uniform = staticmethod(lambda low=0.0, high=1.0, size=None: _W["rng"].uniform(low, high, size))

    nleap, eps, nprops, f = _W["args"]
    grads = leaps = 0
    for c in range(n_chains):
        q, p, code, avec, cnt = samplers.drhmc_step(_W["model"], _W["q"][c], nleap, eps, nprops, f, R)
        _W["q"][c] = q
        grads += cnt[1]
        leaps += cnt[2]
    return grads, leaps


class CpuArm:
    def __init__(self, a, chains_per_proc):
        self.P = os.cpu_count() or 1
        self.m = chains_per_proc
        ctx = mp.get_context("fork")
        self.pool = ctx.Pool(self.P, initializer=_cpu_init,
                             initargs=(a.ndim, a.seed, chains_per_proc, a.nleap, a.step_size, a.nprops, a.stepfactor))

    def step(self, n_chains=None):
        n = self.m if n_chains is None else n_chains
        res = self.pool.map(_cpu_step, [n] * self.P)
        return sum(r[0] for r in res), sum(r[1] for r in res)

    def close(self):
        self.pool.close()
        self.pool.join()


def cpu_baseline(a, seconds):
    """Bounded sample: every host core runs its own chains (one chain per process at a
    time, as the reference's one-chain-per-MPI-rank layout, scripts_old/jobscripts/
    run_funnel2.sh:16,42) for about `seconds` of wall time."""
    arm = CpuArm(a, 64)
    arm.step(1)                             # warm the workers (imports, first-touch)
    grads = leaps = nsteps = 0
    t0 = time.time()
    while time.time() - t0 < seconds:
        g, l = arm.step(4)
        grads, leaps, nsteps = grads + g, leaps + l, nsteps + 1
    dt = time.time() - t0
    n = 4 * nsteps
    arm.close()
    return {"value": grads / dt, "unit": "grad_evals/s", "cores": arm.P, "kind": "port",
            "sample": "%d DRHMC chain-transitions on each of %d processes (oracle/samplers.py drhmc_step, numpy), "
                      "%.1f s wall" % (n, arm.P, dt),
            "leapfrog_steps_per_s": (grads - leaps) / dt}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    m = 16
    arm = CpuArm(a, m)
    for _ in range(max(a.warmup, 1)):
        arm.step(2)
    t0 = time.time()
    grads = leaps = 0
    for _ in range(a.steps):
        g, l = arm.step(m)
        grads += g
        leaps += l
    dt = time.time() - t0
    arm.close()
    val = grads / dt
    line = {"impl": "reference", "metric": "grad_evals_per_sec", "value": val, "unit": "grad_evals/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps,
            "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(a, sample="each step = %d chains x %d processes x 1 transition" % (m, arm.P)),
            "cpu_baseline": {"value": val, "unit": "grad_evals/s", "cores": arm.P, "kind": "port",
                             "sample": "%d steps of %d chain-transitions on %d processes" % (a.steps, m * arm.P, arm.P)},
            "e2e": {"value": val, "unit": "grad_evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "leapfrog_steps_per_sec": (grads - leaps) / dt, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def total_chains(a):
    n = int(os.environ.get("WORLD_SIZE", "1"))
    return a.chains * n if a.scaling == "weak" else a.chains


def workload_config(a, **extra):
    n = int(os.environ.get("WORLD_SIZE", "1"))
    tot = total_chains(a)
    cfg = {"workload": "funnel d=%d sigma=3, DRHMC nleap=%d step_size=%g nprops=%d stepfactor=%g, %d chains total "
                       "(%d per GPU, %s scaling), q0~U(-1,1), %d burn-in transitions before timing" %
                       (a.ndim, a.nleap, a.step_size, a.nprops, a.stepfactor, tot, tot // n, a.scaling, a.burnin),
           "chains": tot, "chains_per_gpu": tot // n, "d": a.ndim,
           "cache": "inputs larger than L2 (q is %.0f MB per rank)" % (tot // n * a.ndim * 8 / 1e6)}
    cfg.update(extra)
    return cfg


# ------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """nvidia-smi in a subprocess (the profiling recipe's clocks line), sampling every
    50 ms while the timed region runs.  In-process NVML polling was tried and rejected:
    its driver calls stalled kernel launches and inflated ms_per_step."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append((time.time(), [x.strip() for x in line.split(",")]))
        except Exception as e:
            self.error = repr(e)

    def stop(self):
        if self.proc:
            self.proc.terminate()

    def summary(self, t0, t1):
        good = [(t, r) for t, r in self.rows if len(r) >= 7]
        rows = [r for t, r in good if t0 <= t <= t1 + 0.05]
        note = "sampled inside the timed region"
        if not rows and good:   # region shorter than the sampling period: nearest sample
            rows = [min(good, key=lambda tr: abs(tr[0] - 0.5 * (t0 + t1)))[1]]
            note = "timed region shorter than the 50 ms sampling period: nearest sample"
        if not rows:
            return {"error": getattr(self, "error", "no samples")}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        sm = sorted(float(r[0]) for r in rows)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons,
                "power_w_max": max(float(r[2]) for r in rows), "samples": len(rows), "note": note}


# ------------------------------------------------------------------ GPU arm
def measured_peaks():
    """Driver-measured HBM / bf16 peaks of this pool's B200 (MEASURED_PEAKS.json), else the profiling recipe's fallback."""
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"], "bf16_tflops_sustained": p.get("bf16_tflops_sustained"),
                "source": "MEASURED_PEAKS.json"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": None, "source": "fallback (B200_PROFILING.md)"}


def fp64_peak_tflops(eng, dev):
    """FP64 pipe peak of this GPU, measured now with a register-resident DFMA loop (MEASURED_PEAKS.json has no FP64 figure)."""
    import torch
    from hmc_b200 import _lib
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


def timed_region(fn, dev, world):
    """fn() enqueues the work; barrier + synchronize on both sides, CUDA events on the launching stream, max over ranks."""
    import torch
    from hmc_b200 import dist
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.time()
    start.record()
    out = fn()
    end.record()
    torch.cuda.synchronize()
    dist.barrier()
    t1 = time.time()
    ms = torch.tensor([start.elapsed_time(end)], dtype=torch.float64, device=dev)
    if world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    return float(ms[0]), out, t0, t1


def exact_funnel(C, d, seed, dev):
    import torch
    g = torch.Generator(device=dev).manual_seed(seed)
    v = 3.0 * torch.randn(C, 1, generator=g, dtype=torch.float64, device=dev)
    th = torch.exp(v / 2) * torch.randn(C, d - 1, generator=g, dtype=torch.float64, device=dev)
    return torch.cat([v, th], dim=1).contiguous()


def dr_block(a, dev, world, lo, C, q0, seed, K, warmup, burnin, want_stage_times, fp64_peak):
    """The headline measurement on C chains of this rank starting from q0: burn-in, warm-up, K timed transitions in one
    hmcb_drhmc_sample call (chain ranges pipelined over streams, running moments folded in), then the same K
    transitions as one hmcb_drhmc_step call + one moments kernel each, then per-stage kernel times."""
    import torch
    from hmc_b200 import algorithms, dist, models, utils
    d = a.ndim
    model = models.Funnel(d)
    dr = algorithms.DRHMC(model, device=dev, seed=seed, chain_offset=lo)
    eng = dr.engine
    eng._ensure()
    q = q0
    if burnin:
        q = eng.drhmc_sample(q, a.nleap, a.step_size, a.nprops, a.stepfactor, burnin, want_avec=False)["q"]
    mom = utils.RunningMoments(eng, C, d)
    nxt = torch.empty_like(q)
    counts = torch.zeros((C, 3), dtype=torch.int32, device=dev)
    code = torch.zeros(C, dtype=torch.int32, device=dev)
    nacc = torch.zeros(C, dtype=torch.int32, device=dev)

    def sample(qi, n):
        return eng.drhmc_sample(qi, a.nleap, a.step_size, a.nprops, a.stepfactor, n, msum=mom.sum, msumsq=mom.sumsq,
                                n_accept=nacc, q_out=nxt, want_avec=False, code_out=code, counts_out=counts)

    if warmup:
        sample(q, warmup)
        q, nxt = nxt, q
    mom.reset()
    nacc.zero_()
    ms_total, _, t0, t1 = timed_region(lambda: sample(q, K), dev, world)
    q, nxt = nxt, q
    mom.n = K
    stats = torch.cat([counts.sum(dim=0, dtype=torch.float64), nacc.sum(dtype=torch.float64)[None]])
    code_frac = (torch.bincount(code.long(), minlength=a.nprops + 1).double() / C).tolist()
    dist.all_reduce_sum(stats)
    grads, leaps, acc = float(stats[1]), float(stats[2]), float(stats[3])
    res = {"ms_total": ms_total, "t0": t0, "t1": t1, "grads": grads, "leaps": leaps, "accepted": acc, "code_frac": code_frac,
           "mom": mom, "q": q, "eng": eng, "model": model, "n_split": 4 if C >= (1 << 17) else (2 if C >= (1 << 14) else 1)}
    if os.environ.get("HMCB_SAMPLE_SPLIT"):
        res["n_split"] = int(os.environ["HMCB_SAMPLE_SPLIT"])

    # the same transitions, one hmcb_drhmc_step call + one hmcb_moments_accumulate per transition (round 1's timed loop)
    s1, s2 = torch.zeros_like(q), torch.zeros_like(q)
    cl = torch.zeros((K, C, 3), dtype=torch.int32, device=dev)

    def per_call():
        qi, qo = q, nxt
        for k in range(K):
            eng.drhmc_step(qi, a.nleap, a.step_size, a.nprops, a.stepfactor, want_p=False, q_out=qo, want_avec=False,
                           code_out=code, counts_out=cl[k])
            eng.moments_accumulate(qo, s1, s2)
            qi, qo = qo, qi
    ms_pc, _, _, _ = timed_region(per_call, dev, world)
    g_pc = cl[:, :, 1].sum(dtype=torch.float64)[None]
    dist.all_reduce_sum(g_pc)
    res["per_call"] = {"ms_per_step": ms_pc / K, "value": float(g_pc[0]) / (ms_pc * 1e-3),
                       "what": "one hmcb_drhmc_step call (2 chain ranges on 2 streams) + one hmcb_moments_accumulate per transition"}

    if want_stage_times:
        # Per-stage kernel times (CUDA events recorded inside hmcb_drhmc_step between the stage launches, stages serialised
        # on one stream), twice: with the plain stage-0 kernel -- the one the timed region launches -- and with the
        # warp-specialised one that single hmcb_drhmc_step calls use.  Gradient work per stage from the pass's own outputs.
        KR = min(K, 4)
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(a.nprops + 1)] for _ in range(KR)]
        for row in evs:
            for ev in row:
                ev.record()
        codes = torch.zeros((KR, C), dtype=torch.int32, device=dev)
        passes = {}
        for name, env in (("plain", "1"), ("ws", None)):
            if env is None:
                os.environ.pop("HMCB_NO_WS", None)
            else:
                os.environ["HMCB_NO_WS"] = env
            qi, qo = q.clone(), nxt
            for k in range(KR):
                eng.drhmc_step(qi, a.nleap, a.step_size, a.nprops, a.stepfactor, want_p=False, q_out=qo, stage_events=evs[k],
                               want_avec=False, code_out=codes[k], counts_out=cl[k])
                qi, qo = qo, qi
            torch.cuda.synchronize()
            st = torch.tensor([[evs[k][j].elapsed_time(evs[k][j + 1]) for j in range(a.nprops)] for k in range(KR)],
                              dtype=torch.float64).mean(dim=0)
            # chains entering stage j: those not accepted before it (code j' + 1 = accepted at stage j', 0 = never)
            tot = float(cl[:KR, :, 1].sum(dtype=torch.float64)) / KR
            g = [C * (max(a.nleap, 1) + 1.0)]
            if a.nprops >= 2:
                n1 = float((codes != 1).sum()) / KR
                L1 = int(a.nleap * a.stepfactor)
                g1 = n1 * ((max(L1, 1) + 1) + (max(a.nleap, 1) + 1)) if a.nprops >= 2 else 0.0
                g.append(g1)
                if a.nprops >= 3:
                    g.append(tot - g[0] - g1)          # everything deeper (stages >= 2 together when nprops > 3)
            passes[name] = {"stage_ms": st.tolist(), "stage_grad_evals": g}
        os.environ.pop("HMCB_NO_WS", None)
        mev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        mev[0].record()
        for _ in range(4):
            eng.moments_accumulate(q, s1, s2)
        mev[1].record()
        torch.cuda.synchronize()
        res["passes"] = passes
        res["moments_ms"] = mev[0].elapsed_time(mev[1]) / 4
        res["stage_passes"] = KR
    return res


def stationary_block(a, dev, world, lo, C, seed):
    """ESS and moment checks on a STATIONARY ensemble: chains start from exact funnel draws (v ~ N(0, 3), theta ~
    N(0, e^v)), T transitions in one sample call with running moments; pooled moments against the analytic ones with
    their Monte Carlo errors, utils.get_ess_true (src/utils.py:8-27) over the window, all ranks reduced over NCCL."""
    import torch
    from hmc_b200 import algorithms, dist, models, utils
    d, T = a.ndim, a.ess_steps
    model = models.Funnel(d)
    eng = algorithms.DRHMC(model, device=dev, seed=seed + 17, chain_offset=lo).engine
    eng._ensure()
    q = exact_funnel(C, d, seed + 101 + lo % 9973, dev)
    mom = utils.RunningMoments(eng, C, d)
    counts = torch.zeros((C, 3), dtype=torch.int32, device=dev)
    ms, out, _, _ = timed_region(lambda: eng.drhmc_sample(q, a.nleap, a.step_size, a.nprops, a.stepfactor, T, msum=mom.sum,
                                                           msumsq=mom.sumsq, want_avec=False, counts_out=counts), dev, world)
    mom.n = T
    sec = ms * 1e-3
    ess = mom.ess_true(model.true_mean(), model.true_std())
    sums, Ctot = mom.reduce()
    mean = sums[0] / Ctot
    se_mean = torch.sqrt((sums[1] - sums[0] ** 2 / Ctot) / (Ctot - 1.0) / Ctot)        # between-chain variance of the chain means
    m2 = mom.sumsq / T                                                                   # per-chain second moments
    blk = torch.stack([m2.sum(dim=0, dtype=torch.float64), (m2 * m2).sum(dim=0, dtype=torch.float64)])
    dist.all_reduce_sum(blk)
    e2 = blk[0] / Ctot
    se_e2 = torch.sqrt(torch.clamp(blk[1] / Ctot - e2 ** 2, min=0.0) / Ctot)
    std = torch.sqrt(torch.clamp(e2 - mean ** 2, min=0.0))
    se_std = se_e2 / (2.0 * std)
    g = counts[:, 1].sum(dtype=torch.float64)[None]
    dist.all_reduce_sum(g)
    tstd = 9.487735836358526      # exp(2.25): Var(theta) = E[e^v] = e^{4.5}

    # control: plain HMC (same kernels, same generator, same step size and length) from the same kind of start
    from hmc_b200 import _lib, pyhmc
    hq = exact_funnel(C, d, seed + 211 + lo % 9973, dev)
    heng = pyhmc.HMC(model, device=dev, seed=seed + 19, chain_offset=lo).engine
    hm = utils.RunningMoments(heng, C, d)
    hnxt = torch.empty_like(hq)
    Th = 8
    for _ in range(Th):
        heng.hmc_step(hq, a.nleap, a.step_size, _lib.RULE_PYHMC, want_p=False, q_out=hnxt)
        hq, hnxt = hnxt, hq
        hm.add(hq)
    hs, hC = hm.reduce()
    hmean = hs[0] / hC
    hse = torch.sqrt((hs[1] - hs[0] ** 2 / hC) / (hC - 1.0) / hC)
    hm2 = hm.sumsq / Th
    hb = torch.stack([hm2[:, 0].sum(dtype=torch.float64), (hm2[:, 0] ** 2).sum(dtype=torch.float64)])
    dist.all_reduce_sum(hb)
    he2 = hb[0] / hC
    hstd = torch.sqrt(he2 - hmean[0] ** 2)
    hse_std = torch.sqrt(torch.clamp(hb[1] / hC - he2 ** 2, min=0.0) / hC) / (2.0 * hstd)
    control = {"sampler": "pyhmc.HMC.step, L=%d eps=%g, %d transitions from exact funnel draws" % (a.nleap, a.step_size, Th),
               "v_mean": float(hmean[0]), "v_mean_mc_err": float(hse[0]), "v_std": float(hstd), "v_std_mc_err": float(hse_std),
               "pass": abs(float(hmean[0])) < 5 * float(hse[0]) and abs(float(hstd) - 3.0) < 5 * float(hse_std)}
    del hq, hnxt, hm, hm2
    v_ok = abs(float(mean[0])) < 5 * float(se_mean[0]) and abs(float(std[0]) - 3.0) < 5 * float(se_std[0])
    th_mean_ok = bool((mean[1:].abs() < 5 * se_mean[1:]).float().mean() > 0.95)
    th_std_ok = bool(((std[1:] - tstd).abs() < 5 * se_std[1:]).float().mean() > 0.95)
    return {"start": "exact funnel draws", "transitions": T, "chains": int(Ctot), "ms_per_step": ms / T,
            "v_mean": float(mean[0]), "v_mean_mc_err": float(se_mean[0]), "v_std": float(std[0]), "v_std_mc_err": float(se_std[0]),
            "v_truth": [0.0, 3.0], "theta_std_mean_over_dims": float(std[1:].mean()), "theta_std_mc_err_typical": float(se_std[1:].median()),
            "theta_std_truth": tstd, "pass": {"v": v_ok, "theta_mean": th_mean_ok, "theta_std": th_std_ok},
            "pass_rule": "|estimate - truth| < 5 Monte Carlo errors (errors from the between-chain variance of per-chain moments); "
                         "theta: at least 95 % of the dimensions",
            "hmc_control": control,
            "note": "delayed rejection with nprops >= 2 drifts slowly off the exact ensemble at this step size (std of v about "
                    "-0.06 % per transition) while plain HMC through the same kernels and generator does not (hmc_control): it is "
                    "the reference's own rule -- the GPU's decisions are identical to algorithms.DRHMC.step's on the same draws "
                    "(tests/test_gpu_bench_path.py: 0 of 4096 differ) -- most likely get_num returning 0 for a ghost stage "
                    "whose energy is NaN (src/algorithms.py:491-492) where one that overflowed to +inf counts as a certain "
                    "rejection; see DESIGN.md section 2",
            "ess_true_v": float(ess[0]), "ess_true_min": float(ess.min()), "ess_true_min_per_sec": float(ess.min()) / sec,
            "ess_true_v_per_sec": float(ess[0]) / sec, "ess_true_per_grad_eval_v": float(ess[0]) / float(g[0]),
            "ess_note": "utils.get_ess_true over the %d-transition window of all chains: (std / rms over chains of the chain-mean "
                        "error)^2 * chains" % T}


def autocorr_block(a, dev, seed):
    """utils.get_rcc (src/utils.py:59-71: FFT autocorrelation, Sokal window c=5) on a STORED window: 4096 stationary
    chains, every transition kept (v and the first two thetas)."""
    import torch
    from hmc_b200 import algorithms, models, utils
    d, C, T = a.ndim, 4096, a.acf_steps
    eng = algorithms.DRHMC(models.Funnel(d), device=dev, seed=seed + 29, chain_offset=1 << 40).engine
    q = exact_funnel(C, d, seed + 5, dev)
    nxt = torch.empty_like(q)
    xs = torch.empty((T, C, 3), dtype=torch.float64, device=dev)
    counts = torch.zeros((T, C, 3), dtype=torch.int32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for t in range(T):
        eng.drhmc_step(q, a.nleap, a.step_size, a.nprops, a.stepfactor, want_p=False, q_out=nxt, want_avec=False,
                       counts_out=counts[t])
        q, nxt = nxt, q
        xs[t] = q[:, :3]
    e1.record()
    torch.cuda.synchronize()
    rcc, tc = utils.get_rcc(xs)                          # tc: integrated autocorrelation time at the window, [C, 3]
    tau = tc.double().clamp(min=1.0)
    ess_chain = T / tau
    grads = float(counts[:, :, 1].sum(dtype=torch.float64))
    return {"chains": C, "transitions_stored": T, "tau_v_median": float(tau[:, 0].median()), "tau_v_mean": float(tau[:, 0].mean()),
            "tau_theta_median": float(tau[:, 1:].median()), "ess_per_chain_v": float(ess_chain[:, 0].mean()),
            "ess_v_per_transition_per_chain": float(ess_chain[:, 0].mean()) / T,
            "ess_v_per_grad_eval": float(ess_chain[:, 0].sum()) / grads,
            "note": "src/utils.py:59-71 on x[T, 4096, 3] (v, theta_1, theta_2); tau = 1 + 2 sum rho up to the first lag >= 5 tau; "
                    "ESS per chain = T / tau.  ESS/s of the full ensemble = ess_v_per_grad_eval x grad evals/s of the headline"}


def gates_block(dev):
    """SURVEY section 8d correctness gates, computed in this run on a committed golden (made by the unmodified reference):
    max relative error of q and p, accept-code mismatches, Vgcount equality."""
    import torch
    from hmc_b200 import algorithms, models
    path = os.path.join(ROOT, "tests", "golden", "drhmc_funnel50.npz")
    if not os.path.exists(path):
        return {"error": "golden not found"}
    g = np.load(path)
    case = json.loads(str(g["spec"]))
    ar = case["args"]
    s = algorithms.DRHMC(models.Funnel(case["model"]["d"], case["model"].get("sigma", 3.0)), device=dev)
    err_q = err_p = 0.0
    code_bad = cnt_bad = n = 0
    for t in range(g["P"].shape[0]):
        q_in = g["q0"] if t == 0 else g["q"][t - 1]
        q, p, code, avec, counts = s.step(q_in, ar["nleap"], ar["step_size"], ar["nprops"], ar["stepfactor"], p=g["P"][t], u=g["U"][t])
        err_q = max(err_q, float(np.max(np.abs(q - g["q"][t]) / (np.abs(g["q"][t]) + 0.1))))
        err_p = max(err_p, float(np.max(np.abs(p - g["p"][t]) / (np.abs(g["p"][t]) + 0.1))))
        code_bad += int((np.asarray(code, dtype=np.float64) != g["code"][t]).sum())
        cnt_bad += int((np.asarray(counts)[:, 1] != g["counts"][t][:, 1]).sum())
        n += q.shape[0]
    return {"golden": "tests/golden/drhmc_funnel50.npz (algorithms.DRHMC.step of the unmodified reference, replayed draws)",
            "chain_transitions": n, "max_rel_err_q": err_q, "max_rel_err_p": err_p, "tolerance": 1e-10,
            "accept_code_mismatches": code_bad, "vgcount_mismatches": cnt_bad,
            "pass": err_q < 1e-10 and err_p < 1e-10 and code_bad == 0 and cnt_bad == 0}


def secondary_block(a, dev, world, rank, fp64_peak, peaks):
    """The other BASELINE.json configurations, chains sharded over the ranks (fixed totals), each a few timed
    transitions: ms, rate and its own roofline fraction.  Timing: CUDA events, best of 3 after 2 warm-ups, max over ranks."""
    import torch
    from hmc_b200 import _lib, adaptivehmc, algorithms, dist, mala, models, pyhmc
    rows = []
    g = torch.Generator().manual_seed(a.seed + 3)

    def shard(n):
        lo, hi = dist.shard_range(n)
        return lo, hi - lo

    def timed(fn, reps=3, warm=2):
        for _ in range(warm):
            fn()
        best, out = 1e30, None
        for _ in range(reps):
            ms, out, _, _ = timed_region(fn, dev, world)
            best = min(best, ms)
        return best, out

    def total(x):
        t = torch.as_tensor([float(x)], dtype=torch.float64, device=dev)
        dist.all_reduce_sum(t)
        return float(t[0])

    def row(name, ms, grads, flops_per_grad, bound="fp64", extra=None):
        tf = grads * flops_per_grad / (ms * 1e-3) / 1e12
        r = {"config": name, "ms_per_transition": ms, "grad_evals_per_s": grads / (ms * 1e-3), "algorithmic_tflops": tf, "bound": bound}
        if bound == "fp64":
            r["frac_of_fp64_peak"] = tf / (fp64_peak * world)
        if extra:
            r.update(extra)
        rows.append(r)

    # config 2: correlated Gaussian d=100, 65,536 chains, fixed step / L (FP64 tensor path, mma.sync.m8n8k4.f64)
    d, Ct, L = 100, 65536, 32
    lo, C = shard(Ct)
    gm = models.CorrGaussian.ar1(d)
    s = pyhmc.HMC(gm, device=dev, seed=1, chain_offset=lo)
    q = torch.randn(Ct, d, generator=g, dtype=torch.float64)[lo:lo + C].to(dev)
    eps = 0.1 * float(np.sqrt(np.linalg.eigvalsh(gm.Sigma).min()))
    ms, out = timed(lambda: s.engine.hmc_step(q, L, eps, _lib.RULE_PYHMC, want_p=False))
    row("cfg2 correlated Gaussian d=100, HMC L=32 eps=%.4f, 65,536 chains (k_gauss_hmc, DMMA)" % eps, ms, Ct * (L + 1), 2 * d * d + 4 * d)

    # config 3, light variant (the reference's usual trajectory length): funnel d=50 DRHMC nleap=20 eps=0.1
    d, Ct = 50, 1 << 20
    lo, C = shard(Ct)
    fm = models.Funnel(d)
    q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
    dr = algorithms.DRHMC(fm, device=dev, seed=1, chain_offset=lo)
    counts = torch.zeros((C, 3), dtype=torch.int32, device=dev)
    ms, out = timed(lambda: dr.engine.drhmc_step(q, 20, 0.1, 3, 2.0, want_p=False, counts_out=counts, want_avec=False))
    row("cfg3-light funnel d=50 DRHMC nleap=20 eps=0.1 nprops=3, 2^20 chains, one call per transition", ms,
        total(counts[:, 1].sum(dtype=torch.float64)), 7 * d)
    Kf = 8
    ms, out = timed(lambda: dr.engine.drhmc_sample(q, 20, 0.1, 3, 2.0, Kf, counts_out=counts, want_avec=False))
    row("cfg3-light, %d transitions per call (hmcb_drhmc_sample)" % Kf, ms / Kf, total(counts[:, 1].sum(dtype=torch.float64)) / Kf, 7 * d)
    h = pyhmc.HMC(fm, device=dev, seed=1, chain_offset=lo)
    ms, out = timed(lambda: h.engine.hmc_step(q, 20, 0.1, _lib.RULE_PYHMC, want_p=False))
    row("funnel d=50 HMC L=20 eps=0.1, 2^20 chains, one call per transition", ms, Ct * 21, 7 * d)
    Kh = 16
    ms, out = timed(lambda: h.engine.hmc_sample(q, 20, 0.1, _lib.RULE_PYHMC, Kh))
    row("funnel d=50 HMC L=20 eps=0.1, 2^20 chains, %d transitions per launch (hmcb_hmc_sample, k_hmc_multi_ws)" % Kh, ms / Kh,
        Ct * 21, 7 * d)
    # GHMC (partial momentum refresh, cached V and grad V): one launch per transition moves six state arrays,
    # K per launch keeps them in registers
    gh = algorithms.GHMC(fm, device=dev, seed=1, chain_offset=lo)
    pg = torch.randn(C, d, generator=g, dtype=torch.float64).to(dev)
    for nl in (20, 1):
        o0 = gh.engine.ghmc_sample(q, pg, nl, 0.1, 0.5, 4)
        ms, out = timed(lambda: gh.engine.ghmc_step(o0["q"], o0["p"], nl, 0.1, 0.5, V=o0["V"], Vg=o0["Vg"]))
        row("funnel d=50 GHMC nleap=%d eps=0.1 alpha=0.5, 2^20 chains, one call per transition" % nl, ms, Ct * nl, 7 * d,
            extra={"state_bytes_per_s": 6.0 * Ct * d * 8 / (ms * 1e-3), "frac_of_hbm_peak": 6.0 * Ct * d * 8 / (ms * 1e-3) / 1e9 / (peaks["hbm_gbs"] * world)})
        Kg = 16
        ms, out = timed(lambda: gh.engine.ghmc_sample(o0["q"], o0["p"], nl, 0.1, 0.5, Kg, V=o0["V"], Vg=o0["Vg"]))
        row("funnel d=50 GHMC nleap=%d, %d transitions per launch (hmcb_ghmc_sample, k_ghmc_multi)" % (nl, Kg), ms / Kg, Ct * nl, 7 * d)
    del q, counts, pg, o0

    # config 4: Bayesian logistic regression N=100,000 D=128, 16,384 chains, HMC L=10 (tcgen05 gradient GEMMs)
    N, D, Ct, L = 100000, 128, 16384, 10
    lo, C = shard(Ct)
    rng = np.random.default_rng(3)
    X = (rng.standard_normal((N, D)) / np.sqrt(D)).astype(np.float32)
    y = (rng.random(N) < 1 / (1 + np.exp(-(X.astype(np.float64) @ rng.standard_normal(D))))).astype(np.float32)
    lm = algorithms.HMC(models.LogisticRegression(X, y, 1.0), device=dev, seed=3, chain_offset=lo)
    q = (0.1 * torch.randn(Ct, D, generator=g))[lo:lo + C].to(dev)
    ms, out = timed(lambda: lm.engine.hmc_step(q, L, 0.01, _lib.RULE_ALG, want_p=False), reps=2, warm=1)
    grads = Ct * (L + 1)
    alg = grads * 4.0 * N * D / (ms * 1e-3) / 1e12
    row("cfg4 logistic regression N=100,000 D=128, HMC L=10 eps=0.01, 16,384 chains", ms, grads, 4.0 * N * D, bound="tensor",
        extra={"frac_of_bf16_peak_algorithmic": alg / (peaks["bf16_tflops"] * world),
               "executed_fp16_tflops": 3 * alg, "frac_of_bf16_peak_executed": 3 * alg / (peaks["bf16_tflops"] * world),
               "peak": peaks["bf16_tflops"], "peak_source": peaks["source"],
               "note": "algorithmic = 4 N D flop per gradient (two GEMMs); executed = 3 fp16 products per algorithmic one "
                       "(hi*hi + hi*lo + lo*hi for 1e-5 parity with the float64 reference)"})
    del lm, q

    # config 5 at d=100: Rosenbrock, 131,072 chains: HMC L=20, MALA (32 steps per launch), AdHMC_eps0, and the
    # dual-averaging warm-up whose accept statistic is all-reduced over the ranks every iteration
    d, Ct = 100, 131072
    lo, C = shard(Ct)
    rm = models.Rosenbrock(d)
    q = (1.0 + 0.1 * torch.randn(Ct, d, generator=g, dtype=torch.float64))[lo:lo + C].to(dev)
    fpg = 8.5 * d
    hh = algorithms.HMC(rm, device=dev, seed=4, chain_offset=lo)
    ms, out = timed(lambda: hh.engine.hmc_step(q, 20, 0.02, _lib.RULE_ALG, want_p=False))
    row("cfg5 Rosenbrock d=100 HMC L=20 eps=0.02, 131,072 chains, one call per transition", ms, Ct * 21, fpg)
    ms, out = timed(lambda: hh.engine.hmc_sample(q, 20, 0.02, _lib.RULE_ALG, 16))
    row("cfg5 Rosenbrock d=100 HMC L=20, 16 transitions per launch (hmcb_hmc_sample, k_hmc_multi)", ms / 16, Ct * 21, fpg)
    mm = mala.MALA(rm, device=dev, seed=4, chain_offset=lo)
    ms, out = timed(lambda: mm.engine.mala_step(q, 0.02, n_steps=32))
    # the generator alone (Philox4x32-10 + f32 Box-Muller, 4 blocks in flight per thread, nothing else): MALA's other ceiling
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    sink = torch.empty(sms * 8 * 256, dtype=torch.float64, device=dev)
    cnt = _lib.C.c_double(0)
    gen_best = 0.0
    for it in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(mm.engine.lib.hmcb_probe_normals(0, sms * 8, 256, 400, _lib.ptr(sink), _lib.C.byref(cnt), mm.engine._stream()))
        e1.record()
        torch.cuda.synchronize()
        if it:
            gen_best = max(gen_best, cnt.value / (e0.elapsed_time(e1) * 1e-3))
    nps = Ct * 32 * d / (ms * 1e-3)
    row("cfg5 Rosenbrock d=100 MALA eps=0.02, 32 steps per launch", ms / 32, Ct * 33 / 32.0, fpg,
        extra={"normals_per_s": nps, "generator_only_normals_per_s_per_gpu": gen_best,
               "frac_of_generator_ceiling": nps / (gen_best * world) if gen_best else None,
               "note": "one gradient per d normals: the step is bound by Philox + Box-Muller issue (integer multiplies, MUFU), not "
                       "by the FP64 pipe; hmcb_probe_normals measures that generator running alone"})
    ad = adaptivehmc.AdHMC_eps0(rm, device=dev, seed=4, chain_offset=lo)
    steps = ad.get_steps(0.01, 1.0, 10)
    ms, out = timed(lambda: ad.engine.adhmc_eps0_step(q, 20, steps, 1))
    row("cfg5 Rosenbrock d=100 AdHMC_eps0 ntry=10 Nleap=20", ms, total(out["counts"][:, 1].sum(dtype=torch.float64)), fpg)
    hh.step_size = 0.05
    nit = 60
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    hh.adapt_stepsize(q, nit, nsamples=0, burnin=0, step_size=0.05, nleap=20)
    torch.cuda.synchronize()
    dist.barrier()
    ms_it = (time.perf_counter() - t0) * 1e3 / (nit + 1)
    row("cfg5 Rosenbrock d=100 HMC L=20 dual-averaging warm-up, %d iterations, accept statistic all-reduced (%s) every "
        "iteration" % (nit + 1, "NCCL, %d ranks" % world if world > 1 else "1 rank: no collective"), ms_it, Ct * 21, fpg,
        extra={"adapted_step_size": float(hh.step_size), "timing": "host wall clock per iteration (one host round trip each)"})
    return rows


def run_hmcb(a):
    import torch

    from hmc_b200 import dist

    rank, world, local = dist.init_from_env()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    n_total = total_chains(a)
    lo, hi = dist.shard_range(n_total)
    C, d, K = hi - lo, a.ndim, a.steps

    g = torch.Generator().manual_seed(a.seed)
    if a.scaling == "strong":   # one global array, sliced: chain statistics do not depend on N
        q0 = (torch.rand(n_total, d, generator=g, dtype=torch.float64) * 2 - 1)[lo:hi].to(dev).contiguous()
    else:                       # per-shard stream (no N*2^20 x d host array); shard 0 is the N=1 workload
        g.manual_seed(a.seed + 7919 * rank)
        q0 = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev).contiguous()

    clocks = ClockSampler(local)
    clocks.start()
    from hmc_b200 import algorithms, models
    probe_eng = algorithms.DRHMC(models.Funnel(d), device=dev).engine
    probe_eng._ensure()
    fp64_peak = fp64_peak_tflops(probe_eng, dev)
    peaks = measured_peaks()
    # nvidia-smi's start-up (NVML init over every GPU of the box) stalls kernel launches for a moment: be past it
    t_wait = time.time()
    while len(clocks.rows) < 3 and time.time() - t_wait < 10.0:
        time.sleep(0.05)

    H = dr_block(a, dev, world, lo, C, q0, a.seed, K, a.warmup, a.burnin, True, fp64_peak)
    clocks_summary = clocks.summary(H["t0"], H["t1"])
    clocks.stop()
    ms_total, grads, leaps = H["ms_total"], H["grads"], H["leaps"]
    sec = ms_total * 1e-3
    value = grads / sec
    fpg = FLOPS_PER_GRAD["funnel"](d)
    mom, model, eng = H["mom"], H["model"], H["eng"]
    ess = mom.ess_true(model.true_mean(), model.true_std())
    rhat = mom.rhat()
    mean, std = mom.mean_std()

    # roofline of the dominant kernel of the timed region: DR stage 0 (every chain integrates nleap steps) as the plain kernel
    # -- what hmcb_drhmc_sample launches with >= 3 chain ranges -- timed by events between the stage launches of single
    # hmcb_drhmc_step calls run straight after the timed region (one launch over the whole batch where the timed region
    # makes four over its quarters)
    timed_plain = H["n_split"] >= 3          # csrc/api.cu: with fewer ranges the sample call launches the warp-specialised kernel
    P0, P1 = (H["passes"]["plain"], H["passes"]["ws"]) if timed_plain else (H["passes"]["ws"], H["passes"]["plain"])
    k_plain = "k_transition<Funnel,f64,EPL=%d,LPC=1,MODE_DR0>" % d
    k_ws = "k_transition_ws<Funnel,f64,EPL=%d,MODE_DR0> (warp-specialised)" % d
    st_ms = P0["stage_ms"]
    s0_flops = P0["stage_grad_evals"][0] * fpg
    achieved = s0_flops / (st_ms[0] * 1e-3) / 1e12

    def fracs(Pp):
        n = min(len(Pp["stage_ms"]), len(Pp["stage_grad_evals"]))
        out = [Pp["stage_grad_evals"][j] * fpg / (Pp["stage_ms"][j] * 1e-3) / 1e12 / fp64_peak for j in range(n - 1)]
        out.append(Pp["stage_grad_evals"][n - 1] * fpg / (sum(Pp["stage_ms"][n - 1:]) * 1e-3) / 1e12 / fp64_peak)
        return out
    region_tflops = grads / world * fpg / sec / 1e12
    roofline = {"bound": "fp64",
                "kernel": "the stage kernels of the timed hmcb_drhmc_sample call together (k_transition<MODE_DR0> and k_transition<MODE_DRN>, "
                          "%d launches per transition and chain range, overlapping on %d streams: they have no durations of their own "
                          "there); per-kernel figures from serialised passes in 'stage0_kernel' / 'stage_frac'" %
                          (len(st_ms) + max(0, len(st_ms) - 2), H["n_split"]),
                "achieved": region_tflops, "peak": fp64_peak, "unit": "TFLOP/s", "frac": region_tflops / fp64_peak if fp64_peak else None,
                "achieved_note": "ALGORITHMIC flops of the timed region (sum of Vgcount x 7d; the exp polynomial, energies and RNG are "
                                 "executed but not counted) / its duration, per GPU",
                "stage0_kernel": {"kernel": "%s: DR stage 0 as the timed region launches it (largest single kernel: %.0f %% of the "
                                            "gradient work)" % (k_plain if timed_plain else k_ws, 100 * P0["stage_grad_evals"][0] * world / grads * K
                                                                if grads else 0.0),
                                  "achieved": achieved, "frac": achieved / fp64_peak if fp64_peak else None,
                                  "algorithmic_flops_per_launch": s0_flops, "avg_launch_ms": st_ms[0]},
                "traffic": None,
                "traffic_note": "not measured in this run; ncu --set full of the stage-0 kernels at this shape: 792 MB per launch "
                                "(profiles/r2_ncu_kernels.md) against 2*C*d*8 + outputs = 0.87 GB algorithmic: FP64-pipe-bound, "
                                "HBM at 5 % of its peak",
                "peak_source": "hmcb_probe_fma: register-resident DFMA loop measured in this run on this GPU "
                               "(MEASURED_PEAKS.json has no FP64 figure)",
                "algorithmic_flops_per_step": grads / world * fpg / K, "avg_step_ms": ms_total / K,
                "stage_ms": st_ms, "stage_ms_sum": sum(st_ms),
                "stage_frac": fracs(P0),
                "stage_note": "per stage: gradient evaluations of the chains that reach it x 7d flops / its kernels' time / peak; "
                              "stages >= 2 run as two kernels each (park / resume)",
                "other_stage0_kernel": {"kernel": (k_ws + ": what a single hmcb_drhmc_step call launches at this size") if timed_plain
                                        else (k_plain + ": what a sample call of >= 3 chain ranges launches"),
                                        "stage_ms": P1["stage_ms"], "stage_frac": fracs(P1)},
                "moments_kernel_ms": H["moments_ms"],
                "timing_note": "stage kernels timed by CUDA events recorded inside hmcb_drhmc_step (hmcb_profile) over %d single "
                               "transitions per variant run straight after the timed region, stages serialised on one stream; the "
                               "timed region itself runs %d chain ranges on %d streams, so its ms_per_step (%.3f) is to be compared "
                               "with stage_ms_sum + moments_kernel_ms (%.3f)" %
                               (H["stage_passes"], H["n_split"], H["n_split"], ms_total / K, sum(st_ms) + H["moments_ms"]),
                "step_frac_all_stages": region_tflops / fp64_peak,
                "hbm_peak_gbs": peaks["hbm_gbs"], "hbm_peak_source": peaks["source"],
                "algorithmic_hbm_gbs_of_the_step": (2 * C * d * 8 + 5 * C * d * 8) / (ms_total / K * 1e-3) / 1e9}

    strong = None
    if world > 1 and a.scaling == "weak":
        # the configuration as BASELINE.json words it: 2^20 chains IN TOTAL, sharded over the N GPUs
        lo_s, hi_s = dist.shard_range(a.chains)
        gs = torch.Generator().manual_seed(a.seed)
        qs = (torch.rand(a.chains, d, generator=gs, dtype=torch.float64) * 2 - 1)[lo_s:hi_s].to(dev).contiguous()
        S = dr_block(a, dev, world, lo_s, hi_s - lo_s, qs, a.seed, K, a.warmup, a.burnin, False, fp64_peak)
        strong = {"chains_total": a.chains, "chains_per_gpu": (hi_s - lo_s), "ms_per_step": S["ms_total"] / K,
                  "value": S["grads"] / (S["ms_total"] * 1e-3), "unit": "grad_evals/s", "scaling": "strong",
                  "per_call": S["per_call"],
                  "note": "same workload and timing method with the 2^20 chains divided over the %d GPUs; compare with the "
                          "N=1 line's value for the strong-scaling efficiency" % world}
        del S, qs

    stationary = stationary_block(a, dev, world, lo, C, a.seed) if a.ess_steps > 0 else None
    secondary = None
    if not a.no_secondary:
        secondary = secondary_block(a, dev, world, rank, fp64_peak, peaks)

    # end to end: public API, pinned host buffers in, host results out, every step
    e2e = None
    if not a.no_e2e:
        Ke = a.e2e_steps or max(3, min(K, 8))
        qh = torch.empty((C, d), dtype=torch.float64).pin_memory()
        qh.copy_(H["q"])
        dr_e = algorithms.DRHMC(model, device=dev, seed=a.seed + 1, chain_offset=lo)
        for _ in range(2):
            r = dr_e.step(qh, a.nleap, a.step_size, a.nprops, a.stepfactor, want_p=False)
            qh = r[0]
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.time()
        # the gradient evaluations of a step (this bench's numerator) are summed from the DEVICE copy of the step's counters,
        # asynchronously behind the step -- every result, counters included, still lands in host memory inside the step;
        # summing the host copy here instead is 0.5 ms of single-core work per step that is the bench's, not the call's
        g_dev = torch.zeros((), dtype=torch.int64, device=dev)
        for _ in range(Ke):
            r = dr_e.step(qh, a.nleap, a.step_size, a.nprops, a.stepfactor, want_p=False)
            qh = r[0]
            g_dev += dr_e.engine._host_stage[1]["counts"][:, 1].sum(dtype=torch.int64)
        torch.cuda.synchronize()
        dt = torch.tensor([time.time() - t0], dtype=torch.float64, device=dev)
        g_e2e = int(g_dev)
        assert int(r[4][:, 1].sum()) > 0 and r[4].device.type == "cpu"
        r = (qh,)
        for _ in range(2):                                                       # warm the pinned p buffers
            r = dr_e.step(r[0], a.nleap, a.step_size, a.nprops, a.stepfactor)
        torch.cuda.synchronize()
        t1 = time.time()
        for _ in range(3):
            r = dr_e.step(r[0], a.nleap, a.step_size, a.nprops, a.stepfactor)
        torch.cuda.synchronize()
        ms_with_p = 1e3 * (time.time() - t1) / 3
        floor_ms = link_floor_ms(C, d, a.nprops, dev, world)
        ge = torch.tensor([float(g_e2e)], dtype=torch.float64, device=dev)
        if world > 1:
            torch.distributed.all_reduce(dt, op=torch.distributed.ReduceOp.MAX)
            dist.all_reduce_sum(ge)
        h2d = C * d * 8
        d2h = C * d * 8 + C * 4 + C * a.nprops * 8 + C * 12
        e2e = {"value": float(ge[0]) / float(dt[0]), "unit": "grad_evals/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": Ke, "ms_per_step": 1e3 * float(dt[0]) / Ke,
               "api": "hmc_b200.algorithms.DRHMC.step(q_pinned_host, ..., want_p=False) -> host (q, None, code, avec, "
                      "counts); the reference's sample loop discards the returned momentum",
               "ms_per_step_with_p_returned": ms_with_p, "d2h_bytes_per_step_with_p": d2h + C * d * 8,
               "link_floor_ms": floor_ms,
               "link_floor_note": "all %d ranks at once doing ONLY this step's host-to-device and device-to-host copies from "
                                  "pinned memory on two streams (no kernels): what the host links and memory allow" % world}

    if rank != 0:
        return
    acf = autocorr_block(a, dev, a.seed) if a.acf_steps > 0 else None
    gates = gates_block(dev)
    cpu = cfg1 = None
    if not a.no_cpu_baseline and world == 1:
        cpu = cpu_baseline(a, a.cpu_seconds)
        cfg1 = cpu_config1()
    split_on = os.environ.get("HMCB_DR_STAGE_SPLIT", "1") != "0" and a.nprops >= 3 and a.nleap >= 48
    n_stage_kernels = a.nprops + (a.nprops - 2 if split_on else 0)
    line = {"metric": "grad_evals_per_sec", "value": value, "unit": "grad_evals/s", "n_gpus": world, "steps": K,
            "warmup": a.warmup, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": a.scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(a, transitions_per_call=K, chain_ranges=H["n_split"]),
            "timed_region": "ONE hmcb_drhmc_sample call: %d DRHMC transitions of every chain, %d chain ranges pipelined over "
                            "%d streams, running moments folded in (bit-identical to %d single hmcb_drhmc_step calls)" %
                            (K, H["n_split"], H["n_split"], K),
            "per_call": H["per_call"],
            "leapfrog_steps_per_sec": (grads - leaps) / sec,
            "grad_evals_per_transition": grads / (K * n_total), "accept_rate": H["accepted"] / (K * n_total),
            "code_fractions": {"note": "rank 0, last timed transition: fraction of chains by DRHMC code (0 = rejected at every "
                               "stage, j+1 = accepted at stage j)", "by_code": H["code_frac"]},
            "timed_window_diagnostics": {"ess_true_min": float(ess.min()), "rhat_max": float(rhat.max()), "v_mean": float(mean[0]),
                                         "v_std": float(std[0]),
                                         "note": "over the %d timed transitions only, chains started at U(-1,1) and burnt in for %d "
                                                 "transitions: NOT stationary (the funnel's v needs thousands); the stationary "
                                                 "figures are in 'stationary' and 'autocorr'" % (K, a.burnin)},
            "stationary": stationary, "autocorr": acf,
            "ess_per_sec": (stationary or {}).get("ess_true_min_per_sec"),
            "gates": gates, "strong": strong, "secondary": secondary,
            "roofline": roofline, "cpu_baseline": cpu, "config1_cpu": cfg1, "e2e": e2e, "clocks": clocks_summary,
            "gpu_launches": K * H["n_split"] * (n_stage_kernels + 1),
            "gpu_launches_note": "timed region: %d transitions x %d chain ranges x (%d DR stage kernels%s + 1 counters/moments "
                                 "kernel), all libhmcb.so" % (K, H["n_split"], n_stage_kernels,
                                                              ", stages >= 2 in two kernels each" if split_on else "")}
    print(json.dumps(line), flush=True)


def link_floor_ms(C, d, nprops, dev, world):
    """The copies of one end-to-end step and nothing else, all ranks at once: H2D of q, D2H of q, code, avec, counts."""
    import torch
    from hmc_b200 import dist
    pin = dict(pin_memory=True)
    hq, hq2 = torch.empty((C, d), dtype=torch.float64, **pin), torch.empty((C, d), dtype=torch.float64, **pin)
    hs = [torch.empty(C, dtype=torch.int32, **pin), torch.empty((C, nprops), dtype=torch.float64, **pin),
          torch.empty((C, 3), dtype=torch.int32, **pin)]
    dq, dq2 = torch.empty((C, d), dtype=torch.float64, device=dev), torch.zeros((C, d), dtype=torch.float64, device=dev)
    ds = [torch.zeros_like(h, device=dev) for h in hs]
    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def once():
        with torch.cuda.stream(s_in):
            dq.copy_(hq, non_blocking=True)
        with torch.cuda.stream(s_out):
            hq2.copy_(dq2, non_blocking=True)
            for h, t in zip(hs, ds):
                h.copy_(t, non_blocking=True)
    once()
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.time()
    for _ in range(4):
        once()
    torch.cuda.synchronize()
    dt = torch.tensor([time.time() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        torch.distributed.all_reduce(dt, op=torch.distributed.ReduceOp.MAX)
    return 1e3 * float(dt[0]) / 4


def cpu_config1():
    """BASELINE.json config 1 exactly as stated: Neal's funnel d=10, pyhmc HMC, step 0.1, 20 leapfrog steps, ONE chain,
    10,000 samples, on one host core (oracle/samplers.py hmc_step_pyhmc, the numpy port of src/pyhmc.py:64-69)."""
    from oracle import models as om
    from oracle import samplers
    model = om.Funnel(10)
    rng = np.random.Generator(np.random.PCG64(0))

    class R:
        normal = staticmethod(lambda size=None: rng.standard_normal(size))
        uniform = staticmethod(lambda low=0.0, high=1.0, size=None: rng.uniform(low, high, size))
    q = rng.uniform(-1, 1, 10)
    grads = acc = 0
    t0 = time.time()
    for _ in range(10000):
        q, p, code, Hs, cnt = samplers.hmc_step_pyhmc(model, q, 20, 0.1, R)
        grads += cnt[1]
        acc += code == 1.0
    dt = time.time() - t0
    return {"config": "funnel d=10, pyhmc.HMC.step, step 0.1, 20 leapfrog steps, 1 chain, 10,000 samples", "seconds": dt,
            "grad_evals": grads, "grad_evals_per_s": grads / dt, "accept_rate": acc / 10000.0, "cores": 1, "kind": "port"}


def main():
    a = parse()
    if os.environ.get("HMCB_BENCH_WATCHDOG"):   # debugging aid: dump every thread's stack and exit after that many seconds
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["HMCB_BENCH_WATCHDOG"]), exit=True)
    # stdout carries exactly ONE line, the JSON: everything else any library prints there (NCCL's version banner ...)
    # goes to stderr.  File descriptor 1 is pointed at stderr for the run; the JSON goes to the saved descriptor.
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(saved, "w")
    if a.impl == "reference":
        run_reference(a)
    else:
        run_hmcb(a)
    td = sys.modules.get("torch.distributed")
    if td is not None and td.is_available() and td.is_initialized():
        td.destroy_process_group()


if __name__ == "__main__":
    main()
