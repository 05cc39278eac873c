#!/usr/bin/env python
"""Benchmark of the sampler hot path on the BASELINE.json headline workload:
Neal's funnel d=50, delayed-rejection HMC (nleap=150, step 0.2 -> lpath 30 as
scripts_old/jobscripts/run_funnel2.sh:21-22; nprops=3, stepfactor=2), 2^20 chains.
Chains are independent, so N GPUs shard them by global chain id with no data-path
collective.  Default: weak scaling -- every GPU runs the N=1 workload (2^20 chains per
GPU, N*2^20 in total); `--scaling strong` keeps 2^20 chains in total, as the BASELINE
config words it.

A "step" is one DRHMC transition of every chain.  Prints ONE JSON line.
  value  : gradient evaluations/s (the reference's Vgcount counter, src/pyhmc.py:21-23),
           chain state resident in HBM
  e2e    : same metric through the public API with HOST (pinned) buffers, copies timed
  roofline / cpu_baseline / clocks: see DESIGN.md "Measurement"
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
def run_hmcb(a):
    import torch

    from hmc_b200 import _lib, algorithms, dist, models, utils

    rank, world, local = dist.init_from_env()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    n_total = total_chains(a)
    lo, hi = dist.shard_range(n_total)
    C, d = hi - lo, a.ndim

    model = models.Funnel(d)
    dr = algorithms.DRHMC(model, device=dev, seed=a.seed, chain_offset=lo)
    eng = dr.engine
    eng._ensure()
    g = torch.Generator().manual_seed(a.seed)
    if a.scaling == "strong":   # one global array, sliced: chain statistics do not depend on N
        q = (torch.rand(n_total, d, generator=g, dtype=torch.float64) * 2 - 1)[lo:hi].to(dev).contiguous()
    else:                       # per-shard stream (no N*2^20 x d host array); shard 0 is the N=1 workload
        g.manual_seed(a.seed + 7919 * rank)
        q = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev).contiguous()
    nxt = torch.empty_like(q)

    K = a.steps
    counts_log = torch.zeros((K + 1, C, 3), dtype=torch.int32, device=dev)   # slot K: scratch for untimed steps
    code_log = torch.zeros((K + 1, C), dtype=torch.int32, device=dev)

    def transition(qi, qo, events=None, slot=K):
        # every output buffer is preallocated: no allocator traffic inside the timed loop
        return eng.drhmc_step(qi, a.nleap, a.step_size, a.nprops, a.stepfactor, want_p=False, q_out=qo,
                              stage_events=events, want_avec=False, code_out=code_log[slot],
                              counts_out=counts_log[slot])

    clocks = ClockSampler(local)
    clocks.start()
    for _ in range(a.burnin):
        transition(q, nxt)
        q, nxt = nxt, q
    mom = utils.RunningMoments(eng, C, d)

    # FP64 pipe peak of this GPU, measured now (MEASURED_PEAKS.json carries only HBM and bf16 numbers)
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
    fp64_peak_tflops = best

    for _ in range(a.warmup):
        out = transition(q, nxt)
        q, nxt = nxt, q
        mom.add(q)
    mom.reset()

    # nvidia-smi was started before burn-in; its start-up (NVML init over every GPU of the box) stalls
    # kernel launches for a moment, so it must be past that before the timed region begins
    t_wait = time.time()
    while len(clocks.rows) < 3 and time.time() - t_wait < 10.0:
        time.sleep(0.05)
    KR = min(K, 6)      # transitions of the per-stage timing pass after the timed region
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(a.nprops + 1)] for _ in range(KR)]
    for row in evs:
        for ev in row:
            ev.record()      # creates the cudaEvent_t handles now, not inside the timed loop
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    t_wall0 = time.time()
    start.record()
    for k in range(K):
        transition(q, nxt, None, slot=k)
        q, nxt = nxt, q
        mom.add(q)
    end.record()
    torch.cuda.synchronize()
    dist.barrier()
    t_wall1 = time.time()
    ms = torch.tensor([start.elapsed_time(end)], dtype=torch.float64, device=dev)
    if world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    ms_total = float(ms[0])
    clocks.stop()

    # Per-stage kernel times for the roofline: the timed region above runs the DR step as the library runs it
    # by default -- chain ranges on several streams, stages of different ranges overlapping -- so a stage has no
    # duration of its own there.  Straight after it, the same transition is run KR more times from the same state
    # with hmcb_profile events, which makes the library run the stages one after another on one stream.
    n_split = 2 if C >= (1 << 15) else 1     # dr_split_for() in csrc/api.cu
    if os.environ.get("HMCB_DR_SPLIT"):
        n_split = int(os.environ["HMCB_DR_SPLIT"])
    for k in range(KR):
        transition(q, nxt, evs[k], slot=K)
        q, nxt = nxt, q
    torch.cuda.synchronize()
    tot = counts_log[:K].sum(dim=(0, 1), dtype=torch.float64)
    stage_ms = torch.tensor([[evs[k][j].elapsed_time(evs[k][j + 1]) for j in range(a.nprops)] for k in range(KR)],
                            dtype=torch.float64, device=dev)
    n_stage1 = (code_log[:K] != 1).sum(dim=1).double().mean()
    acc_rate = (code_log[:K] > 0).double().mean()
    code_frac = (torch.bincount(code_log[:K].flatten().long(), minlength=a.nprops + 1).double() / (K * C)).tolist()
    stats = torch.cat([tot, stage_ms.mean(dim=0), n_stage1[None], acc_rate[None] * C])
    dist.all_reduce_sum(stats[:3])
    dist.all_reduce_sum(stats[-1:])
    grads, leaps = float(stats[1]), float(stats[2])
    sec = ms_total * 1e-3
    value = grads / sec
    fpg = FLOPS_PER_GRAD["funnel"](d)

    # ESS over the timed transitions from streamed per-chain means (utils.get_ess_true), all ranks
    ess = mom.ess_true(model.true_mean(), model.true_std())
    rhat = mom.rhat()
    mean, std = mom.mean_std()
    ess_min = float(ess.min())

    # roofline of the dominant kernel: DR stage 0 (every chain integrates nleap steps)
    L0 = max(a.nleap, 1)
    st_ms = stats[3:3 + a.nprops].tolist()
    s0_flops = C * (L0 + 1) * fpg
    achieved = s0_flops / (st_ms[0] * 1e-3) / 1e12
    default_shape = (C == 1 << 20 and d == 50 and a.nleap == 150)
    roofline = {"bound": "fp64", "kernel": "k_transition_ws<Funnel,f64,EPL=50,MODE_DR0> (DR stage 0, warp-specialised)",
                "achieved": achieved, "peak": fp64_peak_tflops, "unit": "TFLOP/s",
                "frac": achieved / fp64_peak_tflops if fp64_peak_tflops else None,
                # dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full (profiles/r1b_ncu_dr_stages.md);
                # algorithmic HBM bytes are 2*C*d*8 + outputs = 0.87 GB, i.e. no re-reads
                "traffic": 447266560 + 347079680 if default_shape else None,
                "traffic_unit": "bytes per launch (HBM; the kernel is FP64-pipe-bound, HBM is at 5 % of its peak)",
                "peak_source": "hmcb_probe_fma: register-resident DFMA loop measured in this run on this GPU "
                               "(MEASURED_PEAKS.json has no FP64 figure)",
                "algorithmic_flops_per_launch": s0_flops, "avg_launch_ms": st_ms[0],
                "stage_ms": st_ms, "stage_ms_sum": sum(st_ms),
                "timing_note": "stage kernels timed by CUDA events recorded inside hmcb_drhmc_step (hmcb_profile) over %d "
                               "transitions run straight after the timed region with the stages serialised on one stream; "
                               "inside the timed region the library overlaps the stages of 2 chain ranges on 2 streams "
                               "(ms_per_step is below the sum of the stage times for that reason)" % KR, "step_frac_all_stages": grads / world * fpg / sec / 1e12 / fp64_peak_tflops}

    # end to end: public API, pinned host buffers in, host results out, every step
    e2e = None
    if not a.no_e2e:
        Ke = a.e2e_steps or max(3, min(K, 8))
        qh = torch.empty((C, d), dtype=torch.float64).pin_memory()
        qh.copy_(q)
        dr_e = algorithms.DRHMC(model, device=dev, seed=a.seed + 1, chain_offset=lo)
        # the momentum is not brought back: the reference's sample loop discards the p its step returns
        # (src/algorithms.py:546-551); the same call with p is timed after it and reported beside it
        for _ in range(2):
            r = dr_e.step(qh, a.nleap, a.step_size, a.nprops, a.stepfactor, want_p=False)
            qh = r[0]
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.time()
        g_e2e = 0
        for _ in range(Ke):
            r = dr_e.step(qh, a.nleap, a.step_size, a.nprops, a.stepfactor, want_p=False)
            qh = r[0]
            g_e2e += int(r[4][:, 1].sum())
        torch.cuda.synchronize()
        dt = torch.tensor([time.time() - t0], dtype=torch.float64, device=dev)
        r = (qh,)
        for _ in range(3):                                                       # warm the pinned p buffers
            r = dr_e.step(r[0], a.nleap, a.step_size, a.nprops, a.stepfactor)
        torch.cuda.synchronize()
        t1 = time.time()
        for _ in range(3):
            r = dr_e.step(r[0], a.nleap, a.step_size, a.nprops, a.stepfactor)
        torch.cuda.synchronize()
        ms_with_p = 1e3 * (time.time() - t1) / 3
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
               "ms_per_step_with_p_returned": ms_with_p, "d2h_bytes_per_step_with_p": d2h + C * d * 8}

    if rank != 0:
        return
    cpu = None
    if not a.no_cpu_baseline and world == 1:
        cpu = cpu_baseline(a, a.cpu_seconds)
    line = {"metric": "grad_evals_per_sec", "value": value, "unit": "grad_evals/s", "n_gpus": world, "steps": K,
            "warmup": a.warmup, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": a.scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(a),
            "leapfrog_steps_per_sec": (grads - leaps) / sec,
            "ess_per_sec": ess_min / sec, "ess_min": ess_min, "ess_note": "min over dims of utils.get_ess_true over the "
            "%d timed transitions of all chains (analytic funnel mean/std); chains started at U(-1,1), so it includes "
            "remaining burn-in bias" % K,
            "grad_evals_per_transition": grads / (K * n_total), "accept_rate": float(stats[-1]) / n_total,
            "code_fractions": {"note": "rank 0: fraction of chain-transitions by DRHMC code (0 = rejected at every "
                               "stage, j+1 = accepted at stage j)", "by_code": code_frac},
            "rhat_max": float(rhat.max()),
            "rhat_note": "R-hat (between- over within-chain variance, all-reduced moments) over the %d timed transitions "
                         "only: with windows that short it mostly says how far a chain moves in %d transitions of the "
                         "funnel, not whether the ensemble has converged" % (K, K), "v_mean": float(mean[0]), "v_std": float(std[0]),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "clocks": clocks.summary(t_wall0, t_wall1),
            "gpu_launches": K * (n_split * a.nprops + 1),
            "gpu_launches_note": "per step: %d chain ranges x %d DR stage kernels + 1 moments kernel (all libhmcb.so)"
                                 % (n_split, a.nprops)}
    print(json.dumps(line), flush=True)


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_hmcb(a)
    td = sys.modules.get("torch.distributed")
    if td is not None and td.is_available() and td.is_initialized():
        td.destroy_process_group()


if __name__ == "__main__":
    main()
