"""Numeric diagnostics of the reference's ``src/utils.py`` on the hot path's
outputs: ``get_ess_true`` (:8-27), ``get_ess_ref`` (:30-55), ``get_rcc`` (:59-71),
``do_hmc`` (:207-237), plus the streamed, multi-GPU forms the 2^20-chain runs need
(samples are never stored; per-chain running sums are reduced to 4*d doubles per
rank and all-reduced over NCCL).

Sample arrays are ``x[Nsamples, Nchains, Nparams]`` as in the reference; numpy in,
numpy out -- torch tensors (any device) in, torch out.
"""
import time

import numpy as np
import torch

from . import dist


def _is_t(x):
    return torch.is_tensor(x)


def _like(x, v):
    if _is_t(x):
        return torch.as_tensor(np.asarray(v), dtype=x.dtype, device=x.device)
    return np.asarray(v, dtype=np.float64)


def get_ess_true(x, mean, std):
    """ESS from the error of per-chain means against known moments (src/utils.py:8-27).
    Returns (ess[Nparams], esschain[Nchains, Nparams])."""
    mean, std = _like(x, mean), _like(x, std)
    err = x.mean(0) - mean
    rms = (err ** 2).mean(0) ** 0.5
    ess = (std / rms) ** 2
    esschain = (std / err) ** 2
    return ess * x.shape[1], esschain * x.shape[1]


def get_ess_ref(x, xref):
    """ESS against reference draws xref[Nsamples, Nparams] (src/utils.py:30-55).
    Returns (ess, ess2, err, err2)."""
    sd = (lambda a: a.std(0, unbiased=False)) if _is_t(x) else (lambda a: a.std(axis=0))
    err = x.mean(0) - xref.mean(0)
    rms = (err ** 2).mean(0) ** 0.5
    ess = (sd(xref) / rms) ** 2
    x2, xref2 = x ** 2, xref ** 2
    err2 = x2.mean(0) - xref2.mean(0)
    rms2 = (err2 ** 2).mean(0) ** 0.5
    ess2 = (sd(xref2) / rms2) ** 2
    return ess * x.shape[1], ess2 * x.shape[1], err, err2


def get_rcc(x, c=5, threshold=0.2):
    """FFT autocorrelation along axis 0 and the integrated time at the Sokal window
    (src/utils.py:59-71): tau = 1 + 2 cumsum(rho); window = first lag with lag >= c*tau."""
    t = x if _is_t(x) else torch.as_tensor(np.asarray(x, dtype=np.float64))
    n = t.shape[0]
    tmp = t - t.mean(0)
    f = torch.fft.rfft(tmp, n=2 * n, dim=0)
    ac = torch.fft.irfft(f * f.conj(), n=2 * n, dim=0)[:n]
    rcc = ac / ac[0]
    idx = torch.arange(n, device=t.device, dtype=t.dtype).reshape(-1, 1, 1)
    taus = 2 * torch.cumsum(rcc, dim=0) + 1
    window = torch.argmin((idx < c * taus).to(torch.int8), dim=0)
    tc = torch.gather(taus, 0, window[None])[0].to(torch.int64)
    if _is_t(x):
        return rcc, tc
    return rcc.numpy(), tc.numpy()


def do_hmc(stepfunc, initstate, nsamples=1000, burnin=100):
    """The reference's generic sample loop (src/utils.py:207-237), batched: ``stepfunc(q)``
    advances ALL chains (q is [C,d]) and returns (q, p, accepted, probs, counts), e.g.
    ``lambda q: hmc.hmc_step(q, N, eps)``.  Returns samples[ns,C,d], accepted[ns,C],
    probs[ns,C,...], counts[ns,C,3] with burn-in dropped."""
    q = initstate
    samples, accepts, probs, counts = [], [], [], []
    start = time.time()
    for i in range(nsamples + burnin):
        q, _, acc, prob, count = stepfunc(q)
        if i >= burnin:
            samples.append(q)
            accepts.append(acc)
            probs.append(prob)
            counts.append(count)
    stack = torch.stack if _is_t(samples[0]) else np.stack
    end = time.time()
    do_hmc.last_seconds = end - start
    return stack(samples), stack(accepts), stack(probs), stack(counts)


# ------------------------------------------------------------------ streamed / sharded diagnostics
class RunningMoments:
    """Per-chain running sums of q and q^2 on the device, folded in after every
    transition (hmcb_moments_accumulate), reduced across chains and ranks on demand."""

    def __init__(self, engine, n_chains, d):
        self.engine = engine
        engine._ensure()
        self.sum = torch.zeros((n_chains, d), dtype=engine.dtype, device=engine.device)
        self.sumsq = torch.zeros_like(self.sum)
        self.n = 0

    def reset(self):
        self.sum.zero_()
        self.sumsq.zero_()
        self.n = 0

    def add(self, q):
        self.engine.moments_accumulate(q, self.sum, self.sumsq)
        self.n += 1

    def reduce(self, true_mean=None):
        """Global (all ranks) [4,d] sums and the global chain count."""
        out = self.engine.moments_reduce(self.sum, self.sumsq, self.n, true_mean)
        return all_reduce_sums(out, self.sum.shape[0])

    def ess_true(self, true_mean, true_std):
        """utils.get_ess_true over all chains of all ranks (src/utils.py:23-27)."""
        sums, C = self.reduce(true_mean)
        return ess_from_sums(sums, C, true_std)

    def rhat(self):
        sums, C = self.reduce()
        return rhat_from_sums(sums, C, float(self.n))

    def mean_std(self):
        """Pooled mean and std over all chains and samples."""
        sums, C = self.reduce()
        return mean_std_from_sums(sums, C, float(self.n))


def local_moment_sums(x, true_mean=None):
    """The [4,d] block hmcb_moments_reduce produces, from stored samples x[T,C,d] (torch):
    sum_c m, sum_c m^2, sum_c s2 (ddof=1), sum_c (m-mu)^2.  Host-side twin used by the
    CPU (gloo) tests of the cross-rank arithmetic."""
    m = x.mean(0)
    s2 = x.var(0, unbiased=True) if x.shape[0] > 1 else torch.zeros_like(m)
    mu = torch.zeros(x.shape[2], dtype=x.dtype) if true_mean is None else torch.as_tensor(
        np.asarray(true_mean, dtype=np.float64)).to(x.dtype)
    return torch.stack([m.sum(0), (m * m).sum(0), s2.sum(0), ((m - mu) ** 2).sum(0)]).to(torch.float64)


def all_reduce_sums(sums, n_chains_local):
    """All-reduce the [4,d] block and the chain count over the process group."""
    packed = torch.cat([sums.reshape(-1), torch.tensor([float(n_chains_local)], dtype=torch.float64,
                                                       device=sums.device)])
    dist.all_reduce_sum(packed)
    return packed[:-1].reshape(4, -1), float(packed[-1])


def ess_from_sums(sums, C, true_std):
    """(std/rms)^2 * C with rms^2 = mean_c (mean_t x - mean)^2 (src/utils.py:23-27)."""
    std = torch.as_tensor(np.asarray(true_std, dtype=np.float64), device=sums.device)
    return (std ** 2 / (sums[3] / C)) * C


def rhat_from_sums(sums, C, T):
    """Potential scale reduction from per-chain means and variances (Gelman-Rubin; not in
    the reference's src/, which only calls arviz.rhat in post-processing)."""
    W = sums[2] / C
    var_m = (sums[1] - sums[0] ** 2 / C) / (C - 1.0)
    return torch.sqrt(((T - 1.0) / T * W + var_m) / W)


def mean_std_from_sums(sums, C, T):
    mean = sums[0] / C
    ex2 = (sums[2] * (T - 1.0) / T + sums[1]) / C   # E_c[ biased var + m^2 ]
    return mean, torch.sqrt(torch.clamp(ex2 - mean ** 2, min=0.0))
