"""Result files and post-processing in the reference's formats, so that its analysis
notebooks and plotting scripts can read a GPU run unchanged.

What the reference writes (scripts_old/funnel/funnelmpi.py:171-211,
scripts_old/tests/test_drhmc_funnel.py:196-279):

  <fpath>/samples/%02d.npy   per MPI rank, [T, C_rank, d]      (T = nsamples, burn-in cut)
  <fpath>/accepts/%02d.npy   [T, C_rank]         accept codes
  <fpath>/probs/%02d.npy     [T, C_rank, ...]    [H0, H1] for HMC, avec[nprops] for DR
  <fpath>/counts/%02d.npy    [T, C_rank, 3]      [Hcount, Vgcount, leapcount]
  <fpath>/samples.npy, accepted.npy, probs.npy, counts.npy
                             rank 0's concatenation along the chain axis (axis=1)
  <fpath>checks.json         ess, mean, std, counts, acceptance, costV_g, cost (+ azess)

Here a "rank" is a torch.distributed rank (one per GPU) owning a contiguous range of global
chain ids (hmc_b200.dist.shard_range), so concatenating the rank files in rank order along
axis 1 restores global chain order exactly as comm.gather + np.concatenate(axis=1) does.

Also the sample post-processing of src/utils.py:75-117 (clean_samples, get_cdf) and
src/utils.py:118-146 (get_ks).  These run on the host in numpy: they are file-format and
notebook helpers applied once per run to a stored window, not the per-transition hot path.
"""
import json
import os
import shutil

import numpy as np

from . import utils as _utils

_KINDS = ("samples", "accepts", "probs", "counts")
# name of the merged file for each per-rank folder (funnelmpi.py:203-206: 'accepted', not 'accepts')
_MERGED = {"samples": "samples", "accepts": "accepted", "probs": "probs", "counts": "counts"}


def _np(x):
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.asarray(x)


def save_rank(fpath, rank, samples, accepted, probs, counts, burnin=0):
    """Per-rank files, as each MPI rank of the reference writes them (funnelmpi.py:166-187).

    Arrays are [T_total, C_rank, ...]; the first `burnin` rows are dropped like the
    reference's `np.array(samples)[burnin:]`.
    """
    arrs = dict(zip(_KINDS, (samples, accepted, probs, counts)))
    for kind, a in arrs.items():
        os.makedirs(os.path.join(fpath, kind), exist_ok=True)
        np.save(os.path.join(fpath, kind, "%02d" % rank), _np(a)[burnin:])


def merge_ranks(fpath, world_size, delete_rank_files=False):
    """What rank 0 does after comm.gather: concatenate along the chain axis and save
    (funnelmpi.py:198-206).  With delete_rank_files the per-rank folders are removed as in
    scripts_old/tests/test_drhmc_funnel.py:232-239.  Returns the four merged arrays."""
    out = []
    for kind in _KINDS:
        parts = [np.load(os.path.join(fpath, kind, "%02d.npy" % r)) for r in range(world_size)]
        merged = np.concatenate(parts, axis=1)
        np.save(os.path.join(fpath, _MERGED[kind]), merged)
        out.append(merged)
    if delete_rank_files:
        for kind in _KINDS:
            shutil.rmtree(os.path.join(fpath, kind))
    return tuple(out)


def gather_to_rank0(arr, group=None):
    """comm.gather(arr, root=0) + np.concatenate(axis=1) over torch.distributed
    (funnelmpi.py:193-201).  Every rank passes its [T, C_rank, ...] block; rank 0 gets the
    [T, C, ...] array, the others None.  Without an initialised process group this is the
    identity (single GPU)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return _np(arr)
    t = arr if hasattr(arr, "detach") else torch.as_tensor(np.asarray(arr))
    t = t.contiguous()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    # chain counts may differ by one between ranks (shard_range), so exchange shapes first
    n_local = torch.tensor([t.shape[1]], dtype=torch.int64, device=t.device)
    counts = [torch.zeros_like(n_local) for _ in range(world)]
    dist.all_gather(counts, n_local, group=group)
    counts = [int(c.item()) for c in counts]
    if rank == 0:
        bufs = [torch.empty((t.shape[0], c) + tuple(t.shape[2:]), dtype=t.dtype, device=t.device) for c in counts]
    else:
        bufs = None
    # point-to-point rather than dist.gather: the blocks differ in size and NCCL's gather wants equal ones
    if rank == 0:
        bufs[0].copy_(t)
        for r in range(1, world):
            dist.recv(bufs[r], src=r, group=group)
    else:
        dist.send(t, dst=0, group=group)
    if rank != 0:
        return None
    return np.concatenate([_np(b) for b in bufs], axis=1)


def funnel_reference_moments(ndim, sigma=3.0, nrep=10_000_000, seed=None):
    """Means and standard deviations the reference compares the funnel against
    (test_drhmc_funnel.py:243-252): v ~ N(0, sigma), theta = exp(v/2) * N(0,1), the theta
    std estimated from `nrep` Monte-Carlo draws exactly as the reference does.  With
    seed=None the closed form std(theta) = exp(sigma^2/4) (9.4877 at sigma=3) is used
    instead of the Monte-Carlo estimate."""
    if seed is None:
        astd = float(np.exp(sigma * sigma / 4.0))
    else:
        rng = np.random.RandomState(seed)
        sigmarep = rng.normal(0, sigma, nrep)
        yvrep = rng.normal(0, 1, nrep)
        astd = float((np.exp(sigmarep / 2.0) * yvrep).std())
    means = np.array([0.0] + [0.0] * (ndim - 1))
    stds = np.array([sigma] + [astd] * (ndim - 1))
    return means, stds


def make_checks(samples, accepted, counts, means, stds, azess=None):
    """The dictionary behind checks.json (test_drhmc_funnel.py:253-272): same keys, same
    formulas.  `azess` (arviz.ess in the reference; arviz is not a dependency here) is
    included only when the caller supplies it."""
    samples, accepted, counts = _np(samples), _np(accepted), _np(counts)
    ess, _ = _utils.get_ess_true(samples, means, stds)
    ess = _np(ess)
    d = {}
    d["ess"] = ess
    d["mean"] = samples.mean(axis=(0, 1))
    d["std"] = samples.std(axis=(0, 1))
    d["counts"] = counts.sum(axis=(0, 1))
    d["acceptance"] = (accepted == 1).sum() / accepted.size
    d["costV_g"] = d["counts"][1] / ess
    d["cost"] = d["counts"][:2].sum() / ess
    if azess is not None:
        d["azess"] = _np(azess)
    return d


def write_checks(fpath, checks):
    """json.dump with the reference's conventions (test_drhmc_funnel.py:268-272): arrays
    become lists of float64, scalars stay numbers, keys sorted, indent 4, and the file name
    is `fpath + 'checks.json'` (string concatenation, as in the reference)."""
    out = {}
    for k, v in checks.items():
        try:
            out[k] = list(np.asarray(v).astype(np.float64))
        except TypeError:
            out[k] = float(v)
    path = fpath + "checks.json"
    with open(path, "w") as fp:
        json.dump(out, fp, sort_keys=True, indent=4)
    return path


def clean_samples(xx, maxv=None, fidsub=1, verbose=False):
    """Thin every chain by its own autocorrelation length and concatenate
    (src/utils.py:75-94).  xx is [T, C, d]; returns (thinned [sum_i ceil(T/subs_i), d], rc, tc).
    The reference's prints are kept behind `verbose`."""
    x = _np(xx)[::fidsub]
    rc, tc = _utils.get_rcc(x)
    rc, tc = _np(rc), _np(tc)
    subs = tc.max(axis=1)
    subs[subs == 0] = 1
    for i in range(subs.size):
        if subs[i] == 1 and (x[0, i] - x[10, i]).sum() == 0:
            subs[i] = x.shape[0]
            if verbose:
                print("Something wrong for correlation length to be 1, setting it to be size of x", i, subs[i])
    if maxv is not None:
        if verbose:
            print("Correlated samples : %d - %0.2f\n" % (maxv, subs.mean()), subs)
        for i in range(subs.size):
            if (subs[i] > maxv) & (subs[i] != x.shape[0]):
                subs[i] = maxv
        if verbose:
            print("Updated subs :  %d\n" % maxv, subs)
    toret = np.concatenate([x[::int(subs[i]), i] for i in range(subs.size)])
    return toret, rc, tc


def get_cdf(x, xref=None, quantiles=None, nbins=20, qs=None, xmin=None, xmax=None):
    """Rank histogram of x [n, d] over the quantile bins of a reference sample
    (src/utils.py:98-117): per dimension, counts of x in bins whose edges are the `qs`
    quantiles of xref (first / last edge optionally pinned to xmin / xmax), and their running
    fraction.  Returns (cdfranks [d, nbins], countsrank [d, nbins]); None when neither xref nor
    quantiles is given, as the reference does."""
    x = _np(x)
    ndim = x.shape[1]
    if quantiles is None:
        if xref is None:
            print("Need either xref or quantiles")
            return None
        if qs is None:
            qs = np.linspace(0.0, 1.0, nbins + 1)
        quantiles = np.quantile(_np(xref)[:, :ndim], qs, axis=0).T.copy()      # [d, nbins+1]
        if xmin is not None:
            quantiles[:, 0] = xmin
        if xmax is not None:
            quantiles[:, -1] = xmax
    countsrank = np.array([np.histogram(x[:, i], bins=quantiles[i])[0] for i in range(ndim)])
    cdfranks = np.cumsum(countsrank, axis=1) / countsrank.sum(axis=1, keepdims=True)
    return cdfranks, countsrank


def get_ks(x, xref):
    """Two-sample Kolmogorov-Smirnov statistic and p-value per dimension: first for the whole
    sample, then for the tails below mu-2s, below mu-1s, above mu+1s, above mu+2s, with mu, s
    taken from the reference sample (src/utils.py:118-146).  A tail on which scipy raises
    (empty) scores (0, 1e-99) like the reference.  Returns (stats, pvals), lists [d][5]."""
    from scipy.stats import ks_2samp

    x, xref = _np(x), _np(xref)
    mu, sd = xref.mean(axis=0), xref.std(axis=0)
    stats, pvals = [], []
    for i in range(x.shape[1]):
        xi, ri = x[:, i], xref[:, i]
        pairs = [(xi, ri)]
        for j in (-2, -1, 1, 2):
            cut = mu[i] + j * sd[i]
            pairs.append((xi[xi < cut], ri[ri < cut]) if j < 0 else (xi[xi > cut], ri[ri > cut]))
        row_s, row_p = [], []
        for a, b in pairs:
            try:
                r = ks_2samp(a, b)
                row_s.append(r.statistic)
                row_p.append(r.pvalue)
            except Exception:
                row_s.append(0)
                row_p.append(1e-99)
        stats.append(row_s)
        pvals.append(row_p)
    return stats, pvals
