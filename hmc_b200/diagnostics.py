"""Drop-in for the reference's ``src/diagnostics.py``: ``plot_hist``, ``plot_trace``,
``plot_scatter`` and ``plot_autorcc`` with the reference's names, arguments
(``mysamples`` is ``[Nsamples, Nchains, Ndim]``, ``fpath`` a directory) and output file names.

What the reference's functions compute -- the thinned samples behind the histograms and the
scatter, the traces, and for ``plot_autorcc`` the normalised autocorrelation of every chain with
its integrated time at the window ``m > 5 (1 + 2 sum_{l<m} rho_l)`` (src/diagnostics.py:69-80) --
is computed here and always written next to where the figure goes, as ``<figure name>.npz``.
The figure itself is drawn with matplotlib when matplotlib can be imported (it is not a
dependency of this package and is absent from the build image); without it the functions
still run and leave the numbers.  The autocorrelation goes through an FFT on whatever device
the samples are on instead of ``np.correlate``'s O(T^2) loop per chain; the values agree to
rounding (tests/test_diagnostics_cpu.py runs the reference's own nested function beside it).
"""
import os

import numpy as np
import torch


def _t(x):
    return x if torch.is_tensor(x) else torch.as_tensor(np.asarray(x, dtype=np.float64))


def _np(x):
    return x.detach().cpu().numpy() if torch.is_tensor(x) else np.asarray(x)


def _plt():
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        return plt
    except Exception:
        return None


def autocorr_window(x, maxlag=500):
    """The numeric core of plot_autorcc (src/diagnostics.py:69-80) for one parameter:
    x is [Nsamples, Nchains]; returns (rcc [Nsamples, Nchains], tcc [Nchains]) with
    rcc the autocorrelation of the standardised chain divided by its lag-0 value and
    tcc = 1 + 2 sum_{l<m} rcc_l at the first m < maxlag with m > 5 (1 + 2 sum_{l<m} rcc_l)
    (m = maxlag-1 if the window never closes, as the reference's loop leaves it)."""
    t = _t(x).to(torch.float64)
    n = t.shape[0]
    xp = (t - t.mean(0)) / t.std(0, unbiased=False)
    f = torch.fft.rfft(xp, n=2 * n, dim=0)
    rcc = torch.fft.irfft(f * f.conj(), n=2 * n, dim=0)[:n]
    rcc = rcc / rcc[0]
    # S[m] = 1 + 2 * sum_{l<m} rcc_l for m = 0 .. maxlag-1; past the end of the chain the slice rcc[:m] of
    # the reference simply stops growing
    zero = torch.zeros((1,) + tuple(rcc.shape[1:]), dtype=rcc.dtype, device=rcc.device)
    c = torch.cumsum(rcc[:maxlag - 1], dim=0)
    if c.shape[0] < maxlag - 1:
        c = torch.cat([c, c[-1:].expand(maxlag - 1 - c.shape[0], *c.shape[1:])], dim=0)
    S = 1 + 2 * torch.cat([zero, c], dim=0)
    m = torch.arange(maxlag, device=rcc.device, dtype=rcc.dtype).reshape(-1, 1)
    closed = m > 5 * S
    first = torch.where(closed.any(dim=0), closed.to(torch.int8).argmax(dim=0),
                        torch.full((rcc.shape[1],), maxlag - 1, device=rcc.device))
    tcc = S.gather(0, first[None])[0]
    return rcc, tcc


def _save(fpath, name, **arrays):
    os.makedirs(fpath, exist_ok=True)
    np.savez_compressed(os.path.join(fpath, name + ".npz"), **{k: _np(v) for k, v in arrays.items()})


def plot_hist(mysamples, fpath, sub=100):
    """Histograms of the thinned samples: parameter 0 to hist_sigma.png, parameters 1..4 (log
    counts) to hist_alpha.png (src/diagnostics.py:6-23)."""
    s = _t(mysamples)
    ns, nc, ndim = s.shape
    thin = s[::sub]
    sig = thin[..., 0].flatten()
    alph = [thin[..., i].flatten() for i in range(1, min(5, ndim))]
    _save(fpath, "hist_sigma", samples=sig)
    _save(fpath, "hist_alpha", **{"alpha%d" % (i + 1): a for i, a in enumerate(alph)})
    plt = _plt()
    if plt is None:
        return
    plt.hist(_np(sig), bins="auto", alpha=0.5)
    plt.grid(which="both")
    plt.savefig(fpath + "/hist_sigma.png")
    plt.close()
    fig, ax = plt.subplots(ndim - 1, 1, figsize=(5, ndim * 3 - 3), sharex=True, sharey=True)
    if ndim == 2:
        ax = [ax]
    for i, a in enumerate(alph):
        ax[i].hist(_np(a), bins="auto", alpha=0.5)
        ax[i].grid(which="both")
        ax[i].set_yscale("log")
    plt.tight_layout()
    plt.savefig(fpath + "/hist_alpha.png")
    plt.close()


def plot_trace(mysamples, fpath):
    """Trace of every chain: parameter 0 to trace_sigma.png, parameter i to trace_alpha<i>.png
    (src/diagnostics.py:26-47)."""
    s = _t(mysamples)
    ns, nc, ndim = s.shape
    _save(fpath, "trace_sigma", trace=s[..., 0])
    for i in range(1, ndim):
        _save(fpath, "trace_alpha%d" % i, trace=s[..., i])
    plt = _plt()
    if plt is None:
        return
    x = _np(s)
    for i in range(ndim):
        fig, ax = plt.subplots(nc, 1, figsize=(5, nc * 3), sharex=True)
        if nc == 1:
            ax = [ax]
        for j in range(nc):
            ax[j].plot(x[:, j, i])
            ax[j].grid(which="both")
        plt.tight_layout()
        plt.savefig(fpath + ("/trace_sigma.png" if i == 0 else "/trace_alpha%d.png" % i))
        plt.close()


def plot_scatter(mysamples, fpath, sub=100):
    """Thinned parameter 0 against each other parameter, scatter.png (src/diagnostics.py:51-61)."""
    s = _t(mysamples)
    ns, nc, ndim = s.shape
    thin = s[::sub]
    _save(fpath, "scatter", sigma=thin[..., 0].flatten(), alpha=thin[..., 1:].reshape(-1, ndim - 1))
    plt = _plt()
    if plt is None:
        return
    x = _np(thin)
    fig, ax = plt.subplots(ndim - 1, 1, figsize=(5, ndim * 3 - 3), sharex=True, sharey=True)
    if ndim == 2:
        ax = [ax]
    for i in range(1, ndim):
        ax[i - 1].scatter(x[..., 0], x[..., i].flatten(), marker=".")
        ax[i - 1].grid(which="both")
    plt.tight_layout()
    plt.savefig(fpath + "/scatter.png")
    plt.close()


def plot_autorcc(mysamples, fpath):
    """Autocorrelation of every chain and its windowed integrated time, per parameter:
    rcc_sigma.png for parameter 0, rcc_alpha.png for the others (src/diagnostics.py:65-101).
    Returns {parameter index: (rcc[:500], tcc)}, which the reference only draws."""
    s = _t(mysamples)
    ns, nc, ndim = s.shape
    out = {}
    for i in range(ndim):
        rcc, tcc = autocorr_window(s[..., i])
        out[i] = (rcc[:500], tcc)
    _save(fpath, "rcc_sigma", rcc=out[0][0], tcc=out[0][1])
    _save(fpath, "rcc_alpha", **{"rcc%d" % i: out[i][0] for i in range(1, ndim)},
          **{"tcc%d" % i: out[i][1] for i in range(1, ndim)})
    plt = _plt()
    if plt is None:
        return out
    fig, ax = plt.subplots()
    plt.plot(_np(out[0][0]))
    plt.grid()
    plt.title("%d ($\\pm$%d)" % (float(out[0][1].mean()), float(out[0][1].std())))
    plt.savefig(fpath + "/rcc_sigma.png")
    plt.close()
    fig, ax = plt.subplots(ndim - 1, 1, figsize=(5, ndim * 3 - 3), sharex=True, sharey=True)
    if ndim == 2:
        ax = [ax]
    for i in range(1, ndim):
        ax[i - 1].plot(_np(out[i][0]))
        ax[i - 1].grid()
        ax[i - 1].set_title("%d ($\\pm$%d)" % (float(out[i][1].mean()), float(out[i][1].std())))
    plt.savefig(fpath + "/rcc_alpha.png")
    plt.close()
    return out
