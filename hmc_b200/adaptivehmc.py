"""Batched drop-in for the reference's ``src/adaptivehmc.py``:

``HMC`` (:5-99)           the third copy of vanilla HMC, entry point ``hmc_step``
``AdHMC_eps0`` (:105-192) step size drawn per transition from a distribution built by trial
                          trajectories, with the matching Hastings correction; ``Nleap`` steps, or
                          with ``Tint`` given int(Tint/eps)+1 steps of the drawn step size
``AdHMC`` (:196-354)      the step size walks inside the trajectory: after every ``nleap`` steps it
                          is multiplied by one of ``ratios``, chosen by the energy error of the
                          candidate segments, with the reverse-walk Hastings factor
``AdHMC_tint`` (:362-519) the same with segments of fixed integration time Tint/ntimes

``hmc_step`` returns the reference's 6-tuple ``(q, p, accepted, [prob, prob/hastings, hastings],
checks, [Hcount, Vgcount, leapcount])`` with ``checks = stack([pf_den, pb_num, eps])``; batched
inputs give ``probs[C,3]``, ``checks[C,3]`` (eps0) or ``[C,3,N]`` (walks), ``counts[C,3]``.
"""
import numpy as np
import torch

from .engine import Batch
from .pyhmc import HMC as _Base


class HMC(_Base):
    """adaptivehmc.HMC (src/adaptivehmc.py:5-99): codes 1./0./2., entry ``hmc_step``."""

    def leapfrog1(self, q, p, step_size, Vgq=None):
        """One leapfrog step (src/adaptivehmc.py:50-61)."""
        return self.leapfrog(q, p, 1, step_size)

    def hmc_step(self, q, N, step_size, p=None, u=None):
        return self._step(q, N, step_size, p=p, u=u)


class AdHMC_eps0(HMC):
    def get_steps(self, smin=0.01, smax=1.0, ntry=20, logspace=True):
        """The trial step sizes of get_stepsize (src/adaptivehmc.py:114-115)."""
        if logspace:
            return np.logspace(np.log10(smin), np.log10(smax), ntry)
        return np.linspace(smin, smax, ntry)

    def get_stepsize(self, q0, p0, smin=0.01, smax=1.0, ntry=20, logspace=True, nsteps=1, eps=None, u=None):
        """The step-size distribution on its own (src/adaptivehmc.py:111-134; the same method in AdHMC
        :202-224 and AdHMC_tint :368-390): ntry trial trajectories of nsteps steps from (q0, p0), weights
        exp(H0 - H_i) sqrt(eps_i) normalised to pp.  Without ``eps`` it draws a step size -- a bin from the
        CDF, then uniformly inside the bin -- and returns (step size, pp[bin + 1]); with ``eps`` it returns
        the probability of the bin above eps.  Batched over chains; ``u`` [C, 2] injects the two uniforms
        (else torch's generator on the device is used).  hmc_step does all of this inside one kernel;
        this method is ntry leapfrog launches and exists for callers that use it directly."""
        e = self.engine
        bq = Batch(q0, e.dev, e.dtype)
        bp = Batch(p0, e.dev, e.dtype)
        e.check_dim(bq)
        steps = self.get_steps(smin, smax, ntry, logspace)
        H0 = self.H(bq.q, bp.q)
        Hs = []
        for ss in steps:
            q1, p1 = self.leapfrog(bq.q, bp.q, nsteps, float(ss))
            Hs.append(self.H(q1, p1))
        Hs = torch.stack(Hs, dim=1)                                            # [C, ntry]
        st = torch.as_tensor(steps, dtype=torch.float64, device=Hs.device)
        pp = torch.exp(H0.double()[:, None] - Hs.double()) * st ** 0.5
        pp = torch.where(torch.isnan(pp) | torch.isinf(pp), torch.zeros_like(pp), pp)
        pp = pp / pp.sum(dim=1, keepdim=True)
        if eps is not None:
            epst = torch.as_tensor(eps, dtype=torch.float64, device=pp.device).reshape(-1, 1).expand(pp.shape[0], 1)
            j = (st[None, :] > epst).to(torch.int8).argmax(dim=1)              # first bin edge above eps
            return bq.out(pp.gather(1, j[:, None])[:, 0])
        cdf = torch.cumsum(pp, dim=1)
        if u is None:
            ut = torch.rand((pp.shape[0], 2), dtype=torch.float64, device=pp.device)
        else:
            ut = torch.as_tensor(np.asarray(u, dtype=np.float64) if not torch.is_tensor(u) else u).to(pp.device).reshape(-1, 2)
        lo = cdf.min(dim=1).values
        sx = lo + (1.0 - lo) * ut[:, 0]
        above = sx[:, None] > cdf                                              # isx = last index with sx > cdf
        isx = (above.to(torch.int64) * torch.arange(ntry, device=pp.device)[None, :]).max(dim=1).values
        if bool(((~above.any(dim=1)) | (isx > ntry - 2)).any()):
            raise IndexError("get_stepsize: no bin below the draw (all trial weights NaN, or the draw fell in the "
                             "last bin); the reference raises here too")
        sx2 = st[isx] + (st[isx + 1] - st[isx]) * ut[:, 1]
        return bq.out(sx2), bq.out(pp.gather(1, (isx + 1)[:, None])[:, 0])

    def hmc_step(self, q0, Nleap, smin=0.01, smax=1.0, Tint=0, ntry=10, nsteps=1, p=None, u=None):
        if (Tint == 0) and (Nleap == 0):
            raise ValueError("Tint and Nleap cannot be both zeros")
        if (Tint != 0) and (Nleap != 0):
            raise ValueError("Tint and Nleap both given and are inconsistent")
        e = self.engine
        b = Batch(q0, e.dev, e.dtype)
        e.check_dim(b)
        out = e.adhmc_eps0_step(b.q, Nleap, self.get_steps(smin, smax, ntry), nsteps, p=p, u=u, tint=float(Tint))
        code = out["code"].to(torch.float64)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            return (b.out(out["q"]), b.out(out["p"]), float(code[0]), out["probs"][0].tolist(), b.out(out["aux"]), cnt)
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return (b.out(out["q"]), b.out(out["p"]), b.out(code), b.out(out["probs"]), b.out(out["aux"]),
                b.out(out["counts"]))


class AdHMC(AdHMC_eps0):
    """adaptivehmc.AdHMC (src/adaptivehmc.py:196-354).  ``u``, if given, holds N+2 uniforms per chain
    (N = Nleap // nleap) in the order the reference consumes them: bin draw, within-bin draw, one per
    step-size choice (np.random.choice), accept test."""

    _tint = False

    def _walk(self, q0, N, nleap, tint, ratios, pwts0, smin, smax, ntry_eps0, nsteps_eps0, logeps, p, u):
        if N < 2:
            raise ValueError("need at least two segments (the reference leaves its step-size walk undefined for one)")
        e = self.engine
        b = Batch(q0, e.dev, e.dtype)
        e.check_dim(b)
        out = e.adhmc_walk_step(b.q, N, nleap, tint, list(ratios), list(pwts0), smin, smax,
                                self.get_steps(smin, smax, ntry_eps0, logeps), nsteps_eps0, p=p, u=u)
        code = out["code"].to(torch.float64)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            return (b.out(out["q"]), b.out(out["p"]), float(code[0]), out["probs"][0].tolist(), b.out(out["checks"]), cnt)
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return (b.out(out["q"]), b.out(out["p"]), b.out(code), b.out(out["probs"]), b.out(out["checks"]),
                b.out(out["counts"]))

    def hmc_step(self, q0, Nleap=100, nleap=10, ratios=(1 / np.sqrt(2), np.sqrt(2)), pwts0=(1.0, 1.0), smin=0.01,
                 smax=1.0, ntry_eps0=10, nsteps_eps0=1, logeps=True, getstepfunc=None, verbose=False, p=None, u=None):
        return self._walk(q0, int(Nleap // nleap), nleap, 0.0, ratios, pwts0, smin, smax, ntry_eps0, nsteps_eps0, logeps,
                          p, u)


class AdHMC_tint(AdHMC):
    """adaptivehmc.AdHMC_tint (src/adaptivehmc.py:362-519): ``ntimes`` segments, each of integration
    time Tint/ntimes, i.e. max(1, int(tint/eps)) steps at the segment's step size."""

    def hmc_step(self, q0, Tint=10, ntimes=10, ratios=(1 / np.sqrt(2), np.sqrt(2)), pwts0=(1.0, 1.0), smin=0.01,
                 smax=1.0, ntry_eps0=10, nsteps_eps0=1, logeps=True, verbose=False, p=None, u=None):
        return self._walk(q0, int(ntimes), 0, Tint / ntimes, ratios, pwts0, smin, smax, ntry_eps0, nsteps_eps0, logeps,
                          p, u)
