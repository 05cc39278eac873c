"""Batched drop-in for the hot-path part of the reference's ``src/adaptivehmc.py``:
``HMC`` (:5-99, the third copy of vanilla HMC, entry point ``hmc_step``) and
``AdHMC_eps0`` (:105-192), HMC whose step size is drawn per transition from a
distribution built by trial trajectories, with the matching Hastings correction.

``AdHMC_eps0.hmc_step(q0, Nleap, smin, smax, Tint, ntry, nsteps)`` returns the
reference's 6-tuple ``(q, p, accepted, [prob, prob/hastings, hastings],
stack([pf_den, pb_num, eps]), [Hcount, Vgcount, leapcount])``; batched inputs give
``probs[C,3]``, ``aux[C,3]``, ``counts[C,3]``.  Only ``Tint == 0`` (fixed number of
leapfrog steps) runs on the GPU: with ``Tint != 0`` the reference makes the number
of steps a function of the drawn step size, which is a later row.
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

    def hmc_step(self, q0, Nleap, smin=0.01, smax=1.0, Tint=0, ntry=10, nsteps=1, p=None, u=None):
        if (Tint == 0) and (Nleap == 0):
            raise ValueError("Tint and Nleap cannot be both zeros")
        if (Tint != 0) and (Nleap != 0):
            raise ValueError("Tint and Nleap both given and are inconsistent")
        if Tint != 0:
            raise NotImplementedError("AdHMC_eps0 with Tint != 0 (step count depends on the drawn step size) "
                                      "is not on the GPU path yet")
        e = self.engine
        b = Batch(q0, e.dev, e.dtype)
        e.check_dim(b)
        out = e.adhmc_eps0_step(b.q, Nleap, self.get_steps(smin, smax, ntry), nsteps, p=p, u=u)
        code = out["code"].to(torch.float64)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            return (b.out(out["q"]), b.out(out["p"]), float(code[0]), out["probs"][0].tolist(), b.out(out["aux"]), cnt)
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return (b.out(out["q"]), b.out(out["p"]), b.out(code), b.out(out["probs"]), b.out(out["aux"]),
                b.out(out["counts"]))
