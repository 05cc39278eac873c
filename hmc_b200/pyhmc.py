"""Batched drop-in for the reference's ``src/pyhmc.py`` vanilla samplers.

``HMC`` mirrors pyhmc.HMC (src/pyhmc.py:5-69) and ``PyHMC`` mirrors pyhmc.PyHMC
(src/pyhmc.py:73-138): same constructor order, same method names, same return
tuples and accept codes (1./0./2.), but ``q`` may be ``[C, d]`` and one call
advances all C chains in one fused CUDA kernel (hmcb_hmc_step).  A 1-D ``q`` is
one chain and returns the reference's unbatched shapes.

Extra keyword arguments, not in the reference: ``p=`` / ``u=`` inject the
momentum and uniform draws the reference would have taken from
``np.random.normal`` / ``np.random.uniform`` (src/pyhmc.py:66,59) -- parity
mode; ``device=``, ``dtype=``, ``seed=``, ``chain_offset=`` at construction.
"""
import torch

from . import _lib
from .engine import Batch, Engine


class HMC:
    _rule = _lib.RULE_PYHMC

    def __init__(self, log_prob, grad_log_prob=None, invmetric_diag=None, log_prob_and_grad=None, **engine_kw):
        self.engine = Engine(log_prob, grad_log_prob, log_prob_and_grad, invmetric_diag, **engine_kw)
        self.model = self.engine.model
        self.log_prob, self.grad_log_prob = self.model.log_prob, self.model.grad_log_prob
        self.invmetric_diag = 1.0 if invmetric_diag is None else invmetric_diag
        self.metricstd = self.invmetric_diag ** -0.5
        self.leapcount, self.Vgcount, self.Hcount = 0, 0, 0

    # ---- energies (batched on the device) -----------------------------------
    def _minv(self):
        e = self.engine
        e._ensure()
        if e.invmetric_diag is None:
            return None
        return e._keep[-1]

    def V(self, x):
        b = Batch(x, self.engine.dev, self.engine.dtype)
        logp, _ = self.engine.logp_grad(b.q)
        return b.out(-logp)

    def V_g(self, x):
        b = Batch(x, self.engine.dev, self.engine.dtype)
        self.Vgcount += 1
        _, g = self.engine.logp_grad(b.q)
        return b.out(-g)

    def KE(self, p):
        b = Batch(p, self.engine.dev, self.engine.dtype)
        m = self._minv()
        ke = 0.5 * ((b.q * b.q) if m is None else (b.q * b.q * m)).sum(dim=1)
        return b.out(ke)

    def KE_g(self, p):
        b = Batch(p, self.engine.dev, self.engine.dtype)
        m = self._minv()
        return b.out(b.q if m is None else b.q * m)

    def unit_norm_KE(self, p):
        """0.5 * sum(p**2) whatever the metric (src/pyhmc.py:24-25)."""
        b = Batch(p, self.engine.dev, self.engine.dtype)
        return b.out(0.5 * (b.q * b.q).sum(dim=1))

    def unit_norm_KE_g(self, p):
        return p

    def H(self, q, p):
        self.Hcount += 1
        bq = Batch(q, self.engine.dev, self.engine.dtype)
        bp = Batch(p, self.engine.device, self.engine.dtype)
        logp, _ = self.engine.logp_grad(bq.q)
        m = self._minv()
        ke = 0.5 * ((bp.q * bp.q) if m is None else (bp.q * bp.q * m)).sum(dim=1)
        return bq.out(-logp + ke)

    # ---- trajectory -----------------------------------------------------------
    def leapfrog(self, q, p, N, step_size):
        """N leapfrog steps for every chain (src/pyhmc.py:35-48): N+1 gradient evaluations."""
        self.leapcount += 1
        self.Vgcount += max(int(N), 1) + 1
        bq = Batch(q, self.engine.dev, self.engine.dtype)
        bp = Batch(p, self.engine.device, self.engine.dtype)
        self.engine.check_dim(bq)
        q1, p1 = self.engine.leapfrog(bq.q, bp.q, N, step_size)
        return bq.out(q1), bq.out(p1)

    def metropolis(self, qp0, qp1, u=None):
        """Accept/reject between two phase-space points (src/pyhmc.py:50-61).
        Standalone version of the test the fused step kernel applies."""
        dev = self.engine.dev
        b0q, b0p = Batch(qp0[0], dev, self.engine.dtype), Batch(qp0[1], dev, self.engine.dtype)
        b1q, b1p = Batch(qp1[0], dev, self.engine.dtype), Batch(qp1[1], dev, self.engine.dtype)
        H0 = Batch(self.H(b0q.q, b0p.q), dev, self.engine.dtype).q.reshape(-1)
        H1 = Batch(self.H(b1q.q, b1p.q), dev, self.engine.dtype).q.reshape(-1)
        prob = torch.exp(H0 - H1)
        bad = torch.isnan(prob) | torch.isinf(prob) | ((b0q.q - b1q.q).sum(dim=1) == 0)
        if u is None:
            u = torch.rand(b0q.C, dtype=torch.float64, device=dev)
        else:
            u = torch.as_tensor(u, dtype=torch.float64).reshape(-1).to(dev)
        acc = ~bad & ~(u > torch.clamp(prob.double(), max=1.0))
        bad_code = 2.0 if self._rule == _lib.RULE_PYHMC else -1.0
        code = torch.where(bad, torch.full_like(H0, bad_code), acc.to(H0.dtype))
        qn = torch.where(acc[:, None], b1q.q, b0q.q)
        pn = torch.where(acc[:, None], b1p.q, b0p.q)
        Hs = torch.stack([H0, H1], dim=1)
        if b0q.unbatched:
            return b0q.out(qn), b0q.out(pn), float(code[0]), [float(H0[0]), float(H1[0])]
        return b0q.out(qn), b0q.out(pn), b0q.out(code), b0q.out(Hs)

    # ---- transition -------------------------------------------------------------
    def _step(self, q, N, step_size, p=None, u=None):
        e = self.engine
        e._ensure()
        b = Batch(q, e.device, e.dtype)
        e.check_dim(b)
        out = e.hmc_step(b.q, N, step_size, self._rule, p=p, u=u)
        code = out["code"].to(torch.float64)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            Hs = out["Hs"][0].tolist()
            return b.out(out["q"]), b.out(out["p"]), float(code[0]), Hs, cnt
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return b.out(out["q"]), b.out(out["p"]), b.out(code), b.out(out["Hs"]), b.out(out["counts"])

    def step(self, q, N, step_size, p=None, u=None):
        """One HMC transition per chain (src/pyhmc.py:64-69).  Returns
        (q, p, accepted, [H0, H1], [Hcount, Vgcount, leapcount]); batched inputs give
        accepted[C], Hs[C,2], counts[C,3]."""
        return self._step(q, N, step_size, p=p, u=u)


class PyHMC(HMC):
    """pyhmc.PyHMC (src/pyhmc.py:73-138): same sampler, entry point ``hmc_step``.
    (The reference passes invmetric_diag=None to its base and then re-applies the
    caller's metric itself, src/pyhmc.py:76,83-85 -- the metric IS honoured.)"""

    def hmc_step(self, q, N, step_size, p=None, u=None):
        return self._step(q, N, step_size, p=p, u=u)


class PyHMC_multistep(HMC):
    """pyhmc.PyHMC_multistep (src/pyhmc.py:489-598), the delayed-rejection sampler the
    reference's funnel experiments run (scripts_old/tests/test_drhmc_funnel.py:157-159).
    Same proposals and acceptance probabilities as algorithms.DRHMC, with this class's own
    bookkeeping: the first stage accepts iff u <= alpha_0 (:580), codes are the accepting
    stage j >= 0 or -1, and Hcount counts the energy that get_num re-evaluates at its base
    (:541)."""

    def multi_step(self, m, q0, N, ss, fsub, p=None, u=None):
        """Returns (q, p, code, avec[m], [Hcount, Vgcount, leapcount])."""
        e = self.engine
        b = Batch(q0, e.dev, e.dtype)
        e.check_dim(b)
        out = e.drhmc_step(b.q, N, ss, m, fsub, p=p, u=u, convention=_lib.DR_MULTISTEP)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            return b.out(out["q"]), b.out(out["p"]), int(out["code"][0]), b.out(out["avec"]), cnt
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return b.out(out["q"]), b.out(out["p"]), b.out(out["code"]), b.out(out["avec"]), b.out(out["counts"])


class _ProbMultistep(HMC):
    """Probabilistic delayed rejection: after a rejection the next, finer stage is only
    attempted with probability ptries[j-1] (one more uniform per stage; code -2 when the
    stage is declined), and the ghost products carry prod(ptries).  multi_step returns the
    reference's six values (q, p, code, avec[m], [Hcount, Vgcount, leapcount], ptries[m])."""
    _convention = None

    def multi_step(self, m, q0, N, ss, fsub, p=None, u=None):
        """u, if given, holds 2m-1 uniforms per chain in the order the reference consumes them:
        u_acc0, u_try1, u_acc1, u_try2, ..."""
        e = self.engine
        b = Batch(q0, e.dev, e.dtype)
        e.check_dim(b)
        out = e.drhmc_step(b.q, N, ss, m, fsub, p=p, u=u, convention=self._convention)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            return (b.out(out["q"]), b.out(out["p"]), int(out["code"][0]), b.out(out["avec"]), cnt,
                    b.out(out["ptries"]))
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return (b.out(out["q"]), b.out(out["p"]), b.out(out["code"]), b.out(out["avec"]), b.out(out["counts"]),
                b.out(out["ptries"]))


class PyHMC_multistep_tries(_ProbMultistep):
    """pyhmc.PyHMC_multistep_tries (src/pyhmc.py:233-349): p_try_j = 1 - min(1, e^{H0-Hj}),
    i.e. without the balancing Hastings factor of the delayed attempts."""
    _convention = _lib.DR_PROB_METROP


class PyHMC_multistep_tries3(_ProbMultistep):
    """pyhmc.PyHMC_multistep_tries3 (src/pyhmc.py:362-483): p_try_j = 1 - alpha_j, the acceptance
    probability of the stage that just failed."""
    _convention = _lib.DR_PROB_HASTINGS
