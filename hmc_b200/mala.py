"""Batched drop-in for the reference's ``src/mala.py`` (MALA, :4-60).

``MALA(log_prob, grad_log_prob, invmetric_diag=None)`` keeps the reference's
constructor; ``step(q, eps, V_q=None, V_gq=None)`` returns ``(q2, accepted, V_q2,
V_gq2)`` with codes 1./0./2. and carries the cached potential and gradient exactly
as the reference does.  ``q`` may be ``[C, d]``.  Extra keywords: ``z=`` / ``u=``
inject the normal and uniform draws (parity mode); ``n_steps=K`` fuses K steps in
one kernel launch (the chain state stays in registers), returning the state after
the K-th step and ``accepted`` = code of the last step.
"""
import torch

from .engine import Batch, Engine


class MALA:
    def __init__(self, log_prob, grad_log_prob=None, invmetric_diag=None, **engine_kw):
        # the reference stores invmetric_diag but its step never uses it (src/mala.py:45-60)
        self.engine = Engine(log_prob, grad_log_prob, None, None, **engine_kw)
        self.model = self.engine.model
        self.log_prob, self.grad_log_prob = self.model.log_prob, self.model.grad_log_prob
        self.invmetric_diag = 1.0 if invmetric_diag is None else invmetric_diag
        self.metricstd = self.invmetric_diag ** -0.5
        self.leapcount, self.Vgcount, self.Hcount = 0, 0, 0
        self.n_accept = None

    def V(self, x):
        b = Batch(x, self.engine.dev, self.engine.dtype)
        logp, _ = self.engine.logp_grad(b.q)
        return b.out(-logp)

    def V_g(self, x):
        b = Batch(x, self.engine.dev, self.engine.dtype)
        self.Vgcount += 1
        _, g = self.engine.logp_grad(b.q)
        return b.out(-g)

    def step(self, q, eps, V_q=None, V_gq=None, z=None, u=None, n_steps=1):
        e = self.engine
        b = Batch(q, e.dev, e.dtype)
        e.check_dim(b)
        vdt = torch.float64 if e.is_logistic else e.dtype      # the logistic path carries V(q) in float64 (hmcb_logreg_mala_step)
        V = None if V_q is None else torch.as_tensor(V_q).reshape(-1).to(device=e.device, dtype=vdt).contiguous()
        Vg = None if V_gq is None else Batch(V_gq, e.device, e.dtype).q
        out = e.mala_step(b.q, eps, n_steps=n_steps, V=V, Vg=Vg, z=z, u=u)
        self.n_accept = b.out(out["n_accept"])
        code = out["code"].to(torch.float64)
        if b.unbatched:
            self.Hcount, self.Vgcount, self.leapcount = out["counts"][0].tolist()
            return b.out(out["q"]), float(code[0]), float(out["V"][0]), b.out(out["Vg"])
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return b.out(out["q"]), b.out(code), b.out(out["V"]), b.out(out["Vg"])
