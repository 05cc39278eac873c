"""Batched drop-in for the reference's ``src/algorithms.py`` on the hot path:
``Sampler`` (:3-30), ``DualAveragingStepSize`` (:34-75), ``HMC`` (:79-254),
``DRHMC`` (:467-555) and the ``alg_dict`` registry (:823-832).

Constructors keep the reference's positional order
``Cls(log_prob, grad_log_prob=None, log_prob_and_grad=None, invmetric_diag=None)``
and also accept ``Cls(model)``.  ``step`` / ``sample`` keep their names and
keyword arguments (``nsamples, burnin, step_size, nleap, [nprops, stepfactor],
[epsadapt]``); ``q`` may be ``[C, d]`` and every per-chain output gains a
leading C axis.  Accept codes follow the reference per class: HMC 1/0/-1,
DRHMC j+1/0.  The debug prints of the reference (:518-521,:547) are not
reproduced.
"""
import numpy as np
import torch

from . import _lib, dist
from .engine import Batch
from .pyhmc import HMC as _PyHMCBase


class Sampler:
    """Result container of ``sample`` (src/algorithms.py:3-30): lists while
    sampling, stacked arrays after ``to_array``."""

    def __init__(self):
        self.samples = []
        self.accepts = []
        self.Hs = []
        self.counts = []
        self.i = 0

    @staticmethod
    def _stack(v):
        if len(v) and torch.is_tensor(v[0]):
            return torch.stack(v)
        return np.array(v)

    def to_array(self):
        for key, val in self.__dict__.items():
            if isinstance(val, list):
                self.__dict__[key] = self._stack(val)

    def to_list(self):
        for key, val in self.__dict__.items():
            if isinstance(val, np.ndarray) or torch.is_tensor(val):
                self.__dict__[key] = list(val)

    def appends(self, q, acc, Hs, count):
        self.i += 1
        self.accepts.append(acc)
        self.samples.append(q)
        self.Hs.append(Hs)
        self.counts.append(count)

    def save(self, path):
        pass


class DualAveragingStepSize:
    """Nesterov dual averaging of the log step size (src/algorithms.py:34-75).
    Scalar host arithmetic; with many chains the caller feeds it the chain-mean
    acceptance statistic (all-reduced over ranks), see HMC.adapt_stepsize."""

    def __init__(self, initial_step_size, target_accept=0.65, gamma=0.05, t0=10.0, kappa=0.75, nadapt=0):
        self.initial_step_size = initial_step_size
        self.mu = np.log(10 * initial_step_size)  # shrinkage point, biased upwards
        self.target_accept = target_accept
        self.gamma = gamma
        self.t = t0
        self.kappa = kappa
        self.error_sum = 0
        self.log_averaged_step = 0
        self.nadapt = nadapt

    def update(self, p_accept):
        """Feed one acceptance statistic; returns (noisy step size, averaged step size)."""
        if np.isnan(p_accept) or np.isinf(p_accept):
            p_accept = 0.0
        if p_accept > 1:
            p_accept = 1.0
        self.error_sum += self.target_accept - p_accept
        log_step = self.mu - self.error_sum / (np.sqrt(self.t) * self.gamma)
        eta = self.t ** -self.kappa
        self.log_averaged_step = eta * log_step + (1 - eta) * self.log_averaged_step
        self.t += 1
        return np.exp(log_step), np.exp(self.log_averaged_step)

    def __call__(self, i, p_accept):
        if i == 0:
            return self.initial_step_size
        if i < self.nadapt:
            step_size, _ = self.update(p_accept)
        elif i == self.nadapt:
            _, step_size = self.update(p_accept)
        else:
            step_size = np.exp(self.log_averaged_step)
        return step_size


class HMC(_PyHMCBase):
    """algorithms.HMC (src/algorithms.py:79-254): codes 1/0/-1, the uniform is
    always consumed (:184)."""

    _rule = _lib.RULE_ALG

    def __init__(self, log_prob, grad_log_prob=None, log_prob_and_grad=None, invmetric_diag=None, **engine_kw):
        super().__init__(log_prob, grad_log_prob, invmetric_diag, log_prob_and_grad=log_prob_and_grad, **engine_kw)
        self.log_prob_and_grad = self.model.log_prob_and_grad

    def step(self, q, nleap, step_size, p=None, u=None, **kwargs):
        """One transition per chain (src/algorithms.py:192-198)."""
        return self._step(q, nleap, step_size, p=p, u=u)

    def V_vandg(self, x):
        """(V, grad V) from one evaluation (src/algorithms.py:109-115)."""
        b = self._device_q(x)
        self.Vgcount += 1
        logp, g = self.engine.logp_grad(b.q)
        return b.out(-logp), b.out(-g)

    def leapfrog_Vgq(self, q, p, N, step_size, V_q=None, V_gq=None):
        """Leapfrog trajectory that may reuse a cached gradient for its first half kick and returns the
        potential and gradient at its end point for the next one (src/algorithms.py:150-172):
        (q1, p1, V(q1), grad V(q1)).  The fused kernel always evaluates its first gradient itself, so a
        supplied V_gq enters as the difference to it (identical when V_gq is grad V(q), as in every use
        the reference makes of it); the counters follow the reference (N gradients with a cache, N+1
        without)."""
        self.leapcount += 1
        e = self.engine
        bq = self._device_q(q)
        bp = Batch(p, e.dev, e.dtype)
        pd = bp.q
        n = max(int(N), 1)
        if V_gq is not None:
            g = Batch(V_gq, e.dev, e.dtype).q
            _, gl = e.logp_grad(bq.q)
            pd = pd - 0.5 * step_size * g - 0.5 * step_size * gl     # gl = -grad V(q)
            self.Vgcount += n
        else:
            self.Vgcount += n + 1
        q1, p1 = e.leapfrog(bq.q, pd, N, step_size)
        logp, g1 = e.logp_grad(q1)
        return bq.out(q1), bq.out(p1), bq.out(-logp), bq.out(-g1)

    def _sample_fused(self, b, state, run):
        """store=False form of sample(): burn-in and sampling as one fused call each."""
        qd = b.q
        if self.burnin > 0:
            qd = run(qd, self.burnin)["q"]
        C_, d = qd.shape
        msum, msq = torch.zeros_like(qd), torch.zeros_like(qd)
        nacc = torch.zeros(C_, dtype=torch.int32, device=qd.device)
        out = run(qd, self.nsamples, msum=msum, msumsq=msq, n_accept=nacc)
        state.i = self.nsamples + self.burnin
        state.q, state.n_accept, state.counts = b.out(out["q"]), b.out(nacc), b.out(out["counts"])
        state.moment_sum, state.moment_sumsq = b.out(msum), b.out(msq)
        state.samples = state.accepts = state.Hs = None
        return state

    def _parse_kwargs_sample(self, **kwargs):
        self.nsamples = kwargs["nsamples"]
        self.burnin = kwargs["burnin"]
        self.step_size = kwargs["step_size"]
        self.nleap = kwargs["nleap"]

    def _device_q(self, q):
        b = Batch(q, self.engine.dev, self.engine.dtype)
        self.engine.check_dim(b)
        return b

    def adapt_stepsize(self, q, epsadapt, p=None, u=None, **kwargs):
        """Dual-averaging warm-up (src/algorithms.py:209-230): epsadapt+1 transitions,
        each feeding the update with prob = exp(H0-H1) clipped to [0,1], or 0 for a
        chain that did not move.  With C chains (and R ranks) the statistic is the
        mean over all chains -- one shared step size, all-reduced over NCCL; with one
        chain it is the reference's update.  p [epsadapt+1, C, d] / u [epsadapt+1, C, 1]
        replay the draws of each iteration (parity runs)."""
        step_size = self.step_size
        da = DualAveragingStepSize(step_size)
        self._parse_kwargs_sample(**kwargs)
        b = self._device_q(q)
        qd = b.q
        self.adapt_trace = []
        for i in range(epsadapt + 1):
            self.adapt_trace.append(step_size)
            out = self.engine.hmc_step(qd, self.nleap, step_size, self._rule, want_p=False, want_prob=True,
                                       p=None if p is None else p[i], u=None if u is None else u[i])
            qd = out["q"]
            stat = torch.stack([out["prob"].sum(dtype=torch.float64),
                                torch.tensor(float(qd.shape[0]), dtype=torch.float64, device=qd.device)])
            stat = dist.all_reduce_sum(stat)
            prob = float(stat[0] / stat[1])
            if i < epsadapt:
                step_size, _ = da.update(prob)
            else:
                _, step_size = da.update(prob)
                self.step_size = step_size
        return b.out(qd)

    def sample(self, q, p=None, callback=None, skipburn=True, epsadapt=0, draws=None, store=True, **kwargs):
        """src/algorithms.py:233-254.  ``accepts`` is kept for every iteration, samples /
        Hs / counts only for i > burnin (strict, as the reference): nsamples-1 rows.
        ``p`` is accepted and ignored, as in the reference.  draws=(P, U), P [T, C, d] and
        U [T, C, 1] with T = (epsadapt+1 if epsadapt else 0) + nsamples + burnin, replays the
        draws in the order the reference would consume them (parity runs).

        store=False (not in the reference; for batches whose history does not fit): the same
        transitions through ONE hmcb_hmc_sample call per phase (burn-in, then sampling) with the
        chain state on the device; the Sampler then carries, instead of per-iteration lists,
        ``q`` (final state), ``n_accept`` [C] and ``counts`` [C,3] over the nsamples kept
        transitions, and the running moments ``moment_sum`` / ``moment_sumsq`` [C,d] of those."""
        self._parse_kwargs_sample(**kwargs)
        state = Sampler()
        P, U = (None, None) if draws is None else draws
        t0 = 0
        if epsadapt:
            t0 = epsadapt + 1
            q = self.adapt_stepsize(q, epsadapt, p=None if P is None else P[:t0], u=None if U is None else U[:t0],
                                    **kwargs)
        b = self._device_q(q)
        qd = b.q
        if not store:
            if draws is not None:
                raise ValueError("replayed draws need store=True (one transition per call)")
            return self._sample_fused(b, state, lambda qq, n, **kw: self.engine.hmc_sample(
                qq, self.nleap, self.step_size, self._rule, n, **kw))
        for i in range(self.nsamples + self.burnin):
            out = self.engine.hmc_step(qd, self.nleap, self.step_size, self._rule, want_p=False,
                                       p=None if P is None else P[t0 + i], u=None if U is None else U[t0 + i])
            qd = out["q"]
            state.i += 1
            state.accepts.append(b.out(out["code"].to(torch.float64)))
            if skipburn & (i > self.burnin):
                state.samples.append(b.out(qd))
                state.Hs.append(b.out(out["Hs"]))
                state.counts.append(b.out(out["counts"]))
                if callback is not None:
                    callback(state)
        state.to_array()
        return state


class DRHMC(HMC):
    """Delayed-rejection HMC, algorithms.DRHMC (src/algorithms.py:467-555).  One
    ``step`` is hmcb_drhmc_step: stage kernels over device-compacted lists of the
    chains still rejecting, ghost recursion (get_num, :474-497) included."""

    def step(self, q, nleap, step_size, nprops, stepfactor, p=None, u=None, want_p=True):
        """Returns (q, p, code, avec, [Hcount, Vgcount, leapcount]); code = j+1 if
        accepted at stage j, else 0 (src/algorithms.py:500-530).  want_p=False returns None
        for p and skips its device write and device-to-host copy: the reference's own sample
        loop throws the returned momentum away (src/algorithms.py:546-551)."""
        e = self.engine
        if torch.is_tensor(q) and q.device.type == "cpu" and q.dim() == 2 and p is None and u is None \
                and q.shape[0] >= 8192:
            # host-resident batch: chunked copy/compute pipeline (same results, see Engine.drhmc_step_host)
            e._ensure()
            if q.shape[1] != self.model.d:
                raise ValueError("q has dimension %d but the model has d=%d" % (q.shape[1], self.model.d))
            h = e.drhmc_step_host(q.to(e.dtype), nleap, step_size, nprops, stepfactor, want_p=want_p)
            self.Hcount, self.Vgcount, self.leapcount = h["counts"][:, 0], h["counts"][:, 1], h["counts"][:, 2]
            return h["q"], h["p"], h["code"], h["avec"], h["counts"]
        b = self._device_q(q)
        out = e.drhmc_step(b.q, nleap, step_size, nprops, stepfactor, p=p, u=u, want_p=want_p)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            return b.out(out["q"]), b.out(out["p"]), int(out["code"][0]), b.out(out["avec"]), cnt
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return b.out(out["q"]), b.out(out["p"]), b.out(out["code"]), b.out(out["avec"]), b.out(out["counts"])

    def sample(self, q, p=None, callback=None, draws=None, store=True, **kwargs):
        """src/algorithms.py:534-555: every iteration (burn-in included) is stored.  (The reference draws
        one unused normal vector when ``p`` is None, :538; nothing is drawn for it here.)  draws=(P, U),
        P [T, C, d] and U [T, C, nprops] with T = nsamples + burnin, replays the draws (parity runs).
        store=False: as HMC.sample(store=False), through hmcb_drhmc_sample (burn-in, then nsamples
        transitions with running moments); the kept statistics cover the nsamples transitions."""
        self._parse_kwargs_sample(**kwargs)
        state = Sampler()
        state.avecs = []
        b = self._device_q(q)
        qd = b.q
        if not store:
            if draws is not None:
                raise ValueError("replayed draws need store=True (one transition per call)")
            return self._sample_fused(b, state, lambda qq, n, **kw: self.engine.drhmc_sample(
                qq, self.nleap, self.step_size, kwargs["nprops"], kwargs["stepfactor"], n, want_avec=False, **kw))
        P, U = (None, None) if draws is None else draws
        for i in range(self.nsamples + self.burnin):
            out = self.engine.drhmc_step(qd, self.nleap, self.step_size, kwargs["nprops"], kwargs["stepfactor"],
                                         want_p=False, p=None if P is None else P[i], u=None if U is None else U[i])
            qd = out["q"]
            state.i += 1
            state.samples.append(b.out(qd))
            state.accepts.append(b.out(out["code"]))
            state.avecs.append(b.out(out["avec"]))
            state.counts.append(b.out(out["counts"]))
            if callback is not None:
                callback(state)
        state.to_array()
        return state


class GHMC(HMC):
    """Generalised HMC, algorithms.GHMC (src/algorithms.py:259-303): the momentum is only partly
    refreshed, p <- alpha p + sqrt(1-alpha^2) pn, and comes back flipped unless the proposal is
    accepted; V(q) and grad V(q) are carried between transitions.  Work counters follow the
    reference built as GHMC(log_prob, grad_log_prob, log_prob_and_grad)."""

    def _as_dev(self, x, b):
        if x is None:
            return None
        t = torch.as_tensor(np.asarray(x)) if not torch.is_tensor(x) else x
        return t.reshape(b.C, -1).to(device=self.engine.device, dtype=self.engine.dtype).contiguous()

    def step(self, q, p, nleap, step_size, alpha, V_q=None, V_gq=None, pn=None, u=None):
        """Returns (q, p, V_q, V_gq, accepted, Hs, counts) (src/algorithms.py:265-281)."""
        e = self.engine
        b = self._device_q(q)
        V = None if V_q is None else torch.as_tensor(np.asarray(V_q) if not torch.is_tensor(V_q) else V_q).reshape(-1).to(
            device=e.device, dtype=e.dtype).contiguous()
        out = e.ghmc_step(b.q, self._as_dev(p, b), nleap, step_size, alpha, V=V, Vg=self._as_dev(V_gq, b), pn=pn, u=u)
        code = out["code"].to(torch.float64)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            return (b.out(out["q"]), b.out(out["p"]), float(out["V"][0]), b.out(out["Vg"]), float(code[0]),
                    out["Hs"][0].tolist(), cnt)
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return (b.out(out["q"]), b.out(out["p"]), b.out(out["V"]), b.out(out["Vg"]), b.out(code), b.out(out["Hs"]),
                b.out(out["counts"]))

    def _sample_fused_ghmc(self, b, state, pd, alpha, pu=None, delta=0.0):
        """store=False form of sample(): burn-in and sampling as one hmcb_ghmc_sample call each; the state keeps the
        final q, p (and u), the accept counts and summed work counters of the nsamples kept transitions."""
        e = self.engine
        qd, V, Vg = b.q, None, None
        for K, keep in ((self.burnin, False), (self.nsamples, True)):
            if K < 1:
                continue
            out = e.ghmc_sample(qd, pd, self.nleap, self.step_size, alpha, K, V=V, Vg=Vg, pu=pu, delta=delta)
            qd, pd, V, Vg, pu = out["q"], out["p"], out["V"], out["Vg"], out["u"]
            if keep:
                state.n_accept, state.counts = b.out(out["n_accept"]), b.out(out["counts"])
        state.i = self.nsamples + self.burnin
        state.q, state.p = b.out(qd), b.out(pd)
        if pu is not None:
            state.u = b.out(pu)
        state.samples = state.accepts = state.Hs = None
        return state

    def sample(self, q, p=None, callback=None, store=True, draws=None, **kwargs):
        """src/algorithms.py:284-303: every iteration is stored; p starts as a fresh normal draw.  store=False keeps no
        history and runs each phase as one fused call (see _sample_fused_ghmc).  draws = (pn [T,C,d], u [T,C,1]) replays
        the refresh normals and Metropolis uniforms of the T = nsamples + burnin iterations (parity runs)."""
        self._parse_kwargs_sample(**kwargs)
        state = Sampler()
        b = self._device_q(q)
        qd = b.q
        pd = torch.randn(qd.shape, dtype=qd.dtype, device=qd.device) if p is None else self._as_dev(p, b)
        if not store:
            return self._sample_fused_ghmc(b, state, pd, kwargs["alpha"])
        V, Vg = None, None
        for i in range(self.nsamples + self.burnin):
            out = self.engine.ghmc_step(qd, pd, self.nleap, self.step_size, kwargs["alpha"], V=V, Vg=Vg,
                                        pn=None if draws is None else draws[0][i], u=None if draws is None else draws[1][i])
            qd, pd, V, Vg = out["q"], out["p"], out["V"], out["Vg"]
            state.i += 1
            state.samples.append(b.out(qd))
            state.accepts.append(b.out(out["code"].to(torch.float64)))
            state.Hs.append(b.out(out["Hs"]))
            state.counts.append(b.out(out["counts"]))
            if callback is not None:
                callback(state)
        state.to_array()
        return state


class Persistent_GHMC(GHMC):
    """algorithms.Persistent_GHMC (src/algorithms.py:307-357): GHMC whose Metropolis uniform u is a
    persistent variable, u <- (u + 1 + delta) % 2 - 1 each transition, divided by exp(H0-H1) on
    acceptance.  (As in the reference, u itself -- not |u| -- is compared with min(1, prob).)"""

    def step(self, q, p, u, nleap, step_size, alpha, delta, V_q=None, V_gq=None, pn=None):
        """Returns (q, p, u, V_q, V_gq, accepted, Hs, counts) (src/algorithms.py:314-333)."""
        e = self.engine
        b = self._device_q(q)
        V = None if V_q is None else torch.as_tensor(np.asarray(V_q) if not torch.is_tensor(V_q) else V_q).reshape(-1).to(
            device=e.device, dtype=e.dtype).contiguous()
        ut = torch.as_tensor(np.asarray(u, dtype=np.float64) if not torch.is_tensor(u) else u, dtype=torch.float64)
        ut = ut.reshape(-1).to(e.device).contiguous()
        if ut.numel() == 1 and b.C > 1:
            ut = ut.expand(b.C).contiguous()
        out = e.ghmc_step(b.q, self._as_dev(p, b), nleap, step_size, alpha, V=V, Vg=self._as_dev(V_gq, b), pn=pn, pu=ut,
                          delta=delta)
        code = out["code"].to(torch.float64)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            return (b.out(out["q"]), b.out(out["p"]), float(out["u"][0]), float(out["V"][0]), b.out(out["Vg"]),
                    float(code[0]), out["Hs"][0].tolist(), cnt)
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return (b.out(out["q"]), b.out(out["p"]), b.out(out["u"]), b.out(out["V"]), b.out(out["Vg"]), b.out(code),
                b.out(out["Hs"]), b.out(out["counts"]))

    def sample(self, q, p=None, callback=None, store=True, u=None, draws=None, **kwargs):
        """src/algorithms.py:336-357.  store=False: no history, each phase one fused call; u [C] sets the initial
        persistent variable (default: a normal draw per chain, as the reference draws one); draws = pn [T,C,d] replays
        the refresh normals (parity runs)."""
        self._parse_kwargs_sample(**kwargs)
        state = Sampler()
        state.u = []
        b = self._device_q(q)
        qd = b.q
        pd = torch.randn(qd.shape, dtype=qd.dtype, device=qd.device) if p is None else self._as_dev(p, b)
        if u is None:
            ud = torch.randn(qd.shape[0], dtype=torch.float64, device=qd.device)
        else:
            ud = torch.as_tensor(np.asarray(u, dtype=np.float64) if not torch.is_tensor(u) else u, dtype=torch.float64)
            ud = ud.reshape(-1).to(qd.device).contiguous()
        if not store:
            return self._sample_fused_ghmc(b, state, pd, kwargs["alpha"], pu=ud, delta=kwargs["delta"])
        V, Vg = None, None
        for i in range(self.nsamples + self.burnin):
            out = self.engine.ghmc_step(qd, pd, self.nleap, self.step_size, kwargs["alpha"], V=V, Vg=Vg, pu=ud,
                                        delta=kwargs["delta"], pn=None if draws is None else draws[i])
            qd, pd, ud, V, Vg = out["q"], out["p"], out["u"], out["V"], out["Vg"]
            state.i += 1
            state.samples.append(b.out(qd))
            state.accepts.append(b.out(out["code"].to(torch.float64)))
            state.Hs.append(b.out(out["Hs"]))
            state.counts.append(b.out(out["counts"]))
            state.u.append(b.out(ud))
            if callback is not None:
                callback(state)
        state.to_array()
        return state


class DRGHMC(DRHMC):
    """algorithms.DRGHMC (src/algorithms.py:561-646): delayed rejection started from the partially
    refreshed momentum p0 = alpha p + sqrt(1-alpha^2) pn; p0 itself is returned on rejection."""

    def step(self, q, p, nleap, step_size, alpha, nprops, stepfactor, pn=None, u=None):
        """Returns (q, p, code, avec, counts) (src/algorithms.py:596-625)."""
        e = self.engine
        b = self._device_q(q)
        pv = torch.as_tensor(np.asarray(p)) if not torch.is_tensor(p) else p
        pv = pv.reshape(b.C, -1).to(device=e.device, dtype=e.dtype).contiguous()
        out = e.drhmc_step(b.q, nleap, step_size, nprops, stepfactor, p=pn, u=u, p_prev=pv, alpha=alpha)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            return b.out(out["q"]), b.out(out["p"]), int(out["code"][0]), b.out(out["avec"]), cnt
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return b.out(out["q"]), b.out(out["p"]), b.out(out["code"]), b.out(out["avec"]), b.out(out["counts"])

    def sample(self, q, p=None, callback=None, **kwargs):
        """src/algorithms.py:629-646."""
        self._parse_kwargs_sample(**kwargs)
        state = Sampler()
        state.avecs = []
        b = self._device_q(q)
        qd = b.q
        pd = torch.randn(qd.shape, dtype=qd.dtype, device=qd.device) if p is None else torch.as_tensor(p).reshape(
            qd.shape).to(device=qd.device, dtype=qd.dtype)
        for i in range(self.nsamples + self.burnin):
            out = self.engine.drhmc_step(qd, self.nleap, self.step_size, kwargs["nprops"], kwargs["stepfactor"],
                                         p_prev=pd, alpha=kwargs["alpha"])
            qd, pd = out["q"], out["p"]
            state.i += 1
            state.samples.append(b.out(qd))
            state.accepts.append(b.out(out["code"]))
            state.avecs.append(b.out(out["avec"]))
            state.counts.append(b.out(out["counts"]))
            if callback is not None:
                callback(state)
        state.to_array()
        return state


def dr_stage_lengths(nleap, nprops, stepfactor):
    """(L_j, eps scale 1/f^j) per stage as the reference computes them
    (src/algorithms.py:510-511): int() truncation of nleap*f^j with a float f."""
    return [int(nleap * stepfactor ** j) for j in range(nprops)], [1.0 / stepfactor ** j for j in range(nprops)]


class _ProbDRHMC(HMC):
    """Probabilistic delayed rejection in src/algorithms.py: the classes add get_num and
    multi_step to algorithms.HMC and inherit its step / sample unchanged."""
    _convention = None

    def multi_step(self, m, q0, N, ss, stepfactor, p=None, u=None):
        """Returns (q, p, code, avec[m], [Hcount, Vgcount, leapcount], ptries[m]); code = accepting
        stage j, -1 (rejected at every stage) or -2 (a stage was declined).  u, if given, holds
        2m-1 uniforms per chain in consumption order: u_acc0, u_try1, u_acc1, ..."""
        e = self.engine
        b = self._device_q(q0)
        out = e.drhmc_step(b.q, N, ss, m, stepfactor, p=p, u=u, convention=self._convention)
        if b.unbatched:
            cnt = out["counts"][0].tolist()
            self.Hcount, self.Vgcount, self.leapcount = cnt
            return (b.out(out["q"]), b.out(out["p"]), int(out["code"][0]), b.out(out["avec"]), cnt,
                    b.out(out["ptries"]))
        self.Hcount, self.Vgcount, self.leapcount = out["counts"][:, 0], out["counts"][:, 1], out["counts"][:, 2]
        return (b.out(out["q"]), b.out(out["p"]), b.out(out["code"]), b.out(out["avec"]), b.out(out["counts"]),
                b.out(out["ptries"]))


class Prob_DRHMC_Metrop(_ProbDRHMC):
    """algorithms.Prob_DRHMC_Metrop (src/algorithms.py:649-728): p_try_j = 1 - min(1, pi(prop_j)/pi(init))."""
    _convention = _lib.DR_PROB_METROP


class Prob_DRHMC_Hastings(_ProbDRHMC):
    """algorithms.Prob_DRHMC_Hastings (src/algorithms.py:734-818): p_try_j = 1 - alpha_j.  The reference's
    constructor takes (log_prob, grad_log_prob, invmetric_diag) only (:738-739)."""
    _convention = _lib.DR_PROB_HASTINGS

    def __init__(self, log_prob, grad_log_prob=None, invmetric_diag=None, **engine_kw):
        super().__init__(log_prob, grad_log_prob, invmetric_diag=invmetric_diag, **engine_kw)


# name -> class registry (src/algorithms.py:823-832), the reference's keys
alg_dict = {"hmc": HMC, "ghmc": GHMC, "pghmc": Persistent_GHMC, "drhmc": DRHMC, "drghmc": DRGHMC,
            "pdrhmc_metrop": Prob_DRHMC_Metrop, "pdrhmc_hasting": Prob_DRHMC_Hastings}

