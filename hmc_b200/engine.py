"""Host-side plumbing shared by the sampler classes: argument normalisation
(numpy/torch, 1-D/batched), device buffers, and the calls into libhmcb.so.

Chain state at this API is row-major ``[C, d]`` (the reference's batched shape
precedent, src/pyhmc.py:604-684 and src/utils.py:232-235); kernels address it
through (stride_chain, stride_dim), so no transposes are made.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .models import as_model


def _require_cuda(device):
    if not torch.cuda.is_available():
        raise _lib.HmcbError("hmc_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.type != "cuda":
        raise _lib.HmcbError("hmc_b200 samplers run on CUDA devices only, got %r" % (device,))
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    return dev


class Batch:
    """Remembers how the caller passed ``q`` so results go back the same way:
    numpy in -> numpy out; CPU torch tensor in (pinned or not) -> pinned CPU torch
    tensors out; CUDA tensor in -> CUDA tensors out (no copies)."""

    def __init__(self, q, device, dtype):
        self.was_numpy = isinstance(q, np.ndarray) or not torch.is_tensor(q)
        t = torch.as_tensor(np.asarray(q)) if self.was_numpy else q
        self.host = (not self.was_numpy) and t.device.type == "cpu"
        self.unbatched = t.dim() == 1
        if self.unbatched:
            t = t.unsqueeze(0)
        if t.dim() != 2:
            raise ValueError("q must be [d] or [C, d]")
        self.q = t.to(device=device, dtype=dtype, non_blocking=True).contiguous()
        self.C, self.d = self.q.shape

    def out(self, t, per_chain=True):
        """Give a per-chain device tensor back in the caller's convention."""
        if t is None:
            return None
        if self.unbatched and per_chain:
            t = t[0]
        if self.was_numpy:
            return t.cpu().numpy()
        if self.host:
            # pinned staging from torch's caching host allocator; copy_ synchronises
            buf = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            buf.copy_(t)
            return buf
        return t


class Engine:
    """One model on one device: owns the hmcb_model block, the RNG position and
    scratch buffers, and wraps each C-ABI entry point."""

    def __init__(self, log_prob, grad_log_prob=None, log_prob_and_grad=None, invmetric_diag=None, device=None,
                 dtype=None, seed=0, chain_offset=0, rng_f64=False):
        self.model = as_model(log_prob, grad_log_prob, log_prob_and_grad)
        self.is_logistic = self.model.kind == _lib.MODEL_LOGISTIC
        if dtype is None:   # the reference computes in float64; the tensor-core logistic gradient is float32
            dtype = torch.float32 if self.is_logistic else torch.float64
        if self.is_logistic and dtype != torch.float32:
            raise ValueError("LogisticRegression runs in float32 (tensor-core gradient): pass dtype=torch.float32")
        if dtype not in (torch.float64, torch.float32):
            raise ValueError("dtype must be torch.float64 or torch.float32")
        self.dtype = dtype
        self.invmetric_diag = invmetric_diag
        self._device_arg = device
        self.seed = int(seed)
        self.chain_offset = int(chain_offset)
        self.iteration = 0
        self.rng_flags = _lib.RNG_NORMAL_F64 if rng_f64 else 0
        self._struct = None
        self._keep = None
        self._ws = None          # adhmc scratch
        self._ws_pool = {}       # drhmc scratch, one per concurrent stream slot
        self._streams = None
        self._host_stage = None  # device staging of drhmc_step_host, kept between calls
        self._host_out = None    # two alternating sets of pinned result buffers of drhmc_step_host
        self.device = None

    # ------------------------------------------------------------------ set-up
    def _ensure(self):
        if self._struct is None:
            self.device = _require_cuda(self._device_arg)
            self.lib = _lib.load()
            self._struct, self._keep = self.model.to_struct(self.device, self.dtype, self.invmetric_diag)
            if self.is_logistic:
                self._prepare_logistic()
        return self._struct

    def _prepare_logistic(self):
        """Build the fp16 operand blocks of X once (hmcb_logreg_prepare)."""
        m = self._struct
        with torch.cuda.device(self.device):
            need = self.lib.hmcb_logreg_prepared_bytes(C.byref(m))
            if need == 0:
                _lib.check(_lib.E_ARG)
            self._lr_prepared = torch.empty(need, dtype=torch.uint8, device=self.device)
            _lib.check(self.lib.hmcb_logreg_prepare(C.byref(m), C.c_void_p(self._lr_prepared.data_ptr()), need,
                                                    self._stream()))

    def _lr_workspace(self, n_chains):
        need = self.lib.hmcb_logreg_workspace_bytes(C.byref(self._struct), int(n_chains))
        if need == 0:
            _lib.check(_lib.E_ARG)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._ws

    @property
    def dev(self):
        self._ensure()
        return self.device

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _chains(self, q_in, q_out, extra_offset=0):
        c = _lib.Chains()
        c.n_chains, c.chain_offset = q_in.shape[0], self.chain_offset + int(extra_offset)
        c.q_in, c.q_out = q_in.data_ptr(), q_out.data_ptr()
        c.stride_chain, c.stride_dim = q_in.stride(0), q_in.stride(1)
        return c

    def _rng(self, p, u, nu, C_, p_prev=None, alpha=0.0):
        r = _lib.Rng()
        r.seed, r.iteration, r.flags = self.seed, self.iteration, self.rng_flags
        keep = []
        if p_prev is not None:
            pv = p_prev.reshape(C_, -1).to(device=self.device, dtype=self.dtype).contiguous()
            keep.append(pv)
            r.p_prev, r.alpha = pv.data_ptr(), float(alpha)
        if p is not None:
            pt = torch.as_tensor(np.asarray(p)) if not torch.is_tensor(p) else p
            pt = pt.reshape(C_, -1).to(device=self.device, dtype=self.dtype).contiguous()
            keep.append(pt)
            r.p_inject = pt.data_ptr()
        if u is not None:
            ut = torch.as_tensor(np.asarray(u, dtype=np.float64)) if not torch.is_tensor(u) else u
            ut = ut.reshape(C_, -1).to(device=self.device, dtype=torch.float64).contiguous()
            if ut.shape[1] < nu:
                raise ValueError("u must hold %d uniforms per chain" % nu)
            keep.append(ut)
            r.u_inject, r.u_stride = ut.data_ptr(), ut.shape[1]
        return r, keep

    def check_dim(self, b):
        if b.d != self.model.d:
            raise ValueError("q has dimension %d but the model has d=%d" % (b.d, self.model.d))

    # ------------------------------------------------------------------ calls
    def hmc_step(self, q, n_leap, step_size, rule, p=None, u=None, want_p=True, want_prob=False, q_out=None,
                 want_p0=False):
        """hmcb_hmc_step on device tensor q [C,d]; returns dict of device tensors.  want_p0 also returns
        the momentum every chain started from (hmcb_step_out.p0_out)."""
        m = self._ensure()
        C_ = q.shape[0]
        with torch.cuda.device(self.device):
            q_out = torch.empty_like(q) if q_out is None else q_out
            out = {"q": q_out,
                   "p": torch.empty_like(q) if want_p else None,
                   "p0": torch.empty_like(q) if want_p0 else None,
                   "code": torch.empty(C_, dtype=torch.int32, device=self.device),
                   "Hs": torch.empty((C_, 2), dtype=self.dtype, device=self.device),
                   "counts": torch.empty((C_, 3), dtype=torch.int32, device=self.device),
                   "prob": torch.empty(C_, dtype=self.dtype, device=self.device) if want_prob else None}
            so = _lib.StepOut()
            so.p_out, so.code, so.energies = _lib.ptr(out["p"]), _lib.ptr(out["code"]), _lib.ptr(out["Hs"])
            so.counts, so.accept_prob, so.p0_out = _lib.ptr(out["counts"]), _lib.ptr(out["prob"]), _lib.ptr(out["p0"])
            ch = self._chains(q, q_out)
            r, keep = self._rng(p, u, 1, C_)
            if self.is_logistic:
                ws = self._lr_workspace(C_)
                _lib.check(self.lib.hmcb_logreg_hmc_step(C.byref(m), C.c_void_p(self._lr_prepared.data_ptr()),
                                                         C.byref(ch), C.byref(r), int(n_leap), float(step_size),
                                                         int(rule), C.byref(so), C.c_void_p(ws.data_ptr()), ws.numel(),
                                                         self._stream()))
            else:
                _lib.check(self.lib.hmcb_hmc_step(C.byref(m), C.byref(ch), C.byref(r), int(n_leap), float(step_size),
                                                  int(rule), C.byref(so), self._stream()))
        self.iteration += 1
        out["_keep"] = keep
        return out

    def drhmc_step(self, q, n_leap, step_size, nprops, stepfactor, p=None, u=None, want_p=True, q_out=None,
                   stage_events=None, want_avec=True, code_out=None, counts_out=None, extra_offset=0, ws_slot=0,
                   advance=True, convention=0, p_prev=None, alpha=0.0, p_out=None, avec_out=None, want_p0=False):
        """hmcb_drhmc_step on device tensor q [C,d].  q_out / code_out / counts_out let a
        caller reuse its own output buffers (no allocation inside a timed loop)."""
        m = self._ensure()
        C_ = q.shape[0]
        with torch.cuda.device(self.device):
            need = self.lib.hmcb_drhmc_workspace_bytes(C.byref(m), C_, int(nprops))
            if need == 0:
                _lib.check(_lib.E_ARG)
            ws = self._ws_pool.get(ws_slot)
            if ws is None or ws.numel() < need:
                ws = self._ws_pool[ws_slot] = torch.empty(need, dtype=torch.uint8, device=self.device)
            q_out = torch.empty_like(q) if q_out is None else q_out
            out = {"q": q_out,
                   "p": p_out if p_out is not None else (torch.empty_like(q) if want_p else None),
                   "code": torch.empty(C_, dtype=torch.int32, device=self.device) if code_out is None else code_out,
                   "avec": avec_out if avec_out is not None else
                   (torch.empty((C_, nprops), dtype=self.dtype, device=self.device) if want_avec else None),
                   "counts": torch.empty((C_, 3), dtype=torch.int32, device=self.device)
                   if counts_out is None else counts_out}
            prob = convention in (_lib.DR_PROB_METROP, _lib.DR_PROB_HASTINGS)
            if prob:   # the probabilistic classes also return their try probabilities
                out["ptries"] = torch.empty((C_, nprops), dtype=self.dtype, device=self.device)
            so = _lib.StepOut()
            so.p_out, so.code, so.avec = _lib.ptr(out["p"]), _lib.ptr(out["code"]), _lib.ptr(out["avec"])
            so.counts, so.ptries = _lib.ptr(out["counts"]), _lib.ptr(out.get("ptries"))
            if want_p0:
                out["p0"] = torch.empty_like(q)
                so.p0_out = _lib.ptr(out["p0"])
            ch = self._chains(q, q_out, extra_offset)
            r, keep = self._rng(p, u, 2 * nprops - 1 if prob else nprops, C_, p_prev, alpha)
            _lib.check(self.lib.hmcb_drhmc_step(C.byref(m), C.byref(ch), C.byref(r), int(n_leap), float(step_size),
                                                int(nprops), float(stepfactor), int(convention), C.byref(so),
                                                C.c_void_p(ws.data_ptr()), ws.numel(),
                                                self._profile(stage_events), self._stream()))
        if advance:
            self.iteration += 1
        out["_keep"] = keep
        return out

    def hmc_sample(self, q, n_leap, step_size, rule, n_transitions, msum=None, msumsq=None, n_accept=None, q_out=None,
                   want_p=False):
        """hmcb_hmc_sample on device tensor q [C,d]: n_transitions HMC transitions of every chain in one call (one
        launch where the shape has the fused kernel).  Returns the state after the last transition, its code and
        energies, and the work counters of all transitions."""
        m = self._ensure()
        C_ = q.shape[0]
        with torch.cuda.device(self.device):
            q_out = torch.empty_like(q) if q_out is None else q_out
            out = {"q": q_out, "p": torch.empty_like(q) if want_p else None,
                   "code": torch.empty(C_, dtype=torch.int32, device=self.device),
                   "Hs": torch.empty((C_, 2), dtype=self.dtype, device=self.device),
                   "counts": torch.empty((C_, 3), dtype=torch.int32, device=self.device)}
            so = _lib.StepOut()
            so.p_out, so.code, so.energies, so.counts = (_lib.ptr(out["p"]), _lib.ptr(out["code"]), _lib.ptr(out["Hs"]),
                                                         _lib.ptr(out["counts"]))
            io = _lib.SampleIO()
            io.n_transitions = int(n_transitions)
            io.moment_sum, io.moment_sumsq, io.n_accept = _lib.ptr(msum), _lib.ptr(msumsq), _lib.ptr(n_accept)
            ch = self._chains(q, q_out)
            r, keep = self._rng(None, None, 1, C_)
            if self.is_logistic:
                raise _lib.HmcbError("hmc_sample is not available for the logistic-regression model")
            _lib.check(self.lib.hmcb_hmc_sample(C.byref(m), C.byref(ch), C.byref(r), int(n_leap), float(step_size), int(rule),
                                                C.byref(io), C.byref(so), self._stream()))
        self.iteration += int(n_transitions)
        return out

    def drhmc_sample(self, q, n_leap, step_size, nprops, stepfactor, n_transitions, msum=None, msumsq=None,
                     n_accept=None, q_out=None, want_p=False, convention=0, code_out=None, counts_out=None,
                     want_avec=True, p=None, u=None):
        """hmcb_drhmc_sample on device tensor q [C,d]: n_transitions DR transitions of every chain in one launch,
        running moments folded into msum / msumsq ([C,d], updated in place) when given.  Returns the state after the
        last transition, the last transition's code / avec, and the work counters summed over the transitions."""
        m = self._ensure()
        C_ = q.shape[0]
        with torch.cuda.device(self.device):
            need = self.lib.hmcb_drhmc_sample_workspace_bytes(C.byref(m), C_, int(nprops))
            if need == 0:
                _lib.check(_lib.E_ARG)
            ws = self._ws_pool.get("sample")
            if ws is None or ws.numel() < need:
                ws = self._ws_pool["sample"] = torch.empty(need, dtype=torch.uint8, device=self.device)
            q_out = torch.empty_like(q) if q_out is None else q_out
            out = {"q": q_out, "p": torch.empty_like(q) if want_p else None,
                   "code": torch.empty(C_, dtype=torch.int32, device=self.device) if code_out is None else code_out,
                   "avec": torch.empty((C_, nprops), dtype=self.dtype, device=self.device) if want_avec else None,
                   "counts": torch.empty((C_, 3), dtype=torch.int32, device=self.device)
                   if counts_out is None else counts_out}
            prob = convention in (_lib.DR_PROB_METROP, _lib.DR_PROB_HASTINGS)
            if prob:
                out["ptries"] = torch.empty((C_, nprops), dtype=self.dtype, device=self.device)
            so = _lib.StepOut()
            so.p_out, so.code, so.avec = _lib.ptr(out["p"]), _lib.ptr(out["code"]), _lib.ptr(out["avec"])
            so.counts, so.ptries = _lib.ptr(out["counts"]), _lib.ptr(out.get("ptries"))
            io = _lib.SampleIO()
            io.n_transitions = int(n_transitions)
            io.moment_sum, io.moment_sumsq, io.n_accept = _lib.ptr(msum), _lib.ptr(msumsq), _lib.ptr(n_accept)
            ch = self._chains(q, q_out)
            r, keep = self._rng(p, u, 2 * nprops - 1 if prob else nprops, C_)
            _lib.check(self.lib.hmcb_drhmc_sample(C.byref(m), C.byref(ch), C.byref(r), int(n_leap), float(step_size),
                                                  int(nprops), float(stepfactor), int(convention), C.byref(io),
                                                  C.byref(so), C.c_void_p(ws.data_ptr()), ws.numel(), self._stream()))
        self.iteration += int(n_transitions)
        out["_keep"] = keep
        return out

    def drhmc_step_host(self, q_host, n_leap, step_size, nprops, stepfactor, want_p=True, n_chunks=8, run_streams=2):
        """One DR transition for chain state living in HOST memory.  The batch is cut into
        chunks of chains that flow through a copy-in stream, two compute streams and a copy-out
        stream -- host-to-device copies, the stage kernels, device-to-host copies -- chained by
        events, so both copy engines and the SMs
        work at the same time and the inbound copies run back to back.  The first chunks grow from a
        quarter to full size and the last is half-size: the pipeline fills and drains on small pieces.  Chains are
        independent and draws are keyed by global chain id, so the result is bit-identical to
        the one-shot call.  q_host: CPU tensor [C,d] (pinned for full copy speed); returns
        pinned CPU tensors owned by the engine: two sets, used alternately, so a result stays valid
        until the call after the next one overwrites it (copy it if you keep it longer).  Device
        staging for the whole batch is kept between calls."""
        self._ensure()
        C_, d = q_host.shape
        n_chunks = max(1, min(int(n_chunks), (C_ + 4095) // 4096))
        # results land in one of TWO sets of pinned host buffers the engine keeps (alternating per call, so the previous
        # call's results stay valid while this one is produced -- e.g. q of call i is the input of call i+1): no pinned
        # allocation per call
        hkey = (C_, d, nprops, bool(want_p))
        if self._host_out is None or self._host_out[0] != hkey:
            pin = dict(pin_memory=True)
            self._host_out = (hkey, [{"q": torch.empty((C_, d), dtype=self.dtype, **pin),
                                      "p": torch.empty((C_, d), dtype=self.dtype, **pin) if want_p else None,
                                      "code": torch.empty(C_, dtype=torch.int32, **pin),
                                      "avec": torch.empty((C_, nprops), dtype=self.dtype, **pin),
                                      "counts": torch.empty((C_, 3), dtype=torch.int32, **pin)} for _ in range(2)], 0)
        hk, sets, flip = self._host_out
        if q_host.data_ptr() == sets[flip]["q"].data_ptr():
            flip ^= 1               # never write the results over the input
        host = sets[flip]
        self._host_out = (hk, sets, flip ^ 1)
        cur = torch.cuda.current_stream(self.device)
        with torch.cuda.device(self.device):
            if self._streams is None:
                self._streams = [torch.cuda.Stream(device=self.device) for _ in range(6)]
            key = (C_, d, nprops, bool(want_p))
            if self._host_stage is None or self._host_stage[0] != key:
                dev = dict(device=self.device)
                self._host_stage = (key, {
                    "q_in": torch.empty((C_, d), dtype=self.dtype, **dev),
                    "q": torch.empty((C_, d), dtype=self.dtype, **dev),
                    "p": torch.empty((C_, d), dtype=self.dtype, **dev) if want_p else None,
                    "code": torch.empty(C_, dtype=torch.int32, **dev),
                    "avec": torch.empty((C_, nprops), dtype=self.dtype, **dev),
                    "counts": torch.empty((C_, 3), dtype=torch.int32, **dev)})
        stage = self._host_stage[1]
        s_in, s_out = self._streams[0], self._streams[1]
        runs = self._streams[2:2 + max(1, min(4, int(run_streams)))]
        # chunk boundaries
        if n_chunks >= 4:
            # a rising head (the copy-out engine, the slower direction and the one that bounds the call, starts ~0.6 ms
            # sooner than behind one half-size chunk) and a half-size tail
            w = [0.25, 0.5, 0.75] + [1.0] * (n_chunks - 2) + [0.5]
        elif n_chunks == 3:
            w = [0.5, 1.0, 1.0, 0.5]
        else:
            w = [1.0] * n_chunks
        acc, bounds = 0.0, [0]
        for x in w:
            acc += x
            bounds.append(int(round(C_ * acc / sum(w))))
        bounds[-1] = C_
        for st in self._streams:
            st.wait_stream(cur)
        # host_timeline (debug): set engine.host_timeline = [] before the call and every chunk appends
        # [index, chains, copy-in done, run start, run done, copy-out done] as timing events
        tl = getattr(self, "host_timeline", None)

        def _tev(stream):
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(stream)
            return ev
        if tl is not None:
            tl.append(_tev(cur))
        for i in range(len(bounds) - 1):
            lo, hi = bounds[i], bounds[i + 1]
            if hi == lo:
                continue
            with torch.cuda.stream(s_in):
                stage["q_in"][lo:hi].copy_(q_host[lo:hi], non_blocking=True)
                ev_in = _tev(s_in) if tl is not None else s_in.record_event()
            # consecutive chunks alternate between two compute streams (each with its own scratch), so the
            # thin last DR stages of one chunk share the GPU with the first stage of the next
            s_run = runs[i % len(runs)]
            with torch.cuda.stream(s_run):
                s_run.wait_event(ev_in)
                if tl is not None:
                    tl.append([i, hi - lo, ev_in, _tev(s_run)])
                self.drhmc_step(stage["q_in"][lo:hi], n_leap, step_size, nprops, stepfactor, want_p=want_p,
                                q_out=stage["q"][lo:hi], p_out=None if not want_p else stage["p"][lo:hi],
                                code_out=stage["code"][lo:hi], avec_out=stage["avec"][lo:hi],
                                counts_out=stage["counts"][lo:hi], extra_offset=lo, ws_slot=i % len(runs),
                                advance=False)
                ev_run = _tev(s_run) if tl is not None else s_run.record_event()
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_run)
                for k in ("q", "p", "code", "avec", "counts"):
                    if host[k] is not None:
                        host[k][lo:hi].copy_(stage[k][lo:hi], non_blocking=True)
                if tl is not None:
                    tl[-1] += [ev_run, _tev(s_out)]
        s_out.synchronize()
        for s_run in runs:
            cur.wait_stream(s_run)
        self.iteration += 1
        return host

    def ghmc_step(self, q, p_prev, n_leap, step_size, alpha, V=None, Vg=None, pn=None, u=None, pu=None, delta=0.0):
        """hmcb_ghmc_step on device tensors: q, p_prev [C,d]; cached V [C], Vg [C,d] or None;
        pu [C] float64 = the persistent uniform (Persistent_GHMC) or None."""
        m = self._ensure()
        C_ = q.shape[0]
        with torch.cuda.device(self.device):
            out = {"q": torch.empty_like(q), "p": torch.empty_like(q),
                   "V": torch.empty(C_, dtype=self.dtype, device=self.device), "Vg": torch.empty_like(q),
                   "u": torch.empty(C_, dtype=torch.float64, device=self.device) if pu is not None else None,
                   "code": torch.empty(C_, dtype=torch.int32, device=self.device),
                   "Hs": torch.empty((C_, 2), dtype=self.dtype, device=self.device),
                   "counts": torch.empty((C_, 3), dtype=torch.int32, device=self.device)}
            io = _lib.GhmcIO()
            io.p_out, io.V_in, io.Vg_in = _lib.ptr(out["p"]), _lib.ptr(V), _lib.ptr(Vg)
            io.V_out, io.Vg_out = _lib.ptr(out["V"]), _lib.ptr(out["Vg"])
            io.u_in, io.u_out, io.delta = _lib.ptr(pu), _lib.ptr(out["u"]), float(delta)
            io.code, io.energies, io.counts = _lib.ptr(out["code"]), _lib.ptr(out["Hs"]), _lib.ptr(out["counts"])
            ch = self._chains(q, out["q"])
            r, keep = self._rng(pn, u, 1, C_, p_prev, alpha)
            _lib.check(self.lib.hmcb_ghmc_step(C.byref(m), C.byref(ch), C.byref(r), int(n_leap), float(step_size),
                                               C.byref(io), self._stream()))
        self.iteration += 1
        out["_keep"] = keep
        return out

    def ghmc_sample(self, q, p_prev, n_leap, step_size, alpha, n_transitions, V=None, Vg=None, pu=None, delta=0.0,
                    n_accept=None):
        """hmcb_ghmc_sample: n_transitions GHMC (pu=None) or Persistent_GHMC transitions in one call; the returned
        q, p, V, Vg, u are the final ones, code / Hs the last transition's, counts the sums; n_accept [C] int32 (or a
        new zeroed one) is incremented by the number of accepted transitions."""
        m = self._ensure()
        C_ = q.shape[0]
        K = int(n_transitions)
        with torch.cuda.device(self.device):
            if n_accept is None:
                n_accept = torch.zeros(C_, dtype=torch.int32, device=self.device)
            out = {"q": torch.empty_like(q), "p": torch.empty_like(q),
                   "V": torch.empty(C_, dtype=self.dtype, device=self.device), "Vg": torch.empty_like(q),
                   "u": torch.empty(C_, dtype=torch.float64, device=self.device) if pu is not None else None,
                   "code": torch.empty(C_, dtype=torch.int32, device=self.device),
                   "Hs": torch.empty((C_, 2), dtype=self.dtype, device=self.device),
                   "counts": torch.empty((C_, 3), dtype=torch.int32, device=self.device), "n_accept": n_accept}
            io = _lib.GhmcIO()
            io.p_out, io.V_in, io.Vg_in = _lib.ptr(out["p"]), _lib.ptr(V), _lib.ptr(Vg)
            io.V_out, io.Vg_out = _lib.ptr(out["V"]), _lib.ptr(out["Vg"])
            io.u_in, io.u_out, io.delta = _lib.ptr(pu), _lib.ptr(out["u"]), float(delta)
            io.code, io.energies, io.counts = _lib.ptr(out["code"]), _lib.ptr(out["Hs"]), _lib.ptr(out["counts"])
            ch = self._chains(q, out["q"])
            r, keep = self._rng(None, None, 1, C_, p_prev, alpha)
            _lib.check(self.lib.hmcb_ghmc_sample(C.byref(m), C.byref(ch), C.byref(r), int(n_leap), float(step_size),
                                                 C.byref(io), K, C.c_void_p(n_accept.data_ptr()), self._stream()))
        self.iteration += K
        out["_keep"] = keep
        return out

    def mala_step(self, q, eps, n_steps=1, V=None, Vg=None, z=None, u=None, q_out=None):
        """hmcb_mala_step on device tensor q [C,d]; V [C] / Vg [C,d] are the cached
        potential and its gradient (None on the first call)."""
        m = self._ensure()
        C_ = q.shape[0]
        with torch.cuda.device(self.device):
            q_out = torch.empty_like(q) if q_out is None else q_out
            vdt = torch.float64 if self.is_logistic else self.dtype     # the logistic path carries V(q) in float64
            if V is not None:
                V = V.to(vdt).contiguous()
            out = {"q": q_out, "V": torch.empty(C_, dtype=vdt, device=self.device), "Vg": torch.empty_like(q),
                   "code": torch.empty(C_, dtype=torch.int32, device=self.device),
                   "n_accept": torch.empty(C_, dtype=torch.int32, device=self.device),
                   "counts": torch.empty((C_, 3), dtype=torch.int32, device=self.device)}
            io = _lib.MalaIO()
            io.V_in, io.Vg_in = _lib.ptr(V), _lib.ptr(Vg)
            io.V_out, io.Vg_out, io.code = _lib.ptr(out["V"]), _lib.ptr(out["Vg"]), _lib.ptr(out["code"])
            io.n_accept, io.counts = _lib.ptr(out["n_accept"]), _lib.ptr(out["counts"])
            ch = self._chains(q, q_out)
            r, keep = self._rng(z, u, n_steps, C_)
            keep.append(V)
            if self.is_logistic:
                ws = self._lr_workspace(C_)
                _lib.check(self.lib.hmcb_logreg_mala_step(C.byref(m), C.c_void_p(self._lr_prepared.data_ptr()), C.byref(ch),
                                                          C.byref(r), float(eps), int(n_steps), C.byref(io),
                                                          C.c_void_p(ws.data_ptr()), ws.numel(), self._stream()))
            else:
                _lib.check(self.lib.hmcb_mala_step(C.byref(m), C.byref(ch), C.byref(r), float(eps), int(n_steps),
                                                   C.byref(io), self._stream()))
        self.iteration += int(n_steps)
        out["_keep"] = keep
        return out

    def adhmc_eps0_step(self, q, n_leap, steps, nsteps, p=None, u=None, q_out=None, tint=0.0):
        m = self._ensure()
        C_ = q.shape[0]
        steps = np.ascontiguousarray(steps, dtype=np.float64)
        with torch.cuda.device(self.device):
            need = self.lib.hmcb_adhmc_eps0_workspace_bytes(C.byref(m), C_)
            if need == 0:
                _lib.check(_lib.E_ARG)
            if self._ws is None or self._ws.numel() < need:
                self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
            q_out = torch.empty_like(q) if q_out is None else q_out
            out = {"q": q_out, "p": torch.empty_like(q),
                   "code": torch.empty(C_, dtype=torch.int32, device=self.device),
                   "probs": torch.empty((C_, 3), dtype=self.dtype, device=self.device),
                   "aux": torch.empty((C_, 3), dtype=self.dtype, device=self.device),
                   "counts": torch.empty((C_, 3), dtype=torch.int32, device=self.device)}
            so = _lib.AdhmcOut()
            so.p_out, so.code, so.probs = _lib.ptr(out["p"]), _lib.ptr(out["code"]), _lib.ptr(out["probs"])
            so.aux, so.counts = _lib.ptr(out["aux"]), _lib.ptr(out["counts"])
            ch = self._chains(q, q_out)
            r, keep = self._rng(p, u, 3, C_)
            _lib.check(self.lib.hmcb_adhmc_eps0_step(
                C.byref(m), C.byref(ch), C.byref(r), int(n_leap), float(tint), steps.ctypes.data_as(C.POINTER(C.c_double)),
                int(steps.size), int(nsteps), C.byref(so), C.c_void_p(self._ws.data_ptr()), self._ws.numel(),
                self._stream()))
        self.iteration += 1
        out["_keep"] = keep
        return out

    def adhmc_walk_step(self, q, n_segments, n_leap_segment, tint, ratios, pwts0, smin, smax, steps, nsteps, p=None,
                        u=None, q_out=None):
        """hmcb_adhmc_walk_step on device tensor q [C,d] (AdHMC / AdHMC_tint)."""
        m = self._ensure()
        C_ = q.shape[0]
        steps = np.ascontiguousarray(steps, dtype=np.float64)
        nr = len(ratios)
        if not 1 <= nr <= 4 or len(pwts0) != nr:
            raise ValueError("ratios and pwts0 must have the same length, at most 4")
        par = _lib.WalkPar()
        par.n_segments, par.n_leap_segment, par.tint = int(n_segments), int(n_leap_segment), float(tint)
        par.n_ratios, par.ntry, par.nsteps = nr, int(steps.size), int(nsteps)
        for j in range(nr):
            par.ratios[j], par.pwts0[j] = float(ratios[j]), float(pwts0[j])
        par.smin, par.smax = float(smin), float(smax)
        par.steps = steps.ctypes.data_as(C.POINTER(C.c_double))
        with torch.cuda.device(self.device):
            need = self.lib.hmcb_adhmc_walk_workspace_bytes(C.byref(m), C_, nr)
            if need == 0:
                _lib.check(_lib.E_ARG)
            if self._ws is None or self._ws.numel() < need:
                self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
            q_out = torch.empty_like(q) if q_out is None else q_out
            out = {"q": q_out, "p": torch.empty_like(q),
                   "code": torch.empty(C_, dtype=torch.int32, device=self.device),
                   "probs": torch.empty((C_, 3), dtype=self.dtype, device=self.device),
                   "checks": torch.empty((C_, 3, int(n_segments)), dtype=self.dtype, device=self.device),
                   "counts": torch.empty((C_, 3), dtype=torch.int32, device=self.device)}
            so = _lib.WalkOut()
            so.p_out, so.code, so.probs = _lib.ptr(out["p"]), _lib.ptr(out["code"]), _lib.ptr(out["probs"])
            so.checks, so.counts = _lib.ptr(out["checks"]), _lib.ptr(out["counts"])
            ch = self._chains(q, q_out)
            r, keep = self._rng(p, u, int(n_segments) + 2, C_)
            _lib.check(self.lib.hmcb_adhmc_walk_step(C.byref(m), C.byref(ch), C.byref(r), C.byref(par), C.byref(so),
                                                     C.c_void_p(self._ws.data_ptr()), self._ws.numel(), self._stream()))
        self.iteration += 1
        out["_keep"] = keep
        return out

    @staticmethod
    def _profile(stage_events):
        """torch.cuda.Event list (enable_timing=True) -> hmcb_profile, or NULL."""
        if not stage_events:
            return None
        prof = _lib.Profile()
        for i, ev in enumerate(stage_events):
            if not ev.cuda_event:
                ev.record()  # torch creates the cudaEvent_t lazily, at the first record
            prof.events[i] = ev.cuda_event
        return C.byref(prof)

    # ------------------------------------------------------------------ streamed moments
    def moments_accumulate(self, q, msum, msumsq):
        self._ensure()
        dt = _lib.F64 if self.dtype == torch.float64 else _lib.F32
        with torch.cuda.device(self.device):
            _lib.check(self.lib.hmcb_moments_accumulate(dt, q.shape[0], q.shape[1], _lib.ptr(q), q.stride(0),
                                                        q.stride(1), _lib.ptr(msum), _lib.ptr(msumsq), self._stream()))

    def moments_reduce(self, msum, msumsq, n_samples, true_mean=None):
        """[4,d] float64 device tensor: sum_c m, sum_c m^2, sum_c s2, sum_c (m-mu)^2."""
        self._ensure()
        dt = _lib.F64 if self.dtype == torch.float64 else _lib.F32
        d = msum.shape[1]
        with torch.cuda.device(self.device):
            out = torch.empty((4, d), dtype=torch.float64, device=self.device)
            mu = None if true_mean is None else torch.as_tensor(np.asarray(true_mean, dtype=np.float64),
                                                                device=self.device)
            _lib.check(self.lib.hmcb_moments_reduce(dt, msum.shape[0], d, _lib.ptr(msum), _lib.ptr(msumsq),
                                                    float(n_samples), _lib.ptr(mu), _lib.ptr(out), self._stream()))
        return out

    def leapfrog(self, q, p, n_leap, step_size):
        m = self._ensure()
        with torch.cuda.device(self.device):
            q1, p1 = torch.empty_like(q), torch.empty_like(p)
            ch = self._chains(q, q1)
            _lib.check(self.lib.hmcb_leapfrog(C.byref(m), C.byref(ch), _lib.ptr(p), _lib.ptr(p1), int(n_leap),
                                              float(step_size), self._stream()))
        return q1, p1

    def logp_grad(self, q):
        m = self._ensure()
        if self.is_logistic:
            with torch.cuda.device(self.device):
                logp = torch.empty(q.shape[0], dtype=self.dtype, device=self.device)
                grad = torch.empty_like(q)
                ws = self._lr_workspace(q.shape[0])
                _lib.check(self.lib.hmcb_logreg_logp_grad(C.byref(m), C.c_void_p(self._lr_prepared.data_ptr()),
                                                          _lib.ptr(q), q.shape[0], _lib.ptr(logp), _lib.ptr(grad),
                                                          C.c_void_p(ws.data_ptr()), ws.numel(), self._stream()))
            return logp, grad
        with torch.cuda.device(self.device):
            logp = torch.empty(q.shape[0], dtype=self.dtype, device=self.device)
            grad = torch.empty_like(q)
            ch = self._chains(q, q)
            _lib.check(self.lib.hmcb_logp_grad(C.byref(m), C.byref(ch), _lib.ptr(logp), _lib.ptr(grad),
                                               self._stream()))
        return logp, grad
