"""Model descriptors: what the samplers take where the reference takes
``log_prob`` / ``grad_log_prob`` callables (src/algorithms.py:81-89,
src/pyhmc.py:7-10).

A descriptor is two things at once:

* it quacks like the reference's model plug-in -- ``log_prob(x)``,
  ``grad_log_prob(x)``, ``log_prob_and_grad(x)`` on a 1-D float64 numpy array --
  so the reference's own classes (and its scripts, e.g. scripts/test.py:62-81)
  can consume it unchanged;
* it carries the ``model_id`` + parameter block of the in-kernel analytic
  gradient (``hmcb_model`` in include/hmcb.h).

The GPU samplers only ever use the second face: an arbitrary Python callable
cannot run inside a kernel and there is no CPU fallback, so passing one raises
``TypeError``.  The numpy callables here are an API courtesy, never on the
sampling path.

The funnel follows the reference's Stan text (scripts/models.py:9-19); the
other densities are defined by this build (the reference has none).
"""
import numpy as np

from . import _lib


class Model:
    kind = None
    name = None

    # ---- numpy plug-in face (reference API) ---------------------------------
    def log_prob(self, x):
        raise NotImplementedError

    def grad_log_prob(self, x):
        raise NotImplementedError

    def log_prob_and_grad(self, x):
        return self.log_prob(x), self.grad_log_prob(x)

    # ---- kernel face ---------------------------------------------------------
    def _params(self):
        return [0.0, 0.0, 0.0, 0.0]

    def _device_blocks(self, device, dtype):
        """(dev0, dev1) torch tensors or None; kept alive by the struct holder."""
        return None, None

    def to_struct(self, device, dtype, invmetric_diag=None):
        """Build the hmcb_model block for ``device``; returns (struct, keepalive)."""
        import torch

        s = _lib.Model()
        s.kind, s.d = self.kind, self.d
        s.dtype = _lib.F64 if dtype == torch.float64 else _lib.F32
        s.n_data = getattr(self, "N", 0)
        for i, v in enumerate(self._params()):
            s.par[i] = v
        dev0, dev1 = self._device_blocks(device, dtype)
        keep = [dev0, dev1]
        s.dev0 = dev0.data_ptr() if dev0 is not None else None
        s.dev1 = dev1.data_ptr() if dev1 is not None else None
        if invmetric_diag is not None:
            m = torch.as_tensor(np.broadcast_to(np.asarray(invmetric_diag, dtype=np.float64), (self.d,)).copy(),
                                dtype=dtype, device=device).contiguous()
            keep.append(m)
            s.invmetric_diag = m.data_ptr()
        else:
            s.invmetric_diag = None
        return s, keep


def as_model(log_prob, grad_log_prob=None, log_prob_and_grad=None):
    """Resolve the reference-style constructor arguments to a descriptor.

    Accepts ``Cls(model)`` or ``Cls(model.log_prob, model.grad_log_prob)`` (bound
    methods of one descriptor).  Anything else is an opaque callable."""
    if isinstance(log_prob, Model):
        return log_prob
    owners = [getattr(f, "__self__", None) for f in (log_prob, grad_log_prob, log_prob_and_grad) if f is not None]
    if owners and all(isinstance(o, Model) for o in owners) and all(o is owners[0] for o in owners):
        return owners[0]
    raise TypeError(
        "hmc_b200 samplers need a model descriptor (Funnel, CorrGaussian, Rosenbrock, LogisticRegression) or its "
        "bound log_prob/grad_log_prob: an arbitrary Python callable cannot run inside a CUDA kernel and there is "
        "no CPU fallback")


class Funnel(Model):
    """Neal's funnel: x[0] = v ~ N(0, sigma^2), x[1:] = theta ~ N(0, exp(v))."""

    kind, name = _lib.MODEL_FUNNEL, "funnel"

    def __init__(self, d, sigma=3.0):
        if d < 2:
            raise ValueError("funnel needs d >= 2")
        self.d, self.sigma = int(d), float(sigma)

    def log_prob(self, x):
        x = np.asarray(x, dtype=np.float64)
        v, th = x[0], x[1:]
        return -v * v / (2.0 * self.sigma ** 2) - 0.5 * (self.d - 1) * v - 0.5 * np.exp(-v) * (th * th).sum()

    def grad_log_prob(self, x):
        x = np.asarray(x, dtype=np.float64)
        v, th = x[0], x[1:]
        e = np.exp(-v)
        g = np.empty_like(x)
        g[0] = -v / self.sigma ** 2 - 0.5 * (self.d - 1) + 0.5 * e * (th * th).sum()
        g[1:] = -e * th
        return g

    def _params(self):
        return [self.sigma, 0.0, 0.0, 0.0]

    def true_mean(self):
        return np.zeros(self.d)

    def true_std(self):
        s = np.full(self.d, np.exp(self.sigma ** 2 / 4.0))
        s[0] = self.sigma
        return s


class CorrGaussian(Model):
    """Zero-mean Gaussian with dense symmetric precision P: log p = -x^T P x / 2."""

    kind, name = _lib.MODEL_GAUSSIAN, "gaussian"

    def __init__(self, P):
        self.P = np.ascontiguousarray(P, dtype=np.float64)
        if self.P.ndim != 2 or self.P.shape[0] != self.P.shape[1]:
            raise ValueError("P must be square")
        if not np.allclose(self.P, self.P.T, rtol=1e-12, atol=1e-12):
            raise ValueError("P must be symmetric")
        self.d = self.P.shape[0]

    @classmethod
    def ar1(cls, d, rho=0.9):
        idx = np.arange(d)
        Sigma = rho ** np.abs(idx[:, None] - idx[None, :])
        P = np.linalg.inv(Sigma)
        m = cls(0.5 * (P + P.T))
        m.Sigma = Sigma
        return m

    def log_prob(self, x):
        x = np.asarray(x, dtype=np.float64)
        return -0.5 * float(x @ (self.P @ x))

    def grad_log_prob(self, x):
        return -(self.P @ np.asarray(x, dtype=np.float64))

    def _device_blocks(self, device, dtype):
        import torch

        return torch.as_tensor(self.P, dtype=dtype, device=device).contiguous(), None

    def true_mean(self):
        return np.zeros(self.d)

    def true_std(self):
        return np.sqrt(np.diag(np.linalg.inv(self.P)))


class Rosenbrock(Model):
    """Pairwise Rosenbrock: (u_k, w_k) = (x[2k], x[2k+1]),
    log p = -sum_k [(a-u_k)^2 + b (w_k-u_k^2)^2] / T."""

    kind, name = _lib.MODEL_ROSENBROCK, "rosenbrock"

    def __init__(self, d, a=1.0, b=100.0, T=20.0):
        if d < 2 or d % 2:
            raise ValueError("pairwise Rosenbrock needs an even d >= 2")
        self.d, self.a, self.b, self.T = int(d), float(a), float(b), float(T)

    def log_prob(self, x):
        x = np.asarray(x, dtype=np.float64)
        u, w = x[0::2], x[1::2]
        r = w - u * u
        return -(((self.a - u) ** 2).sum() + self.b * (r * r).sum()) / self.T

    def grad_log_prob(self, x):
        x = np.asarray(x, dtype=np.float64)
        u, w = x[0::2], x[1::2]
        r = w - u * u
        g = np.empty_like(x)
        g[0::2] = (2.0 * (self.a - u) + 4.0 * self.b * u * r) / self.T
        g[1::2] = -2.0 * self.b * r / self.T
        return g

    def _params(self):
        return [self.a, self.b, self.T, 0.0]

    def true_mean(self):
        # u ~ N(a, T/2), w|u ~ N(u^2, T/(2b)): E[w] = a^2 + T/2
        m = np.empty(self.d)
        m[0::2] = self.a
        m[1::2] = self.a ** 2 + self.T / 2.0
        return m

    def true_std(self):
        s2u = self.T / 2.0
        s = np.empty(self.d)
        s[0::2] = np.sqrt(s2u)
        # Var[w] = T/(2b) + Var[u^2] = T/(2b) + 2 s2u^2 + 4 a^2 s2u
        s[1::2] = np.sqrt(self.T / (2.0 * self.b) + 2.0 * s2u ** 2 + 4.0 * self.a ** 2 * s2u)
        return s


class LogisticRegression(Model):
    """Bayesian logistic regression: log p = sum_i [y_i z_i - softplus(z_i)] - |theta|^2/(2 tau^2), z = X theta."""

    kind, name = _lib.MODEL_LOGISTIC, "logistic"

    def __init__(self, X, y, tau=1.0):
        self.X = np.ascontiguousarray(X)
        self.y = np.ascontiguousarray(y)
        self.tau = float(tau)
        self.N, self.d = self.X.shape

    def log_prob(self, x):
        x = np.asarray(x, dtype=np.float64)
        z = self.X.astype(np.float64) @ x
        return float((self.y * z - np.logaddexp(0.0, z)).sum() - (x * x).sum() / (2.0 * self.tau ** 2))

    def grad_log_prob(self, x):
        x = np.asarray(x, dtype=np.float64)
        Xd = self.X.astype(np.float64)
        s = 1.0 / (1.0 + np.exp(-(Xd @ x)))
        return Xd.T @ (self.y - s) - x / self.tau ** 2

    def _params(self):
        return [self.tau, 0.0, 0.0, 0.0]

    def _device_blocks(self, device, dtype):
        import torch

        return (torch.as_tensor(self.X, dtype=dtype, device=device).contiguous(),
                torch.as_tensor(self.y, dtype=dtype, device=device).contiguous())
