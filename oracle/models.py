"""numpy statements of the benchmark densities (oracle side; test infrastructure).

Every model exposes the three callables the reference's samplers take
(src/algorithms.py:81-89): ``log_prob(x) -> float``, ``grad_log_prob(x) ->
float64[d]`` and ``log_prob_and_grad(x) -> (float, float64[d])`` on a 1-D
float64 array.

Funnel: restated from the Stan program text at scripts/models.py:9-19
(``v ~ normal(0,3); theta ~ normal(0, exp(v/2))``) and the TFP restatement at
scripts_old/deprecated/funneltf.py:60-67, additive constants dropped (Stan's
``~`` statements drop them; trajectories and accept decisions do not depend on
them).  Correlated Gaussian, Rosenbrock and logistic regression do not exist in
the reference; they are defined here.
"""
import numpy as np


class Funnel:
    """x[0] = v, x[1:] = theta; log p = -v^2/(2 s^2) - (n/2) v - exp(-v)/2 * sum(theta^2)."""

    name = "funnel"

    def __init__(self, d, sigma=3.0):
        assert d >= 2
        self.d, self.sigma = int(d), float(sigma)
        self.n = self.d - 1

    def log_prob(self, x):
        v, th = x[0], x[1:]
        S = (th * th).sum()
        return -v * v / (2.0 * self.sigma ** 2) - 0.5 * self.n * v - 0.5 * np.exp(-v) * S

    def grad_log_prob(self, x):
        v, th = x[0], x[1:]
        e = np.exp(-v)
        S = (th * th).sum()
        g = np.empty_like(x)
        g[0] = -v / self.sigma ** 2 - 0.5 * self.n + 0.5 * e * S
        g[1:] = -e * th
        return g

    def log_prob_and_grad(self, x):
        return self.log_prob(x), self.grad_log_prob(x)

    # analytic truth used by the statistical tests (scripts_old/tests/test_drhmc_funnel.py:244-254)
    def true_mean(self):
        return np.zeros(self.d)

    def true_std(self):
        s = np.full(self.d, np.exp(self.sigma ** 2 / 4.0))
        s[0] = self.sigma
        return s


class CorrGaussian:
    """log p = -x^T P x / 2 with a dense symmetric precision matrix P."""

    name = "gaussian"

    def __init__(self, P):
        self.P = np.ascontiguousarray(P, dtype=np.float64)
        assert self.P.ndim == 2 and self.P.shape[0] == self.P.shape[1]
        self.d = self.P.shape[0]

    @classmethod
    def ar1(cls, d, rho=0.9):
        """Benchmark instance: Sigma_ij = rho^|i-j|, P = Sigma^-1 stored dense."""
        idx = np.arange(d)
        Sigma = rho ** np.abs(idx[:, None] - idx[None, :])
        m = cls(np.linalg.inv(Sigma))
        m.Sigma = Sigma
        return m

    def log_prob(self, x):
        return -0.5 * float(x @ (self.P @ x))

    def grad_log_prob(self, x):
        return -(self.P @ x)

    def log_prob_and_grad(self, x):
        Px = self.P @ x
        return -0.5 * float(x @ Px), -Px


class Rosenbrock:
    """Pairwise Rosenbrock: pairs (u_k, w_k) = (x[2k], x[2k+1]);
    log p = -sum_k[(a-u_k)^2 + b (w_k-u_k^2)^2] / T.  With a=1, b=100, T=20:
    u ~ N(1, 10), w | u ~ N(u^2, 0.1)."""

    name = "rosenbrock"

    def __init__(self, d, a=1.0, b=100.0, T=20.0):
        assert d % 2 == 0 and d >= 2
        self.d, self.a, self.b, self.T = int(d), float(a), float(b), float(T)

    def log_prob(self, x):
        u, w = x[0::2], x[1::2]
        r = w - u * u
        return -(((self.a - u) ** 2).sum() + self.b * (r * r).sum()) / self.T

    def grad_log_prob(self, x):
        u, w = x[0::2], x[1::2]
        r = w - u * u
        g = np.empty_like(x)
        g[0::2] = (2.0 * (self.a - u) + 4.0 * self.b * u * r) / self.T
        g[1::2] = -2.0 * self.b * r / self.T
        return g

    def log_prob_and_grad(self, x):
        return self.log_prob(x), self.grad_log_prob(x)


class LogisticRegression:
    """log p = sum_i [y_i z_i - softplus(z_i)] - |theta|^2/(2 tau^2), z = X theta."""

    name = "logistic"

    def __init__(self, X, y, tau=1.0):
        self.X = np.ascontiguousarray(X, dtype=np.float64)
        self.y = np.ascontiguousarray(y, dtype=np.float64)
        self.tau = float(tau)
        self.N, self.d = self.X.shape

    def log_prob(self, x):
        z = self.X @ x
        return float((self.y * z - np.logaddexp(0.0, z)).sum() - (x * x).sum() / (2.0 * self.tau ** 2))

    def grad_log_prob(self, x):
        z = self.X @ x
        s = 1.0 / (1.0 + np.exp(-z))
        return self.X.T @ (self.y - s) - x / self.tau ** 2

    def log_prob_and_grad(self, x):
        return self.log_prob(x), self.grad_log_prob(x)
