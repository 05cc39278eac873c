"""Product-side numpy model callables agree with the oracle's, and both have
correct gradients (finite differences)."""
import numpy as np
import pytest

from hmc_b200 import models as pm
from oracle import models as om


def _pairs():
    rng = np.random.default_rng(3)
    X = rng.standard_normal((50, 5))
    y = (rng.random(50) < 0.4).astype(float)
    g = om.CorrGaussian.ar1(7)
    return [(pm.Funnel(9, 2.5), om.Funnel(9, 2.5)), (pm.CorrGaussian(0.5 * (g.P + g.P.T)), g),
            (pm.Rosenbrock(6), om.Rosenbrock(6)), (pm.LogisticRegression(X, y, 2.0), om.LogisticRegression(X, y, 2.0))]


@pytest.mark.parametrize("i", range(4))
def test_product_and_oracle_models_agree(i):
    a, b = _pairs()[i]
    x = np.random.default_rng(i).standard_normal(a.d) * 0.7
    assert np.isclose(a.log_prob(x), b.log_prob(x), rtol=1e-12)
    assert np.allclose(a.grad_log_prob(x), b.grad_log_prob(x), rtol=1e-12, atol=1e-12)
    lp, g = a.log_prob_and_grad(x)
    assert lp == a.log_prob(x) and np.array_equal(g, a.grad_log_prob(x))


@pytest.mark.parametrize("i", range(4))
def test_oracle_gradients_match_finite_differences(i):
    _, m = _pairs()[i]
    x = np.random.default_rng(10 + i).standard_normal(m.d) * 0.5
    g = m.grad_log_prob(x)
    h = 1e-6
    fd = np.array([(m.log_prob(x + h * e) - m.log_prob(x - h * e)) / (2 * h) for e in np.eye(m.d)])
    assert np.allclose(g, fd, rtol=1e-6, atol=1e-6)


def test_rosenbrock_moments():
    m = pm.Rosenbrock(4)
    assert np.allclose(m.true_mean(), [1, 11, 1, 11])
    assert np.allclose(m.true_std() ** 2, [10, 240.1, 10, 240.1])


def test_funnel_truth():
    m = pm.Funnel(5)
    assert np.allclose(m.true_std(), [3.0] + [np.exp(2.25)] * 4)
