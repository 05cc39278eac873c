"""Streamed-moment kernels against stored-sample arithmetic (utils.get_ess_true of
the oracle, which is pinned to src/utils.py:8-27)."""
import numpy as np
import pytest
import torch

from hmc_b200 import models, pyhmc, utils
from oracle import samplers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("C,d,T", [(1000, 50, 7), (37, 3, 4), (5000, 300, 3), (1, 10, 5)])
def test_running_moments_match_stored_samples(C, d, T):
    dev = torch.device("cuda:0")
    eng = pyhmc.HMC(models.Funnel(max(d, 2)), device=dev).engine
    x = torch.randn(T, C, d, dtype=torch.float64, device=dev).cumsum(0) * 0.3 + 1.0
    mom = utils.RunningMoments(eng, C, d)
    for t in range(T):
        mom.add(x[t])
    mean, std = np.linspace(0.5, 1.5, d), np.linspace(1.0, 3.0, d)
    sums, Ctot = mom.reduce(mean)
    ref = utils.local_moment_sums(x.cpu(), mean)
    assert Ctot == C
    assert torch.allclose(sums.cpu(), ref, rtol=1e-10, atol=1e-10)
    ess_ref, _ = samplers.get_ess_true(x.cpu().numpy(), mean, std)
    assert np.allclose(mom.ess_true(mean, std).cpu().numpy(), ess_ref, rtol=1e-9)
    if C > 1:
        xm = x.cpu().numpy()
        W = xm.var(axis=0, ddof=1).mean(axis=0)
        rhat = np.sqrt(((T - 1) / T * W + xm.mean(axis=0).var(axis=0, ddof=1)) / W)
        assert np.allclose(mom.rhat().cpu().numpy(), rhat, rtol=1e-9)


def test_moments_on_strided_state_and_float32():
    dev = torch.device("cuda:0")
    eng = pyhmc.HMC(models.Funnel(8), device=dev, dtype=torch.float32).engine
    C, d = 513, 8
    soa = torch.randn(d, C, dtype=torch.float32, device=dev)     # structure-of-arrays state
    q = soa.t()                                                   # [C, d] view with strides (1, C)
    mom = utils.RunningMoments(eng, C, d)
    mom.add(q)
    mom.add(q)
    assert torch.allclose(mom.sum, 2 * q, rtol=1e-6)
    assert torch.allclose(mom.sumsq, 2 * q * q, rtol=1e-6)
