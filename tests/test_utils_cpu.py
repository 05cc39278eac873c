"""hmc_b200.utils diagnostics against the oracle (which is pinned bit-for-bit to the
reference's src/utils.py), and the host-side sampler bookkeeping."""
import numpy as np
import torch

from hmc_b200 import algorithms, dist, utils
from oracle import samplers


def _x(seed=0, T=300, C=6, d=4):
    return np.random.default_rng(seed).standard_normal((T, C, d)).cumsum(axis=0) * 0.1


def test_get_ess_true_matches_oracle():
    x = _x()
    mean, std = np.zeros(4), np.ones(4) * 2
    for a, b in zip(utils.get_ess_true(x, mean, std), samplers.get_ess_true(x, mean, std)):
        assert np.allclose(a, b, rtol=1e-13)
    ta, tb = utils.get_ess_true(torch.as_tensor(x), mean, std)
    assert np.allclose(ta.numpy(), samplers.get_ess_true(x, mean, std)[0], rtol=1e-12)


def test_get_rcc_matches_oracle():
    x = _x(1)
    r1, t1 = utils.get_rcc(x)
    r2, t2 = samplers.get_rcc(x)
    assert np.allclose(r1, r2, rtol=1e-9, atol=1e-10)
    assert np.array_equal(t1, t2)


def test_dual_averaging_known_answers():
    # generated from the reference, numpy 2.3.5
    da = algorithms.DualAveragingStepSize(0.1)
    assert np.allclose(da.update(0.9), (4.8604879304466735, 1.3246796579134237), rtol=1e-15)
    assert np.allclose(da.update(0.3), (0.5471552537388176, 1.1442891199116336), rtol=1e-15)
    assert np.allclose(da.update(0.65), (0.561383913798928, 1.0246294655563801), rtol=1e-15)
    mine, ora = algorithms.DualAveragingStepSize(0.2, 0.8), samplers.DualAveragingStepSize(0.2, 0.8)
    for p in [0.1, np.nan, 3.0, np.inf, 0.7, 0.0]:
        assert mine.update(p) == ora.update(p)
    sched = algorithms.DualAveragingStepSize(0.1, nadapt=3)
    assert sched(0, 0.5) == 0.1 and sched(1, 0.5) != 0.1


def test_sampler_container():
    s = algorithms.Sampler()
    s.appends(np.zeros(3), 1.0, [0.0, 1.0], [2, 21, 1])
    s.appends(np.ones(3), 0.0, [0.0, 2.0], [2, 21, 1])
    s.to_array()
    assert s.samples.shape == (2, 3) and s.counts.shape == (2, 3) and s.i == 2
    s.to_list()
    assert isinstance(s.samples, list)
    # the reference's registry keys (src/algorithms.py:823-832)
    assert set(algorithms.alg_dict) == {"hmc", "ghmc", "pghmc", "drhmc", "drghmc", "pdrhmc_metrop", "pdrhmc_hasting"}
    from oracle import ref_loader
    if ref_loader.available():
        ref = ref_loader.load("algorithms").alg_dict
        assert set(ref) == set(algorithms.alg_dict)
        assert all(ref[k].__name__ == algorithms.alg_dict[k].__name__ for k in ref)


def test_shard_range_partitions():
    for n, R in [(10, 3), (1 << 20, 8), (7, 8), (0, 2)]:
        spans = [dist.shard_range(n, r, R) for r in range(R)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(spans[i][1] == spans[i + 1][0] for i in range(R - 1))
        assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1


def test_dr_stage_lengths_truncate_like_the_reference():
    L, s = algorithms.dr_stage_lengths(7, 4, 1.5)
    assert L == [7, 10, 15, 23] and s[1] == 1 / 1.5
    Lo, _ = samplers.dr_leapfrog_work(7, 4, 1.5)
    assert L == Lo
