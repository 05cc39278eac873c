"""hmc_b200.results -- the reference's result-file layout and sample post-processing
(scripts_old/funnel/funnelmpi.py:171-211, scripts_old/tests/test_drhmc_funnel.py:196-279,
src/utils.py:75-146).  Host-side only; compared with the reference's own utils where the
reference tree is present."""
import json
import os

import numpy as np
import pytest

from hmc_b200 import results
from oracle import ref_loader


def _ar1_chains(T, C, d, rho, seed):
    rng = np.random.default_rng(seed)
    x = np.zeros((T, C, d))
    x[0] = rng.standard_normal((C, d))
    for t in range(1, T):
        x[t] = rho * x[t - 1] + np.sqrt(1 - rho * rho) * rng.standard_normal((C, d))
    return x


def _fake_run(T, C, d, nprops, seed):
    rng = np.random.default_rng(seed)
    samples = rng.standard_normal((T, C, d))
    accepted = rng.integers(0, nprops + 1, (T, C)).astype(np.float64)
    probs = rng.random((T, C, nprops))
    counts = rng.integers(1, 500, (T, C, 3))
    return samples, accepted, probs, counts


def test_rank_files_and_merge_follow_the_reference_layout(tmp_path):
    fpath = str(tmp_path / "run")
    burnin, T, d, nprops = 3, 12, 5, 3
    blocks = [_fake_run(T + burnin, c, d, nprops, 10 + r) for r, c in enumerate((4, 3))]   # ragged shards
    for r, b in enumerate(blocks):
        results.save_rank(fpath, r, *b, burnin=burnin)
    for kind in ("samples", "accepts", "probs", "counts"):
        assert sorted(os.listdir(os.path.join(fpath, kind))) == ["00.npy", "01.npy"]
    assert np.load(os.path.join(fpath, "samples", "01.npy")).shape == (T, 3, d)
    merged = results.merge_ranks(fpath, 2)
    for name, m, k in zip(("samples", "accepted", "probs", "counts"), merged, range(4)):
        on_disk = np.load(os.path.join(fpath, name + ".npy"))
        want = np.concatenate([b[k][burnin:] for b in blocks], axis=1)
        assert np.array_equal(on_disk, want) and np.array_equal(m, want)
    assert merged[0].shape == (T, 7, d) and merged[3].shape == (T, 7, 3)
    # the variant that removes the per-rank folders (test_drhmc_funnel.py:232-239)
    results.merge_ranks(fpath, 2, delete_rank_files=True)
    assert not os.path.exists(os.path.join(fpath, "samples")) and os.path.exists(os.path.join(fpath, "samples.npy"))


def test_gather_to_rank0_is_identity_without_a_process_group():
    x = np.arange(24.0).reshape(2, 3, 4)
    assert np.array_equal(results.gather_to_rank0(x), x)


def test_checks_json_schema_and_formulas(tmp_path):
    T, C, d = 50, 16, 6
    samples, accepted, _, counts = _fake_run(T, C, d, 3, 5)
    means, stds = results.funnel_reference_moments(d)
    assert means.shape == (d,) and stds[0] == 3.0 and abs(stds[1] - 9.4877) < 1e-3
    checks = results.make_checks(samples, accepted, counts, means, stds)
    assert set(checks) == {"ess", "mean", "std", "counts", "acceptance", "costV_g", "cost"}
    err = samples.mean(axis=0) - means
    ess = (stds / np.sqrt((err ** 2).mean(axis=0))) ** 2 * C
    assert np.allclose(checks["ess"], ess, rtol=1e-12)
    tot = counts.sum(axis=(0, 1))
    assert np.allclose(checks["costV_g"], tot[1] / ess) and np.allclose(checks["cost"], (tot[0] + tot[1]) / ess)
    assert checks["acceptance"] == (accepted == 1).mean()
    path = results.write_checks(str(tmp_path) + "/", checks)
    assert path.endswith("/checks.json")
    back = json.load(open(path))
    assert list(back) == sorted(back)                         # sort_keys
    assert isinstance(back["ess"], list) and len(back["ess"]) == d and isinstance(back["acceptance"], float)
    assert back["counts"] == [float(v) for v in tot]


def test_get_cdf_of_the_reference_sample_itself_is_uniform():
    rng = np.random.default_rng(3)
    xref = rng.standard_normal((4000, 3)) * np.array([1.0, 5.0, 0.1])
    cdf, counts = results.get_cdf(xref, xref, nbins=20)
    assert cdf.shape == (3, 20) and counts.shape == (3, 20)
    assert np.all(counts == 200) and np.allclose(cdf[:, -1], 1.0)
    assert results.get_cdf(xref) is None                       # neither xref nor quantiles


def test_clean_samples_thins_by_autocorrelation_length():
    x = _ar1_chains(600, 4, 2, 0.9, 1)
    thinned, rc, tc = results.clean_samples(x)
    assert rc.shape == x.shape and tc.shape == (4, 2)
    subs = tc.max(axis=1)
    assert np.all(subs > 5)                                    # tau of an AR(1) with rho = .9 is 19
    assert thinned.shape == (sum(len(range(0, 600, int(s))) for s in subs), 2)
    capped, _, _ = results.clean_samples(x, maxv=4)
    assert capped.shape == (4 * 150, 2)


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")
def test_post_processing_matches_the_reference_utils():
    ref = ref_loader.load("utils")
    x = _ar1_chains(400, 5, 3, 0.8, 2)
    xref = np.random.default_rng(9).standard_normal((3000, 3))
    flat = x.reshape(-1, 3)
    got, want = results.get_cdf(flat, xref, nbins=15), ref.get_cdf(flat, xref, nbins=15)
    assert np.array_equal(got[1], want[1]) and np.allclose(got[0], want[0], rtol=0, atol=1e-15)
    got, want = results.get_cdf(flat, xref, nbins=10, xmin=-50, xmax=50), ref.get_cdf(flat, xref, nbins=10, xmin=-50, xmax=50)
    assert np.array_equal(got[1], want[1])
    got, want = results.clean_samples(x), ref.clean_samples(x)
    assert np.array_equal(got[2], want[2]) and np.allclose(got[0], want[0]) and np.allclose(got[1], want[1], atol=1e-10)
    got, want = results.clean_samples(x, maxv=3, fidsub=2), ref.clean_samples(x, maxv=3, fidsub=2)
    assert got[0].shape == want[0].shape and np.allclose(got[0], want[0])
    if not hasattr(ref, "ks_2samp"):
        # src/utils.py calls ks_2samp without importing it; give the module the scipy function it means
        from scipy.stats import ks_2samp
        ref.ks_2samp = ks_2samp
    got, want = results.get_ks(flat, xref), ref.get_ks(flat, xref)
    assert np.allclose(np.array(got[0], dtype=float), np.array(want[0], dtype=float))
    assert np.allclose(np.array(got[1], dtype=float), np.array(want[1], dtype=float))
