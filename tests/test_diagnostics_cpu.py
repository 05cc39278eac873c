"""hmc_b200.diagnostics -- the reference's src/diagnostics.py entry points (plot_hist, plot_trace,
plot_scatter, plot_autorcc).  The numbers behind the figures are checked: the autocorrelation window
against the oracle and, where the reference tree is present, against the reference's own nested
get_rcc (lifted out of plot_autorcc's source, which cannot be imported without matplotlib)."""
import ast
import os

import numpy as np
import pytest

from hmc_b200 import diagnostics
from oracle import ref_loader, samplers


def _chains(T, C, d, rho, seed):
    rng = np.random.default_rng(seed)
    x = np.zeros((T, C, d))
    x[0] = rng.standard_normal((C, d))
    for t in range(1, T):
        x[t] = rho * x[t - 1] + np.sqrt(1 - rho * rho) * rng.standard_normal((C, d))
    return x


def _reference_get_rcc():
    """The function nested in plot_autorcc (src/diagnostics.py:69-80), compiled from the reference's text."""
    src = open(os.path.join(ref_loader.ref_src_dir(), "diagnostics.py")).read()
    tree = ast.parse(src)
    outer = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "plot_autorcc")
    inner = next(n for n in outer.body if isinstance(n, ast.FunctionDef) and n.name == "get_rcc")
    mod = ast.Module(body=[inner], type_ignores=[])
    ns = {"np": np}
    exec(compile(mod, "diagnostics.get_rcc", "exec"), ns)
    return ns["get_rcc"]


@pytest.mark.parametrize("T,rho", [(700, 0.9), (300, 0.5), (120, 0.97)])
def test_autocorr_window_matches_oracle(T, rho):
    x = _chains(T, 6, 1, rho, 3)[..., 0]
    rcc, tcc = diagnostics.autocorr_window(x)
    rcc_o, tcc_o = samplers.autocorr_time_window(x)
    assert np.allclose(rcc.numpy(), rcc_o, atol=1e-10)
    assert np.allclose(tcc.numpy(), tcc_o, rtol=1e-9)


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")
def test_autocorr_window_matches_the_reference_code():
    get_rcc = _reference_get_rcc()
    for T, rho in ((700, 0.9), (120, 0.97)):
        x = _chains(T, 5, 1, rho, 8)[..., 0]
        rcc_r, tcc_r = get_rcc(x)
        rcc_o, tcc_o = samplers.autocorr_time_window(x)          # pins the oracle too
        assert np.array_equal(rcc_r, rcc_o) and np.array_equal(tcc_r, tcc_o)
        rcc, tcc = diagnostics.autocorr_window(x)
        assert np.allclose(rcc.numpy(), rcc_r, atol=1e-10) and np.allclose(tcc.numpy(), tcc_r, rtol=1e-9)


def test_entry_points_write_the_numbers_behind_each_figure(tmp_path):
    x = _chains(400, 3, 4, 0.8, 5)
    fpath = str(tmp_path / "figs")
    diagnostics.plot_hist(x, fpath, sub=10)
    diagnostics.plot_trace(x, fpath)
    diagnostics.plot_scatter(x, fpath, sub=10)
    out = diagnostics.plot_autorcc(x, fpath)
    names = set(os.listdir(fpath))
    for n in ("hist_sigma.npz", "hist_alpha.npz", "trace_sigma.npz", "trace_alpha1.npz", "trace_alpha3.npz", "scatter.npz",
              "rcc_sigma.npz", "rcc_alpha.npz"):
        assert n in names, n
    h = np.load(os.path.join(fpath, "hist_sigma.npz"))["samples"]
    assert np.array_equal(h, x[::10, :, 0].flatten())
    sc = np.load(os.path.join(fpath, "scatter.npz"))
    assert sc["sigma"].shape == (40 * 3,) and sc["alpha"].shape == (40 * 3, 3)
    assert np.array_equal(np.load(os.path.join(fpath, "trace_alpha3.npz"))["trace"], x[..., 3])
    r = np.load(os.path.join(fpath, "rcc_alpha.npz"))
    assert r["rcc2"].shape == (400, 3) and r["tcc2"].shape == (3,)
    assert set(out) == {0, 1, 2, 3} and float(out[0][1].mean()) > 3          # tau of an AR(1) with rho = .8 is 9
