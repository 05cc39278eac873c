"""The C-ABI library loads and exports every symbol include/hmcb.h declares
(no compute calls: this runs without a GPU)."""
import os
import re

import pytest

from hmc_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "hmcb.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hmcb_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert _declared_symbols() == sorted(_lib.SIGNATURES)


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    for name in _declared_symbols():
        assert hasattr(lib, name), name
    assert lib.hmcb_version() == 1


def test_argument_errors_do_not_need_a_gpu():
    import ctypes as C
    lib = _lib.load()
    m = _lib.Model()
    m.kind, m.d, m.dtype = _lib.MODEL_FUNNEL, 1, _lib.F64
    m.par[0] = 3.0
    ch, r, so = _lib.Chains(), _lib.Rng(), _lib.StepOut()
    rc = lib.hmcb_hmc_step(C.byref(m), C.byref(ch), C.byref(r), 5, 0.1, 0, C.byref(so), None)
    assert rc == _lib.E_ARG and b"funnel" in lib.hmcb_last_error()
    m.d, m.kind = 10, 99
    rc = lib.hmcb_hmc_step(C.byref(m), C.byref(ch), C.byref(r), 5, 0.1, 0, C.byref(so), None)
    assert rc == _lib.E_UNSUPPORTED


def test_opaque_callables_are_rejected():
    from hmc_b200 import algorithms, models, pyhmc
    with pytest.raises(TypeError):
        pyhmc.HMC(lambda x: 0.0, lambda x: x)
    m = models.Funnel(10)
    assert pyhmc.HMC(m.log_prob, m.grad_log_prob).model is m
    assert algorithms.DRHMC(m.log_prob, log_prob_and_grad=m.log_prob_and_grad).model is m
    with pytest.raises(TypeError):
        pyhmc.HMC(m.log_prob, models.Funnel(10).grad_log_prob)  # two different models


def test_no_cpu_fallback():
    import numpy as np
    import torch
    from hmc_b200 import models, pyhmc
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(_lib.HmcbError):
        pyhmc.HMC(models.Funnel(10)).step(np.zeros(10), 5, 0.1)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "hmc_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f


def test_argument_checks_of_the_later_entry_points():
    """Bad arguments are refused before any CUDA call, with a message that names the argument (runs without a GPU)."""
    import ctypes as C
    import numpy as np
    lib = _lib.load()
    m = _lib.Model()
    m.kind, m.d, m.dtype = _lib.MODEL_FUNNEL, 10, _lib.F64
    m.par[0] = 3.0
    ch, r, so = _lib.Chains(), _lib.Rng(), _lib.StepOut()          # n_chains = 0: nothing would run anyway
    # delayed rejection: stage count, convention
    rc = lib.hmcb_drhmc_step(C.byref(m), C.byref(ch), C.byref(r), 10, 0.1, 7, 2.0, 0, C.byref(so), None, 0, None, None)
    assert rc == _lib.E_ARG and b"nprops" in lib.hmcb_last_error()
    rc = lib.hmcb_drhmc_step(C.byref(m), C.byref(ch), C.byref(r), 10, 0.1, 3, 2.0, 9, C.byref(so), None, 0, None, None)
    assert rc == _lib.E_ARG and b"convention" in lib.hmcb_last_error()
    # injected uniforms: the probabilistic conventions read 2*nprops-1 per chain
    u = np.zeros(8)
    r2 = _lib.Rng()
    r2.u_inject, r2.u_stride = u.ctypes.data, 3
    rc = lib.hmcb_drhmc_step(C.byref(m), C.byref(ch), C.byref(r2), 10, 0.1, 3, 2.0, _lib.DR_PROB_METROP, C.byref(so), None, 0,
                             None, None)
    assert rc == _lib.E_ARG and b"u_stride >= 5" in lib.hmcb_last_error()
    # step-size walk: segments, ratios closed under reciprocal, eps0: exactly one of n_leap / tint
    steps = np.logspace(-2, 0, 6)
    par, wo = _lib.WalkPar(), _lib.WalkOut()
    par.n_segments, par.n_leap_segment, par.n_ratios, par.ntry, par.nsteps = 1, 5, 2, 6, 1
    par.ratios[0], par.ratios[1], par.pwts0[0], par.pwts0[1] = 2 ** -0.5, 2 ** 0.5, 1.0, 1.0
    par.smin, par.smax = 0.01, 1.0
    par.steps = steps.ctypes.data_as(C.POINTER(C.c_double))
    rc = lib.hmcb_adhmc_walk_step(C.byref(m), C.byref(ch), C.byref(r), C.byref(par), C.byref(wo), None, 0, None)
    assert rc == _lib.E_ARG and b"n_segments" in lib.hmcb_last_error()
    par.n_segments = 4
    par.ratios[1] = 3.0
    rc = lib.hmcb_adhmc_walk_step(C.byref(m), C.byref(ch), C.byref(r), C.byref(par), C.byref(wo), None, 0, None)
    assert rc == _lib.E_ARG and b"reciprocal" in lib.hmcb_last_error()
    ao = _lib.AdhmcOut()
    rc = lib.hmcb_adhmc_eps0_step(C.byref(m), C.byref(ch), C.byref(r), 10, 2.0, steps.ctypes.data_as(C.POINTER(C.c_double)),
                                  6, 1, C.byref(ao), None, 0, None)
    assert rc == _lib.E_ARG and b"either n_leap" in lib.hmcb_last_error()
    # with valid arguments and no chains the calls are no-ops
    par.ratios[1] = 2 ** 0.5
    assert lib.hmcb_adhmc_walk_step(C.byref(m), C.byref(ch), C.byref(r), C.byref(par), C.byref(wo), None, 0, None) == 0
    assert lib.hmcb_drhmc_step(C.byref(m), C.byref(ch), C.byref(r), 10, 0.1, 3, 2.0, 3, C.byref(so), None, 0, None, None) == 0
