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
