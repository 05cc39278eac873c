"""Import the REAL reference sampler modules when the tree is present (oracle
side; test infrastructure).

Search order: $HMC_REF_SRC, /root/reference/src.  The GPU box has neither, so
everything that runs there (``-m gpu`` tests, smoke(), bench.py) must work from
oracle/samplers.py and tests/golden/ alone; ``available()`` is how callers ask.
Nothing is copied from the reference and no bytecode is written into its tree.
"""
import importlib
import os
import sys
import warnings

_CANDIDATES = [os.environ.get("HMC_REF_SRC"), "/root/reference/src"]
_cache = {}


def ref_src_dir():
    for c in _CANDIDATES:
        if c and os.path.isfile(os.path.join(c, "pyhmc.py")):
            return c
    return None


def available():
    return ref_src_dir() is not None


def load(name):
    """Return reference module ``name`` in {'pyhmc','algorithms','mala','adaptivehmc','utils'}."""
    if name in _cache:
        return _cache[name]
    src = ref_src_dir()
    if src is None:
        raise ImportError("reference tree not present (looked in %r)" % (_CANDIDATES,))
    old_flag = sys.dont_write_bytecode
    sys.dont_write_bytecode = True
    sys.path.insert(0, src)
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")  # SyntaxWarning for '\p' escapes in docstrings
            spec = importlib.util.spec_from_file_location("_hmc_reference_" + name, os.path.join(src, name + ".py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
    finally:
        sys.path.remove(src)
        sys.dont_write_bytecode = old_flag
    _cache[name] = mod
    return mod
