"""CPU oracle for the batched-chain sampler hot path.  TEST INFRASTRUCTURE ONLY.

Nothing in ``hmc_b200/`` may import this package.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` use it, and there only as the checker or the timed CPU
baseline -- never as the thing shipped.

Contents
--------
models.py     numpy statements of the benchmark densities (funnel restated from
              the reference's Stan text scripts/models.py:9-19; the others are
              builder-defined because the reference has none).
samplers.py   numpy restatement of the reference's sampler classes
              (src/pyhmc.py, src/algorithms.py, src/mala.py,
              src/adaptivehmc.py, src/utils.py), each function citing the
              lines it follows, taking its random draws from an explicit
              ``Draws`` object instead of numpy's global generator.
replay.py     the ``Draws`` stream plus a context manager that patches
              ``np.random.normal/uniform`` so the UNMODIFIED reference consumes
              the same pre-generated draws.
ref_loader.py imports the real reference from /root/reference/src when it is
              present (this container only; it does not travel to the GPU box).
make_goldens.py  runs the REAL reference on replayed draws and writes
              tests/golden/*.npz (committed).

Parity pinning: the reference ships no golden vectors or known-answer tests
(SURVEY.md section 4).  The oracle is pinned instead against outputs of the
reference itself run in the build container: ``tests/test_oracle_vs_reference.py``
checks samplers.py bit-for-bit against the imported reference classes whenever
/root/reference exists, and the committed goldens were produced by the real
reference (make_goldens.py), so the GPU box checks against reference outputs
without needing the reference tree.  The Stan-evaluated model densities that
the reference's scripts use are NOT available anywhere (PyStan/BridgeStan
absent): for those the parity is unpinned, the funnel being restated from its
closed form.
"""
