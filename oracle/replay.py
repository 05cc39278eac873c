"""Draw replay (oracle side; test infrastructure).

"Fed the same momentum and uniform draws" needs a contract for WHICH draw goes
where.  The reference pulls from numpy's global generator in a fixed order per
transition:

  pyhmc.HMC.step       normal(size=d)  then  uniform(0,1,size=1) iff prob is finite and q moved
                       (src/pyhmc.py:66,59)
  algorithms.HMC.step  normal(size=d)  then  uniform(0,1,size=1) always (src/algorithms.py:195,184)
  DRHMC.step           normal(size=d)  then  one bare uniform() per executed stage
                       (src/algorithms.py:506,527)
  MALA.step            normal(size=d)  then  conditional uniform(0,1,size=1) (src/mala.py:47,40)
  AdHMC_eps0.hmc_step  normal(size=d), uniform(low=cdf.min()), uniform(lo,hi), conditional
                       uniform(0,1,size=1) (src/adaptivehmc.py:162,127,129,190)
  AdHMC(_tint).hmc_step normal(size=d), uniform(low=cdf.min()), uniform(lo,hi), one
                       choice(nr, p=pf) per later segment, conditional uniform(0,1,size=1)
                       (src/adaptivehmc.py:259,219,221,303,350)

``Draws`` holds one stream of standard-normal vectors and one stream of U[0,1)
scalars and hands them out in call order; ``low``/``high`` are applied the way
numpy applies them (``low + (high-low)*u``).  ``patched_numpy_random`` swaps
``np.random.normal`` / ``np.random.uniform`` for a ``Draws`` so the unmodified
reference classes consume exactly those values.  The GPU path receives the
same arrays through its ``p=`` / ``u=`` arguments.
"""
import contextlib

import numpy as np


class Draws:
    def __init__(self, normals, uniforms):
        # normals: sequence of 1-D float64 arrays; uniforms: 1-D sequence of floats in [0,1)
        self._normals = [np.asarray(z, dtype=np.float64) for z in normals]
        self._uniforms = [float(u) for u in np.asarray(uniforms, dtype=np.float64).ravel()]
        self.n_normal = 0
        self.n_uniform = 0

    def normal(self, loc=0.0, scale=1.0, size=None):
        z = self._normals[self.n_normal]
        self.n_normal += 1
        n = 1 if size is None else int(np.prod(size))
        assert z.size == n, "draw replay: normal(size=%r) but next stored draw has %d entries" % (size, z.size)
        out = loc + scale * z if (loc != 0.0 or scale != 1.0) else z.copy()
        if size is None:
            return float(out[0])
        return out.reshape(size)

    def uniform(self, low=0.0, high=1.0, size=None):
        u = self._uniforms[self.n_uniform]
        self.n_uniform += 1
        val = low + (high - low) * u
        if size is None:
            return val
        assert int(np.prod(size)) == 1, "draw replay: only scalar uniforms are replayed"
        return np.full(size, val, dtype=np.float64)


    def choice(self, a, p=None):
        """np.random.choice(n, p=p) for one draw, as numpy's legacy generator does it
        (numpy/random/mtrand.pyx, RandomState.choice with replace=True): cdf = cumsum(p),
        cdf /= cdf[-1], index = searchsorted(cdf, u, side='right') with one U[0,1) draw."""
        u = self._uniforms[self.n_uniform]
        self.n_uniform += 1
        cdf = np.asarray(p, dtype=np.float64).cumsum()
        cdf /= cdf[-1]
        return int(cdf.searchsorted(u, side="right"))


@contextlib.contextmanager
def patched_numpy_random(draws):
    """Route np.random.normal / np.random.uniform / np.random.choice to ``draws`` inside the block."""
    saved = (np.random.normal, np.random.uniform, np.random.choice)
    np.random.normal, np.random.uniform, np.random.choice = draws.normal, draws.uniform, draws.choice
    try:
        yield draws
    finally:
        np.random.normal, np.random.uniform, np.random.choice = saved


def make_draws(seed, T, C, d, n_uniform):
    """P[T,C,d] standard normals and U[T,C,n_uniform] uniforms from PCG64(seed)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    P = rng.standard_normal((T, C, d))
    U = rng.random((T, C, n_uniform))
    return P, U
