"""hmc_b200 -- B200-native batched-chain HMC engine behind the sampler/model API
of modichirag/hmc (leapfrog + Metropolis for pyhmc.HMC/PyHMC, delayed-rejection
HMC, MALA, step-size adaptation).  Python/PyTorch here is plumbing (device
memory, streams, torch.distributed); the transitions run in hand-written sm_100a
CUDA kernels reached through the C ABI in include/hmcb.h (libhmcb.so).  There is
no CPU fallback: sampler calls raise if the library or a CUDA device is missing.
"""
from . import _lib
from .models import CorrGaussian, Funnel, LogisticRegression, Model, Rosenbrock  # noqa: F401


def version():
    """ABI version of the loaded libhmcb.so (loads it)."""
    return _lib.load().hmcb_version()
