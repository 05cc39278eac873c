"""ctypes binding of libhmcb.so (include/hmcb.h).  Loading fails loudly: the
product path has no CPU fallback."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhmcb.so")

OK, E_ARG, E_UNSUPPORTED, E_CUDA, E_WORKSPACE = 0, -1, -2, -3, -4
F64, F32 = 0, 1
MODEL_FUNNEL, MODEL_GAUSSIAN, MODEL_ROSENBROCK, MODEL_LOGISTIC = 0, 1, 2, 3
RULE_PYHMC, RULE_ALG = 0, 1
DR_ALGORITHMS, DR_MULTISTEP, DR_PROB_METROP, DR_PROB_HASTINGS = 0, 1, 2, 3
RNG_NORMAL_F64 = 1
MAX_PROPS = 6
STATIC_SCRATCH_BYTES = 32 << 20


class Model(C.Structure):
    _fields_ = [("kind", C.c_int32), ("d", C.c_int32), ("dtype", C.c_int32), ("n_data", C.c_int32),
                ("par", C.c_double * 4), ("dev0", C.c_void_p), ("dev1", C.c_void_p), ("invmetric_diag", C.c_void_p)]


class Chains(C.Structure):
    _fields_ = [("n_chains", C.c_int64), ("chain_offset", C.c_int64), ("q_in", C.c_void_p), ("q_out", C.c_void_p),
                ("stride_chain", C.c_int64), ("stride_dim", C.c_int64)]


class Rng(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("iteration", C.c_uint64), ("p_inject", C.c_void_p), ("u_inject", C.c_void_p),
                ("u_stride", C.c_int32), ("flags", C.c_int32), ("p_prev", C.c_void_p), ("alpha", C.c_double)]


class GhmcIO(C.Structure):
    _fields_ = [("p_out", C.c_void_p), ("V_in", C.c_void_p), ("Vg_in", C.c_void_p), ("V_out", C.c_void_p),
                ("Vg_out", C.c_void_p), ("u_in", C.c_void_p), ("u_out", C.c_void_p), ("delta", C.c_double),
                ("code", C.c_void_p), ("energies", C.c_void_p), ("counts", C.c_void_p)]


class Profile(C.Structure):
    _fields_ = [("events", C.c_void_p * (MAX_PROPS + 1))]


class StepOut(C.Structure):
    _fields_ = [("p_out", C.c_void_p), ("code", C.c_void_p), ("energies", C.c_void_p), ("avec", C.c_void_p),
                ("counts", C.c_void_p), ("accept_prob", C.c_void_p), ("ptries", C.c_void_p), ("p0_out", C.c_void_p)]


class SampleIO(C.Structure):
    _fields_ = [("n_transitions", C.c_int32), ("moment_sum", C.c_void_p), ("moment_sumsq", C.c_void_p),
                ("n_accept", C.c_void_p)]


class MalaIO(C.Structure):
    _fields_ = [("V_in", C.c_void_p), ("Vg_in", C.c_void_p), ("V_out", C.c_void_p), ("Vg_out", C.c_void_p),
                ("code", C.c_void_p), ("n_accept", C.c_void_p), ("counts", C.c_void_p)]


class WalkPar(C.Structure):
    _fields_ = [("n_segments", C.c_int32), ("n_leap_segment", C.c_int32), ("tint", C.c_double), ("n_ratios", C.c_int32),
                ("ntry", C.c_int32), ("nsteps", C.c_int32), ("ratios", C.c_double * 4), ("pwts0", C.c_double * 4),
                ("smin", C.c_double), ("smax", C.c_double), ("steps", C.POINTER(C.c_double))]


class WalkOut(C.Structure):
    _fields_ = [("p_out", C.c_void_p), ("code", C.c_void_p), ("probs", C.c_void_p), ("checks", C.c_void_p),
                ("counts", C.c_void_p)]


class AdhmcOut(C.Structure):
    _fields_ = [("p_out", C.c_void_p), ("code", C.c_void_p), ("probs", C.c_void_p), ("aux", C.c_void_p),
                ("counts", C.c_void_p)]


# every symbol include/hmcb.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "hmcb_version": (C.c_int, []),
    "hmcb_last_error": (C.c_char_p, []),
    "hmcb_philox4x32_10": (None, [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "hmcb_philox_block": (None, [C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]),
    "hmcb_u53": (C.c_double, [C.c_uint32, C.c_uint32]),
    "hmcb_hmc_step": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.POINTER(Rng), C.c_int32, C.c_double, C.c_int32,
                                C.POINTER(StepOut), C.c_void_p]),
    "hmcb_leapfrog": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.c_void_p, C.c_void_p, C.c_int32, C.c_double,
                                C.c_void_p]),
    "hmcb_logp_grad": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.c_void_p, C.c_void_p, C.c_void_p]),
    "hmcb_drhmc_workspace_bytes": (C.c_size_t, [C.POINTER(Model), C.c_int64, C.c_int32]),
    "hmcb_drhmc_step": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.POINTER(Rng), C.c_int32, C.c_double,
                                  C.c_int32, C.c_double, C.c_int32, C.POINTER(StepOut), C.c_void_p, C.c_size_t,
                                  C.POINTER(Profile), C.c_void_p]),
    "hmcb_drhmc_sample_workspace_bytes": (C.c_size_t, [C.POINTER(Model), C.c_int64, C.c_int32]),
    "hmcb_drhmc_sample": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.POINTER(Rng), C.c_int32, C.c_double,
                                    C.c_int32, C.c_double, C.c_int32, C.POINTER(SampleIO), C.POINTER(StepOut),
                                    C.c_void_p, C.c_size_t, C.c_void_p]),
    "hmcb_hmc_sample": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.POINTER(Rng), C.c_int32, C.c_double, C.c_int32,
                                  C.POINTER(SampleIO), C.POINTER(StepOut), C.c_void_p]),
    "hmcb_drhmc_step_static": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.POINTER(Rng), C.c_int32, C.c_double,
                                         C.c_int32, C.c_double, C.c_int32, C.POINTER(StepOut), C.c_void_p]),
    "hmcb_moments_accumulate": (C.c_int, [C.c_int32, C.c_int64, C.c_int32, C.c_void_p, C.c_int64, C.c_int64,
                                          C.c_void_p, C.c_void_p, C.c_void_p]),
    "hmcb_moments_reduce": (C.c_int, [C.c_int32, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_double,
                                      C.c_void_p, C.c_void_p, C.c_void_p]),
    "hmcb_mala_step": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.POINTER(Rng), C.c_double, C.c_int32,
                                 C.POINTER(MalaIO), C.c_void_p]),
    "hmcb_ghmc_step": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.POINTER(Rng), C.c_int32, C.c_double,
                                 C.POINTER(GhmcIO), C.c_void_p]),
    "hmcb_ghmc_sample": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.POINTER(Rng), C.c_int32, C.c_double,
                                   C.POINTER(GhmcIO), C.c_int32, C.c_void_p, C.c_void_p]),
    "hmcb_adhmc_eps0_workspace_bytes": (C.c_size_t, [C.POINTER(Model), C.c_int64]),
    "hmcb_adhmc_eps0_step": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.POINTER(Rng), C.c_int32, C.c_double,
                                       C.POINTER(C.c_double), C.c_int32, C.c_int32, C.POINTER(AdhmcOut), C.c_void_p,
                                       C.c_size_t, C.c_void_p]),
    "hmcb_adhmc_walk_workspace_bytes": (C.c_size_t, [C.POINTER(Model), C.c_int64, C.c_int32]),
    "hmcb_adhmc_walk_step": (C.c_int, [C.POINTER(Model), C.POINTER(Chains), C.POINTER(Rng), C.POINTER(WalkPar),
                                       C.POINTER(WalkOut), C.c_void_p, C.c_size_t, C.c_void_p]),
    "hmcb_logreg_prepared_bytes": (C.c_size_t, [C.POINTER(Model)]),
    "hmcb_logreg_prepare": (C.c_int, [C.POINTER(Model), C.c_void_p, C.c_size_t, C.c_void_p]),
    "hmcb_logreg_workspace_bytes": (C.c_size_t, [C.POINTER(Model), C.c_int64]),
    "hmcb_logreg_hmc_step": (C.c_int, [C.POINTER(Model), C.c_void_p, C.POINTER(Chains), C.POINTER(Rng), C.c_int32,
                                       C.c_double, C.c_int32, C.POINTER(StepOut), C.c_void_p, C.c_size_t, C.c_void_p]),
    "hmcb_logreg_logp_grad": (C.c_int, [C.POINTER(Model), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_size_t, C.c_void_p]),
    "hmcb_logreg_mala_step": (C.c_int, [C.POINTER(Model), C.c_void_p, C.POINTER(Chains), C.POINTER(Rng), C.c_double, C.c_int32,
                                        C.POINTER(MalaIO), C.c_void_p, C.c_size_t, C.c_void_p]),
    "hmcb_probe_normals": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_void_p, C.POINTER(C.c_double),
                                     C.c_void_p]),
    "hmcb_probe_fma": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_void_p, C.POINTER(C.c_double),
                                 C.c_void_p]),
}

_lib = None


class HmcbError(RuntimeError):
    pass


def load():
    """Load libhmcb.so and set prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise HmcbError("libhmcb.so not found at %s -- build it with `python __graft_entry__.py` "
                        "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().hmcb_last_error()
        raise HmcbError("libhmcb error %d: %s" % (rc, msg.decode() if msg else "?"))


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())
