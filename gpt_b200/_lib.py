"""ctypes binding of libgptb200.so — one prototype per symbol declared in include/gptb200.h."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GPTB200_LIB: developer override used to time two builds of the library in the same process environment
LIB_PATH = os.environ.get("GPTB200_LIB") or os.path.join(_HERE, "lib", "libgptb200.so")

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)
c_int64_p = C.POINTER(C.c_int64)
c_float_p = C.POINTER(C.c_float)


class Config(C.Structure):
    _fields_ = [("D", C.c_int32), ("n", C.c_int32), ("r", C.c_int32), ("Q", C.c_int32),
                ("precision", C.c_int32), ("feature_kind", C.c_int32), ("device", C.c_int32),
                ("reserved", C.c_int32), ("seed", C.c_uint64)]


# status codes / enums (include/gptb200.h)
OK, ERR_ARG, ERR_SHAPE, ERR_STATE, ERR_CUDA, ERR_NCCL, ERR_NOMEM, GEODESIC_NAN, ERR_UNSUPPORTED = range(9)
PREC_FP64, PREC_TC3, PREC_TC1 = 0, 1, 2
FEAT_COS, FEAT_SINCOS = 0, 1
FLAG_LANGEVIN, FLAG_STIEFEL, FLAG_FIX_U, FLAG_RMSPROP, FLAG_CLS_SINGLE_UPDATE = 1, 2, 4, 8, 16
ALG_AUTO, ALG_DIRECT, ALG_NODES = 0, 1, 2
OPT_PIPE_GROUPS, OPT_PIPE_RESERVE, OPT_ALGORITHM, OPT_NODE_ORDER, OPT_NODE_THETA, OPT_RMS_ALPHA, OPT_RMS_LAMBDA, OPT_PRED_CHUNK_MB, \
    OPT_FRONT_PREFETCH, OPT_PRED_LAYOUT, OPT_PRED_WARPS = range(11)
PHASES = ("gather", "fwd", "contract", "bwd", "reduce", "update")

H = C.c_void_p
PROTOTYPES = {
    "gptb_version": (C.c_int, []),
    "gptb_create": (C.c_int, [C.POINTER(Config), C.POINTER(H)]),
    "gptb_destroy": (C.c_int, [H]),
    "gptb_last_error": (C.c_char_p, [H]),
    "gptb_set_stream": (C.c_int, [H, C.c_void_p]),
    "gptb_synchronize": (C.c_int, [H]),
    "gptb_set_features": (C.c_int, [H, c_double_p, c_double_p, c_double_p, C.c_int32, C.c_double, C.c_double]),
    "gptb_set_index": (C.c_int, [H, c_int32_p]),
    "gptb_set_train": (C.c_int, [H, c_double_p, c_double_p, C.c_int64]),
    "gptb_set_domain": (C.c_int, [H, c_double_p, c_double_p]),
    "gptb_set_train_phi": (C.c_int, [H, c_double_p, c_double_p, C.c_int64]),
    "gptb_set_params": (C.c_int, [H, c_double_p, c_double_p]),
    "gptb_get_params": (C.c_int, [H, c_double_p, c_double_p]),
    "gptb_init_params": (C.c_int, [H, C.c_uint64, C.c_int32]),
    "gptb_features": (C.c_int, [H, c_double_p, C.c_int64, c_double_p]),
    "gptb_pred": (C.c_int, [H, c_double_p, c_double_p, c_double_p, C.c_int64, c_double_p]),
    "gptb_pred_phi": (C.c_int, [H, c_double_p, c_double_p, c_double_p, C.c_int64, c_double_p]),
    "gptb_pred_samples": (C.c_int, [H, c_double_p, c_double_p, C.c_int32, c_double_p, C.c_int64, c_double_p, c_double_p]),
    "gptb_grad": (C.c_int, [H, C.c_double, c_int64_p, C.c_int64, C.c_int64, c_double_p, c_double_p, c_double_p, C.c_int32]),
    "gptb_proj": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, c_double_p, c_double_p, c_double_p]),
    "gptb_proj_noalloc": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gptb_geod": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, c_double_p, c_double_p, C.c_double, c_double_p]),
    "gptb_geodboth": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, c_double_p, c_double_p, C.c_double, c_double_p, c_double_p]),
    "gptb_gmc_run": (C.c_int, [H, C.c_double, C.c_double, C.c_double, C.c_int32, C.c_int32, C.c_int32, C.c_int32, c_double_p,
                               c_double_p, c_double_p, c_double_p, c_double_p, c_double_p]),
    "gptb_tgp_gibbs": (C.c_int, [H, C.c_double, C.c_int32, C.c_int32, c_double_p, c_double_p, c_double_p, c_double_p]),
    "gptb_classify_run": (C.c_int, [C.POINTER(H), C.c_int32, c_int32_p, C.c_int64, C.c_double, C.c_double, C.c_int32, C.c_int32,
                                    C.c_int32, c_int64_p, c_double_p, c_double_p, c_double_p, c_double_p]),
    "gptb_sgld_step": (C.c_int, [H, C.c_double, c_int64_p, C.c_int64, C.c_int64, C.c_double, C.c_double, C.c_int32,
                                 c_double_p, c_double_p]),
    "gptb_sgld_step_host": (C.c_int, [H, C.c_double, c_double_p, c_double_p, C.c_int64, C.c_int64, C.c_double,
                                      C.c_double, C.c_int32, c_double_p, c_double_p, c_double_p]),
    "gptb_sgld_step_host_submit": (C.c_int, [H, C.c_double, c_double_p, c_double_p, C.c_int64, C.c_int64, C.c_double,
                                      C.c_double, C.c_int32, c_double_p, c_double_p, c_double_p]),
    "gptb_sgld_step_host_wait": (C.c_int, [H]),
    "gptb_sgld_run": (C.c_int, [H, C.c_double, C.c_int64, C.c_double, C.c_double, C.c_int32, C.c_int32, C.c_int32,
                                c_int64_p, c_double_p, c_double_p, c_double_p, c_double_p]),
    "gptb_sgld_advance": (C.c_int, [H, C.c_double, C.c_int64, C.c_double, C.c_double, C.c_int32, C.c_int64]),
    "gptb_comm_unique_id": (C.c_int, [C.c_void_p]),
    "gptb_comm_init_local": (C.c_int, [C.POINTER(H), C.c_int32]),
    "gptb_multi_sgld_run": (C.c_int, [C.POINTER(H), C.c_int32, C.c_double, C.c_int64, C.c_double, C.c_double, C.c_int32, C.c_int32,
                                      C.c_int32, c_int64_p, c_double_p, c_double_p, c_double_p, c_double_p]),
    "gptb_multi_sgld_advance": (C.c_int, [C.POINTER(H), C.c_int32, C.c_double, C.c_int64, C.c_double, C.c_double, C.c_int32, C.c_int64]),
    "gptb_multi_pred_samples": (C.c_int, [C.POINTER(H), C.c_int32, c_double_p, c_double_p, C.c_int32, c_double_p, C.c_int64,
                                          c_double_p, c_double_p]),
    "gptb_comm_init": (C.c_int, [H, C.c_int32, C.c_int32, C.c_void_p]),
    "gptb_comm_destroy": (C.c_int, [H]),
    # full-theta ("no tensor") RFF model
    "gptb_nt_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_uint64, C.POINTER(H)]),
    "gptb_nt_destroy": (C.c_int, [H]),
    "gptb_nt_last_error": (C.c_char_p, [H]),
    "gptb_nt_set_features": (C.c_int, [H, c_double_p, c_double_p, c_double_p, C.c_int32, C.c_double]),
    "gptb_nt_features": (C.c_int, [H, c_double_p, C.c_int64, c_double_p]),
    "gptb_nt_set_train": (C.c_int, [H, c_double_p, c_double_p, C.c_int64]),
    "gptb_nt_pred": (C.c_int, [H, c_double_p, c_double_p, C.c_int64, c_double_p]),
    "gptb_nt_run": (C.c_int, [H, C.c_double, C.c_double, C.c_int64, C.c_double, C.c_double, C.c_int32, C.c_int32,
                              C.c_uint64, c_double_p, c_int64_p, c_double_p, c_double_p]),
    "gptb_nt_last_run_stats": (C.c_int, [H, c_double_p, c_int64_p, c_int64_p]),
    "gptb_set_profiling": (C.c_int, [H, C.c_int32]),
    "gptb_set_option": (C.c_int, [H, C.c_int32, C.c_double]),
    "gptb_get_phase_ms": (C.c_int, [H, c_float_p, c_int64_p, C.c_int32]),
}

_lib = None


def load():
    """Loads libgptb200.so (built by build.sh / __graft_entry__.build()).  Fails loudly if it is missing:
    there is no Python or CPU fallback for any compute entry."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OSError(f"{LIB_PATH} not found — run ./build.sh (or __graft_entry__.build()); "
                          "gpt_b200 has no CPU fallback")
        lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
