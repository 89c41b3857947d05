"""ctypes binding of include/psgd_b200.h.  There is no fallback: if the library cannot be
loaded the import of this module raises, and every compute entry point needs a CUDA device."""
import ctypes
import os

from .build import LIB_PATH

c_i64p = ctypes.POINTER(ctypes.c_int64)
c_i32p = ctypes.POINTER(ctypes.c_int32)
c_f64p = ctypes.POINTER(ctypes.c_double)
c_vpp = ctypes.POINTER(ctypes.c_void_p)
vp = ctypes.c_void_p

N_PHASES = 7
ABI_VERSION = 2  # PSGD_ABI_VERSION of include/psgd_b200.h
PHASE_NAMES = ["pack", "p_from_m", "orthogonalize", "q_from_mt", "q_reduce", "finalize", "finish"]
PHASE_PACK, PHASE_P_FROM_M, PHASE_ORTH, PHASE_Q_FROM_MT, PHASE_Q_REDUCE, PHASE_FINALIZE, PHASE_FINISH = range(7)

# name -> (restype, argtypes); every name here must be declared in include/psgd_b200.h
SIGNATURES = {
    "psgd_abi_version": (ctypes.c_int, []),
    "psgd_last_error": (ctypes.c_char_p, []),
    "psgd_plan_create": (ctypes.c_int, [ctypes.c_int, c_i64p, c_i64p, ctypes.c_int, ctypes.c_int, c_i64p,
                                        ctypes.c_int, ctypes.POINTER(vp)]),
    "psgd_plan_destroy": (ctypes.c_int, [vp]),
    "psgd_p_elems": (ctypes.c_int64, [vp]),
    "psgd_q_elems": (ctypes.c_int64, [vp]),
    "psgd_vec_elems": (ctypes.c_int64, [vp]),
    "psgd_workspace_bytes": (ctypes.c_int64, [vp]),
    "psgd_matrix_info": (ctypes.c_int, [vp, ctypes.c_int, c_i64p]),
    "psgd_tile_count": (ctypes.c_int, [vp, ctypes.c_int]),
    "psgd_tile_get": (ctypes.c_int, [vp, ctypes.c_int, ctypes.c_int, c_i32p]),
    "psgd_set_ptrs": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "psgd_set_ptrs_fused": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "psgd_pack_vec": (ctypes.c_int, [vp, vp, vp]),
    "psgd_pack_vec_fused": (ctypes.c_int, [vp, vp, vp]),
    "psgd_p_from_m": (ctypes.c_int, [vp, vp, vp, ctypes.c_int, ctypes.c_float, vp]),
    "psgd_p_from_m_pack": (ctypes.c_int, [vp, vp, vp, ctypes.c_int, ctypes.c_float, vp, vp]),
    "psgd_p_from_m_orth": (ctypes.c_int, [vp, vp, vp, ctypes.c_int, ctypes.c_float, vp, vp]),
    "psgd_finalize_unpack": (ctypes.c_int, [vp, vp, vp, ctypes.c_float, vp, vp]),
    "psgd_orthogonalize": (ctypes.c_int, [vp, vp, ctypes.c_float, vp]),
    "psgd_q_from_mt": (ctypes.c_int, [vp, vp, vp, vp, vp]),
    "psgd_finalize": (ctypes.c_int, [vp, vp, vp, ctypes.c_float, vp]),
    "psgd_finish": (ctypes.c_int, [vp, vp, vp, ctypes.c_float, vp]),
    "psgd_timing_enable": (ctypes.c_int, [vp, ctypes.c_int]),
    "psgd_timing_read": (ctypes.c_int, [vp, c_f64p, c_i64p]),
    "psgd_launch_count": (ctypes.c_int64, []),
    "psgd_debug_read": (ctypes.c_int, [vp, ctypes.POINTER(ctypes.c_uint64), ctypes.c_int]),
    "psgd_tb_create": (ctypes.c_int, [ctypes.c_int, c_i64p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(vp)]),
    "psgd_tb_destroy": (ctypes.c_int, [vp]),
    "psgd_tb_pack": (ctypes.c_int, [vp, vp, vp, vp]),
    "psgd_tb_unpack": (ctypes.c_int, [vp, vp, vp, vp]),
    "psgd_sgd_create": (ctypes.c_int, [ctypes.c_int, c_i64p, ctypes.c_int, ctypes.POINTER(vp)]),
    "psgd_sgd_destroy": (ctypes.c_int, [vp]),
    "psgd_sgd_step": (ctypes.c_int, [vp, vp, vp, vp, ctypes.c_float, ctypes.c_float, ctypes.c_int, vp]),
    "psgd_peer_allreduce": (ctypes.c_int, [vp, vp, vp, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                           ctypes.c_uint32, ctypes.c_int, vp]),
    "psgd_peer_allreduce_mc": (ctypes.c_int, [vp, vp, vp, vp, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_uint32, ctypes.c_float, ctypes.c_int64, vp, ctypes.c_int, vp]),
    "psgd_peer_create": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, vp, vp, vp, vp, vp, ctypes.c_int,
                                        ctypes.c_int, ctypes.POINTER(vp)]),
    "psgd_peer_create_push": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, vp, ctypes.c_int64, vp, vp, vp,
                                             ctypes.c_int64, vp, ctypes.c_int, ctypes.POINTER(vp)]),
    "psgd_p_from_m_push": (ctypes.c_int, [vp, vp, vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp]),
    "psgd_q_from_mt_push": (ctypes.c_int, [vp, vp, vp, vp, vp]),
    "psgd_peer_destroy": (ctypes.c_int, [vp]),
    "psgd_orthogonalize_peer": (ctypes.c_int, [vp, vp, vp, vp, vp]),
    "psgd_finalize_peer": (ctypes.c_int, [vp, vp, vp, vp, vp, vp]),
    "psgd_p_from_m_peer": (ctypes.c_int, [vp, vp, vp, vp, ctypes.c_int, vp]),
    "psgd_orthogonalize_one": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_int, ctypes.c_int, vp]),
    "psgd_q_from_mt_noalloc": (ctypes.c_int, [vp, vp, vp, ctypes.c_int64, ctypes.c_int64, ctypes.c_int, ctypes.c_int, vp]),
}

_lib = None


class PsgdError(RuntimeError):
    pass


def load():
    """dlopen the in-tree library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("PSGD_LIB_PATH", LIB_PATH)  # development aid: A/B two builds in one process tree
    if not os.path.exists(path):
        raise PsgdError(
            f"{LIB_PATH} is missing: build it with `python -m network_distributed_pytorch_b200.build` "
            "(or __graft_entry__.build()).  There is no CPU or PyTorch fallback for the reducer hot path.")
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    if lib.psgd_abi_version() != ABI_VERSION:
        raise PsgdError("libpsgd_b200.so has an unexpected ABI version")
    _lib = lib
    return lib


def check(rc, lib=None):
    if rc != 0:
        lib = lib or load()
        raise PsgdError(lib.psgd_last_error().decode("utf-8", "replace"))


def i64_array(values):
    arr = (ctypes.c_int64 * max(1, len(values)))(*values)
    return arr


def ptr_array(values):
    return (ctypes.c_void_p * max(1, len(values)))(*values)
