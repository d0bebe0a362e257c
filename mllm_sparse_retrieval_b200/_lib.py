"""ctypes binding of libmsr.so (include/msr.h).  Fails loudly: there is no CPU or PyTorch fallback."""
from __future__ import annotations

import ctypes
import os

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libmsr.so")
LIB_PATH = os.environ.get("MSR_LIB_PATH", LIB_PATH)      # developer knob: a variant build of the same library

MSR_HOST, MSR_DEVICE = 0, 1
MSR_MAX_K = 4096
MSR_ERR_RANGE = -3

_lib = None


class MsrError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libmsr error {code}: {msg}")
        self.code = code


def _declare(lib):
    P, I, U32, U64, F = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_float
    D = ctypes.c_double
    sigs = {
        "msr_version": (I, []),
        "msr_last_error": (ctypes.c_char_p, []),
        "msr_kernel_launches": (U64, []),
        "msr_sparse_index_build": (I, [I, U32, U32, P, P, P, U32, P, ctypes.c_int32, U32, I, P, ctypes.POINTER(P)]),
        "msr_sparse_index_destroy": (I, [P]),
        "msr_sparse_index_info": (I, [P, P]),
        "msr_sparse_search": (I, [P, U32, P, P, P, U32, P, P, P, I, P]),
        "msr_sparse_index_status": (I, [P, P]),
        "msr_sparse_query_work": (I, [P, U32, P, P, P, ctypes.POINTER(U64), ctypes.POINTER(U64), I, P]),
        "msr_sparse_query_traffic": (I, [P, U32, P, P, P, ctypes.POINTER(U64), P]),
        "msr_microbench": (I, [I, I, U64, U32, ctypes.POINTER(D), P]),
        "msr_sparse_profile": (I, [P, I, ctypes.POINTER(D), ctypes.POINTER(D), ctypes.POINTER(U64), P]),
        "msr_topk_merge": (I, [I, U32, U32, U32, P, P, P, U32, P, P, P, P]),
        "msr_topk_rows": (I, [I, U32, U32, P, ctypes.c_size_t, U32, ctypes.c_int32, P, P, P, P]),
        "msr_dense_index_create": (I, [I, U32, ctypes.c_int32, ctypes.POINTER(P)]),
        "msr_dense_index_add": (I, [P, U32, P, I, P]),
        "msr_dense_index_destroy": (I, [P]),
        "msr_dense_index_info": (I, [P, P]),
        "msr_dense_search": (I, [P, U32, P, U32, P, P, I, I, P]),
        "msr_dense_profile": (I, [P, I, ctypes.POINTER(D), ctypes.POINTER(D), ctypes.POINTER(U64)]),
        "msr_dense_last_fallback": (I, [P, ctypes.POINTER(U32)]),
        "msr_dense_tc_scores": (I, [P, U32, P, P, ctypes.c_size_t, P]),
        "msr_dense_search_static": (I, [P, U32, P, U32, P, P, P]),
        "msr_quantise_topk": (I, [I, U32, U32, P, I, ctypes.c_size_t, U32, U32, P, I, P, P, P, P]),
        "msr_quantise_candidates": (I, [I, U32, U32, P, I, ctypes.c_size_t, P, P, U32, U32, U32, P, I, P, U32, P, P, P, P]),
        "msr_provenance_hist": (I, [I, U32, P, P, ctypes.c_size_t, P, U32, P, P]),
        "msr_topk_merge_strided": (I, [I, U32, U32, U32, P, P, P, ctypes.c_size_t, ctypes.c_size_t, U32, P, P, P, P]),
        "msr_sparse_set_peer_outputs": (I, [P, U32, P, P, P]),
        "msr_exchange_publish": (I, [I, U32, U32, P, P, P, U32, P, P, P, P]),
        "msr_sparse_set_peer_theta": (I, [P, U32, P, P, U32, U32, U32, P, P]),
        "msr_topk_merge_exchange": (I, [I, U32, U32, U32, P, P, P, ctypes.c_size_t, ctypes.c_size_t, U32, P, P, P, U32, U32, P, P, P,
                                        U32, P]),
        "msr_jsonl_parse": (I, [ctypes.POINTER(ctypes.c_char_p), U32, ctypes.POINTER(P)]),
        "msr_collection_info": (I, [P, P]),
        "msr_collection_export": (I, [P, P, P, P, P, P, P, P, P]),
        "msr_collection_destroy": (I, [P]),
        "msr_invert_docs": (I, [I, U32, U32, P, P, P, P, P, P, P, I, P]),
        "msr_recall_hits": (I, [I, U32, P, ctypes.c_size_t, P, P, P, U32, P, P, P, P]),
        "msr_order_rows": (I, [I, U32, U32, P, I, P, P, ctypes.c_size_t, P, P, I, P, P, P, P, P]),
        "msr_fuse": (I, [I, U32, U32, P, P, P, U32, P, P, P, P, I, I, D, D, P, P, P, P, P]),
        "msr_fuse_f64": (I, [I, U32, U32, P, P, P, U32, P, P, P, P, I, D, D, P, P, P, P, P]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)        # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    return list(sigs)


def load():
    """Load libmsr.so; raises if it has not been built (python -m mllm_sparse_retrieval_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `python -m mllm_sparse_retrieval_b200.build` "
                              "(nvcc, sm_100a).  There is no CPU fallback.")
        lib = ctypes.CDLL(LIB_PATH)
        _declare(lib)
        _lib = lib
    return _lib


def exported_symbols():
    return _declare(ctypes.CDLL(LIB_PATH))


def check(rc):
    if rc != 0:
        raise MsrError(rc, load().msr_last_error().decode("utf-8", "replace"))


def ptr(t):
    """void* of a torch tensor / numpy array / None."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return ctypes.c_void_p(t.data_ptr())
    return t.ctypes.data_as(ctypes.c_void_p)


def current_stream_ptr(device=None):
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
