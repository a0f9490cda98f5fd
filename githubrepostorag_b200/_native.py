"""ctypes binding of ``include/coderag_b200.h``.

There is no fallback: if the CUDA library has not been built (``python -m
githubrepostorag_b200.build``) loading raises, and every entry point raises ``RuntimeError``
with the library's message on a non-zero return (the reference propagates Python exceptions,
``rag_worker/src/worker/services/agent_graph.py:276-278``).  ctypes drops the GIL for the
duration of each call, which is what lets the worker's thread pool
(``rag_worker/src/worker/worker.py:136-139``) overlap searches.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Optional

_LIB_PATH = Path(__file__).resolve().parent / "_lib" / "libcoderag_b200.so"
_lib: Optional[C.CDLL] = None

ABI_VERSION = 2
METRIC_COSINE = 0
METRIC_IP = 1
FILT_ANY = -1
FILT_IMPOSSIBLE = -2
CODE_MISSING = -1
MAX_KEYS = 16
MAX_TOPK = 128
MAX_DIM = 4096
NOALLOC_MAX_BATCH = 4
MAX_PEERS = 8


class Exchange(C.Structure):
    """``crb_exchange`` of include/coderag_b200.h."""
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("cap_keys", C.c_uint64),
                ("peer_keys", C.c_void_p * MAX_PEERS), ("peer_counters", C.c_void_p * MAX_PEERS),
                ("dest_mask", C.c_uint32)]

class Ctl(C.Structure):
    """``crb_ctl`` of include/coderag_b200.h."""
    _fields_ = [("seq_off", C.c_uint64), ("hdr_off", C.c_uint64), ("ack_off", C.c_uint64), ("status_off", C.c_uint64),
                ("queries_off", C.c_uint64), ("filt_off", C.c_uint64), ("rank", C.c_int32), ("world", C.c_int32)]


# every symbol include/coderag_b200.h declares: name -> (restype, argtypes)
_vp, _i32, _i64, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t
PROTOTYPES = {
    "crb_abi_version": (C.c_int, []),
    "crb_last_error": (C.c_char_p, []),
    "crb_index_create": (C.c_int, [C.POINTER(_vp), _vp, _vp, _vp, _i64, _i32, _i32, _i64, _i32]),
    "crb_index_destroy": (C.c_int, [_vp]),
    "crb_index_reload_tuning": (C.c_int, [_vp]),
    "crb_index_set_k5_stats": (C.c_int, [_vp, _vp]),
    "crb_index_rebind": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64]),
    "crb_index_rebind_columns": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64]),
    "crb_tombstone_rows": (C.c_int, [_vp, _vp, _i32]),
    "crb_encode_rows": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _i32, _vp]),
    "crb_row_inv_norms": (C.c_int, [_vp, _vp, _i64, _i32, _i32, _vp]),
    "crb_workspace_bytes": (_sz, [_vp, _i32, _i32, _i32]),
    "crb_search": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _i32, _vp]),
    "crb_search_masked": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _i32, _vp]),
    "crb_search_host": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _i32]),
    "crb_search_host_vectors": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _i32]),
    "crb_search_host_any_of": (C.c_int, [_vp, _vp, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    "crb_merge_keys": (C.c_int, [_vp, _i32, _i32, _i32, _i32, _vp, _i32, _vp, _vp, _vp, _vp, _i32, _vp]),
    "crb_search_exchange": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _i32, _vp,
                                       _vp, C.c_uint32, C.c_uint32]),
    "crb_search_exchange_host": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _i32, _vp,
                                            C.c_uint32, C.c_uint32]),
    "crb_search_exchange_host_any_of": (C.c_int, [_vp, _vp, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _vp, C.c_uint32, C.c_uint32]),
    "crb_serve_post_search": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp,
                                         C.c_double]),
    "crb_serve_next": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "crb_host_register": (C.c_int, [_vp, _sz]),
    "crb_host_unregister": (C.c_int, [_vp]),
    "crb_fetch_rows_host": (C.c_int, [_vp, _vp, _i32, _vp]),
    "crb_search_noalloc": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp]),
    "crb_launch_count": (_i64, []),
}


def lib_path() -> Path:
    return _LIB_PATH


def load() -> C.CDLL:
    """Load the library once; raise loudly when it is missing or of another ABI."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise RuntimeError(
            f"{_LIB_PATH} is missing: build it with `python -m githubrepostorag_b200.build` "
            "(nvcc, sm_100a). There is no CPU fallback for the retrieval path.")
    lib = C.CDLL(str(_LIB_PATH))
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError here == header and library disagree
        fn.restype = res
        fn.argtypes = args
    if lib.crb_abi_version() != ABI_VERSION:
        raise RuntimeError(f"ABI mismatch: library {lib.crb_abi_version()} != binding {ABI_VERSION}")
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().crb_last_error()
        raise RuntimeError("coderag_b200: " + (msg.decode("utf-8", "replace") if msg else f"error {rc}"))


def launch_count() -> int:
    return int(load().crb_launch_count())
