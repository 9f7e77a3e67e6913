"""ctypes binding of include/docagents_b200.h.

This is the harness-side twin of the cgo binding in INTEGRATION.md: the same C ABI,
the same structs, nothing else.  There is no fallback: if libdocagents_b200.so is not
built, or no B200 is present when a store is created, the caller gets an exception.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdocagents_b200.so")

DA_MAX_DEVICES = 8
DA_MAX_K = 128
DA_MAX_BATCH = 4096
DA_NCCL_ID_BYTES = 128
DA_DEFAULT_MIN_SIMILARITY = 0.7

DA_OK = 0
DA_ERR_INVALID = -1
DA_ERR_DIM = -2
DA_ERR_UNKNOWN_CHUNK = -3
DA_ERR_CAPACITY = -4
DA_ERR_CUDA = -5
DA_ERR_NCCL = -6
DA_ERR_STATE = -7
DA_ERR_NONFINITE = -8

DA_SEARCH_ALL_DOCS = 0x1
DA_SEARCH_READ_ALL_ROWS = 0x2
DA_SEARCH_IVF = 0x4
DA_SEARCH_NO_TENSOR = 0x10
DA_SEARCH_NO_Q8 = 0x20
DA_SEARCH_NO_COMPACT = 0x40

DA_STORE_SPLIT16 = 0x1
DA_STORE_Q8_SIDECAR = 0x2

DA_IVF_SIMT_ASSIGN = 0x1
DA_IVF_NO_SPLIT = 0x2

DA_NOALLOC_MAX_K = 128
DA_NOALLOC_MAX_CTAS = 256


class da_config(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("dim", C.c_int32), ("n_devices", C.c_int32),
                ("device_ids", C.c_int32 * DA_MAX_DEVICES), ("capacity_rows", C.c_uint64),
                ("min_similarity", C.c_float), ("world_size", C.c_int32), ("rank", C.c_int32),
                ("nccl_id", C.c_uint8 * DA_NCCL_ID_BYTES), ("flags", C.c_uint32),
                ("batch_window_us", C.c_uint32), ("batch_max", C.c_uint32), ("min_similarity_f64", C.c_double)]


class da_search_params(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("flags", C.c_uint32),
                ("doc_ids", C.c_void_p), ("n_docs", C.c_size_t),
                ("queries", C.c_void_p), ("batch", C.c_int32), ("dim", C.c_int32),
                ("k", C.c_int32), ("min_similarity", C.c_float), ("ivf_probes", C.c_int32),
                ("doc_offsets", C.c_void_p), ("min_similarity_f64", C.c_double)]


class da_ivf_params(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("lists", C.c_int32), ("iters", C.c_int32),
                ("sample_rows", C.c_uint64), ("seed", C.c_uint64), ("default_probes", C.c_int32),
                ("flags", C.c_uint32)]


class da_synth_params(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("seed_data", C.c_uint64), ("seed_centroid", C.c_uint64),
                ("seed_ids", C.c_uint64), ("chunks_per_doc", C.c_uint32), ("n_clusters", C.c_uint32),
                ("cmul", C.c_int32), ("nmul", C.c_int32), ("row_lo", C.c_uint64), ("n_rows", C.c_uint64),
                ("smul", C.c_int32), ("super_group", C.c_uint32)]


class da_stats(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("dim", C.c_int32), ("n_devices", C.c_int32),
                ("world_size", C.c_int32), ("rank", C.c_int32), ("rows", C.c_uint64),
                ("rows_per_device", C.c_uint64 * DA_MAX_DEVICES), ("hbm_bytes", C.c_uint64),
                ("n_docs", C.c_uint64), ("last_scan_ms", C.c_float), ("last_search_ms", C.c_float),
                ("last_rows_read", C.c_uint64), ("last_bytes_algorithmic", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("ivf_lists", C.c_int32), ("scan_grid", C.c_int32),
                ("scan_ms_sum", C.c_double), ("scan_launches", C.c_uint64),
                ("ivf_build_ms", C.c_float), ("ivf_kmeans_ms", C.c_float), ("ivf_assign_ms", C.c_float),
                ("ivf_assign_tflops", C.c_float), ("peer_gather", C.c_int32), ("tensor_overflows", C.c_uint64),
                ("topk_calls", C.c_uint64), ("topk_passes", C.c_uint64), ("small_searches", C.c_uint64),
                ("last_bench_launches", C.c_uint64)]


# Every symbol include/docagents_b200.h declares, with its signature.
SIGNATURES = {
    "da_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "da_store_create": (C.c_int, [C.POINTER(da_config), C.POINTER(C.c_void_p)]),
    "da_store_destroy": (None, [C.c_void_p]),
    "da_store_save_chunks": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "da_store_save_embeddings": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int32]),
    "da_store_bulk_load": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int32]),
    "da_store_topk": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int32, C.c_int32,
                                C.c_void_p, C.c_void_p, C.POINTER(C.c_int32)]),
    "da_store_export": (C.c_int, [C.c_void_p, C.c_char_p]),
    "da_store_attach": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "da_store_doc_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "da_store_search": (C.c_int, [C.c_void_p, C.POINTER(da_search_params), C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p]),
    "da_store_build_ivf": (C.c_int, [C.c_void_p, C.POINTER(da_ivf_params)]),
    "da_store_generate_synthetic": (C.c_int, [C.c_void_p, C.POINTER(da_synth_params)]),
    "da_synth_query": (C.c_int, [C.POINTER(da_synth_params), C.c_uint64, C.c_uint64, C.c_uint32, C.c_int32, C.c_void_p]),
    "da_synth_ids": (C.c_int, [C.POINTER(da_synth_params), C.c_int32, C.c_uint64, C.c_uint64, C.c_void_p]),
    "da_store_stats": (C.c_int, [C.c_void_p, C.POINTER(da_stats)]),
    "da_store_bench_resident": (C.c_int, [C.c_void_p, C.POINTER(da_search_params), C.c_int32,
                                          C.POINTER(C.c_float)]),
    "da_store_debug_read_rows": (C.c_int, [C.c_void_p, C.c_int32, C.c_uint64, C.c_uint64, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p]),
    "da_store_debug_read_ivf": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "da_store_debug_read_probes": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "da_scan_topk_noalloc": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_uint64, C.c_uint32, C.c_int32, C.c_int32, C.c_float,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "da_last_error": (C.c_char_p, []),
    "da_abi_version": (C.c_int, []),
}

_lib = None


def _preload_newest_nccl() -> None:
    """The library needs libnccl.so.2.  A Python process that later imports torch must end up with
    ONE NCCL, and torch needs its own (newer) bundled build: load that one first when it exists, so
    the loader does not bind the older system copy under the same SONAME.  (A Go host has no such
    concern and simply uses the system libnccl.)"""
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for d in (spec.submodule_search_locations if spec else []):
            cand = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                C.CDLL(cand, mode=C.RTLD_GLOBAL)
                return
    except Exception:
        pass


def load() -> C.CDLL:
    """Load the in-tree shared library; raise if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m doc_agents_b200.build` "
                "(there is no CPU or PyTorch fallback for this path)")
        _preload_newest_nccl()
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error() -> str:
    return (load().da_last_error() or b"").decode("utf-8", "replace")
