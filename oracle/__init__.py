"""ctypes loader for the CPU oracle (oracle/oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(doc_agents_b200) never imports this module.  PARITY UNPINNED: see oracle.c.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

# Fixed seeds of the synthetic corpus.
SEED_DATA = 0xD0C5EED
SEED_CENTROID = 0xC0FFEE
SEED_QUERY = 0xBEEF
SEED_IDS = 0x1D5EED


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in ("oracle.c", "oracle_fast.c", "Makefile")]
    stale = force or not os.path.exists(_LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if stale:
        subprocess.run(["make", "-C", _HERE, "-s", "liboracle.so"], check=True)
    return _LIB_PATH


class _Synth(C.Structure):
    _fields_ = [("seed_data", C.c_uint64), ("seed_centroid", C.c_uint64),
                ("seed_query", C.c_uint64), ("seed_ids", C.c_uint64),
                ("chunks_per_doc", C.c_uint32), ("n_clusters", C.c_uint32),
                ("cmul", C.c_int32), ("nmul", C.c_int32), ("smul", C.c_int32), ("super_group", C.c_uint32)]


@dataclass(frozen=True)
class Synth:
    """Parameters of the synthetic corpus; defaults are the BASELINE configs."""
    seed_data: int = SEED_DATA
    seed_centroid: int = SEED_CENTROID
    seed_query: int = SEED_QUERY
    seed_ids: int = SEED_IDS
    chunks_per_doc: int = 200
    n_clusters: int = 1024
    cmul: int = 2   # centroid multiplier  } sigma = nmul / cmul = 0.5
    nmul: int = 1   # noise multiplier     } same-cluster cosine ~ 0.8
    smul: int = 0          # optional second level: + smul x the direction of the cluster's group of
    super_group: int = 0   # `super_group` clusters (0 = off): clusters of one group overlap

    def c(self) -> _Synth:
        return _Synth(self.seed_data, self.seed_centroid, self.seed_query, self.seed_ids,
                      self.chunks_per_doc, self.n_clusters, self.cmul, self.nmul, self.smul, self.super_group)


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        fp = C.POINTER(C.c_float)
        L.orc_gen_rows.argtypes = [C.POINTER(_Synth), C.c_uint64, C.c_uint64, C.c_int, fp]
        L.orc_gen_rows_at.argtypes = [C.POINTER(_Synth), C.c_void_p, C.c_uint64, C.c_int, fp]
        L.orc_gen_query.argtypes = [C.POINTER(_Synth), C.c_uint64, C.c_uint32, C.c_int, fp]
        L.orc_chunk_id.argtypes = [C.POINTER(_Synth), C.c_uint64, C.c_void_p]
        L.orc_doc_id.argtypes = [C.POINTER(_Synth), C.c_uint64, C.c_void_p]
        L.orc_normalize.argtypes = [fp, C.c_int]
        L.orc_avg_similarity.argtypes = [fp, C.c_int]
        L.orc_avg_similarity.restype = C.c_float
        for n in ("orc_cosine_distance_f32", "orc_cosine_distance_f64"):
            getattr(L, n).argtypes = [fp, fp, C.c_int]
            getattr(L, n).restype = C.c_double
        for n in ("orc_topk", "orc_topk_fast"):
            getattr(L, n).argtypes = [fp, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p, fp, C.c_int,
                                      C.c_double, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
            getattr(L, n).restype = C.c_int
        L.orc_all_scores.argtypes = [fp, C.c_uint64, C.c_int, fp, C.c_int, fp]
        L.orc_table_new.argtypes = [C.c_int]
        L.orc_table_new.restype = C.c_void_p
        L.orc_table_free.argtypes = [C.c_void_p]
        L.orc_table_rows.argtypes = [C.c_void_p]
        L.orc_table_rows.restype = C.c_uint64
        L.orc_table_vecs.argtypes = [C.c_void_p]
        L.orc_table_vecs.restype = C.c_void_p
        L.orc_table_ids.argtypes = [C.c_void_p]
        L.orc_table_ids.restype = C.c_void_p
        L.orc_table_upsert.argtypes = [C.c_void_p, C.c_void_p, fp, C.c_uint64, C.c_int]
        L.orc_table_upsert.restype = C.c_int
        L.orc_num_threads.restype = C.c_int
        L.orc_set_num_threads.argtypes = [C.c_int]
        _lib = L
    return _lib


def _fp(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags.c_contiguous
    return a.ctypes.data_as(C.POINTER(C.c_float))


def gen_rows(p: Synth, row_lo: int, n: int, dim: int) -> np.ndarray:
    out = np.empty((n, dim), dtype=np.float32)
    if n:
        s = p.c()
        lib().orc_gen_rows(C.byref(s), row_lo, n, dim, _fp(out))
    return out


def gen_rows_at(p: Synth, rows, dim: int) -> np.ndarray:
    """Rows of the synthetic corpus by global row number, in the order given."""
    rows = np.ascontiguousarray(rows, dtype=np.uint64)
    out = np.empty((rows.shape[0], dim), dtype=np.float32)
    if rows.shape[0]:
        s = p.c()
        lib().orc_gen_rows_at(C.byref(s), rows.ctypes.data, rows.shape[0], dim, _fp(out))
    return out


def gen_query(p: Synth, qi: int, cluster: int, dim: int) -> np.ndarray:
    out = np.empty(dim, dtype=np.float32)
    s = p.c()
    lib().orc_gen_query(C.byref(s), qi, cluster, dim, _fp(out))
    return out


def chunk_ids(p: Synth, row_lo: int, n: int) -> np.ndarray:
    out = np.empty((n, 16), dtype=np.uint8)
    s = p.c()
    for i in range(n):
        lib().orc_chunk_id(C.byref(s), row_lo + i, out[i].ctypes.data)
    return out


def doc_ids(p: Synth, doc_lo: int, n: int) -> np.ndarray:
    out = np.empty((n, 16), dtype=np.uint8)
    s = p.c()
    for i in range(n):
        lib().orc_doc_id(C.byref(s), doc_lo + i, out[i].ctypes.data)
    return out


def row_docs(p: Synth, row_lo: int, n: int) -> np.ndarray:
    return ((np.arange(row_lo, row_lo + n, dtype=np.uint64)) // p.chunks_per_doc).astype(np.uint32)


def normalize(v: np.ndarray) -> np.ndarray:
    out = np.ascontiguousarray(v, dtype=np.float32).copy()
    lib().orc_normalize(_fp(out), out.shape[0])
    return out


def cosine_distance(a: np.ndarray, b: np.ndarray, f64: bool = False) -> float:
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    fn = lib().orc_cosine_distance_f64 if f64 else lib().orc_cosine_distance_f32
    return fn(_fp(a), _fp(b), a.shape[0])


def avg_similarity(scores) -> float:
    s = np.ascontiguousarray(scores, dtype=np.float32)
    return float(lib().orc_avg_similarity(_fp(s), len(s)))


def topk(mat: np.ndarray, row_doc, doc_allowed, q: np.ndarray, k: int, min_sim: float = 0.7,
         f64: bool = False, threads: int = 0, fast: bool = False):
    """Store.TopK restated.  Returns (rows u32[n], scores f32[n], dist f64[n]) best first.

    doc_allowed: None = every document, else uint8 array indexed by document index.
    """
    mat = np.ascontiguousarray(mat, dtype=np.float32)
    q = np.ascontiguousarray(q, dtype=np.float32)
    n, dim = mat.shape
    assert q.shape[0] == dim
    kk = max(k, 1)
    rows = np.zeros(kk, dtype=np.uint32)
    scores = np.zeros(kk, dtype=np.float32)
    dist = np.zeros(kk, dtype=np.float64)
    rd = None if row_doc is None else np.ascontiguousarray(row_doc, dtype=np.uint32)
    da = None if doc_allowed is None else np.ascontiguousarray(doc_allowed, dtype=np.uint8)
    if da is not None:
        assert rd is not None
    fn = lib().orc_topk_fast if fast else lib().orc_topk
    cnt = fn(_fp(mat), n, dim, None if rd is None else rd.ctypes.data,
             None if da is None else da.ctypes.data, _fp(q), k, float(min_sim), int(f64), threads,
             rows.ctypes.data, scores.ctypes.data, dist.ctypes.data)
    return rows[:cnt].copy(), scores[:cnt].copy(), dist[:cnt].copy()


def all_scores(mat: np.ndarray, q: np.ndarray, f64: bool = False) -> np.ndarray:
    mat = np.ascontiguousarray(mat, dtype=np.float32)
    q = np.ascontiguousarray(q, dtype=np.float32)
    out = np.empty(mat.shape[0], dtype=np.float32)
    lib().orc_all_scores(_fp(mat), mat.shape[0], mat.shape[1], _fp(q), int(f64), _fp(out))
    return out


class Table:
    """SaveEmbeddings restated: rows keyed by chunk_id, overwrite on conflict."""

    def __init__(self, dim: int):
        self.dim = dim
        self._t = lib().orc_table_new(dim)

    def __del__(self):
        if getattr(self, "_t", None):
            lib().orc_table_free(self._t)
            self._t = None

    def upsert(self, ids: np.ndarray, vecs: np.ndarray) -> int:
        ids = np.ascontiguousarray(ids, dtype=np.uint8).reshape(-1, 16)
        vecs = np.ascontiguousarray(vecs, dtype=np.float32)
        n = ids.shape[0]
        dim = vecs.shape[1] if vecs.ndim == 2 else self.dim
        return lib().orc_table_upsert(self._t, ids.ctypes.data, _fp(vecs), n, dim)

    @property
    def rows(self) -> int:
        return lib().orc_table_rows(self._t)

    def vecs(self) -> np.ndarray:
        n = self.rows
        buf = (C.c_float * (n * self.dim)).from_address(lib().orc_table_vecs(self._t))
        return np.frombuffer(buf, dtype=np.float32).reshape(n, self.dim).copy()

    def ids(self) -> np.ndarray:
        n = self.rows
        buf = (C.c_uint8 * (n * 16)).from_address(lib().orc_table_ids(self._t))
        return np.frombuffer(buf, dtype=np.uint8).reshape(n, 16).copy()


def num_threads() -> int:
    return lib().orc_num_threads()


def set_num_threads(n: int) -> None:
    lib().orc_set_num_threads(int(n))
