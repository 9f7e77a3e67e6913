"""Host-side mirror of the reference's `store.Store` interface for the one path this
repository replaces (internal/store/store.go:57-67): SaveChunks, SaveEmbeddings, TopK.

It is the Python twin of the Go decorator sketched in INTEGRATION.md: type conversion and
one C call per method.  The relational side (chunk text, summaries) stays in Postgres in the
reference; here a small in-memory table stands in for it so TopK can hand back the same
`SearchResult` shape (store.go:50-54).  All search arithmetic happens in the CUDA library.
"""
from __future__ import annotations

import ctypes as C
import math
import uuid
from dataclasses import dataclass, field
from typing import Iterable, Sequence

import numpy as np

from . import _lib as L


class StoreError(RuntimeError):
    """A failed C-ABI call: what the Go shim returns as `error` (HTTP 500 upstream)."""

    def __init__(self, code: int, message: str):
        super().__init__(f"docagents_b200 error {code}: {message}")
        self.code = code
        self.message = message


def _check(rc: int) -> None:
    if rc != L.DA_OK:
        raise StoreError(rc, L.last_error())


# --- domain types, field for field (internal/store/store.go:30-54) -------------------------
@dataclass
class Chunk:
    ID: uuid.UUID
    DocumentID: uuid.UUID = uuid.UUID(int=0)
    Index: int = 0
    Text: str = ""
    TokenCount: int = 0


@dataclass
class Summary:
    DocumentID: uuid.UUID = uuid.UUID(int=0)
    Summary: str = ""
    KeyPoints: list = field(default_factory=list)


@dataclass
class Embedding:
    ChunkID: uuid.UUID
    Vector: np.ndarray
    Model: str = ""


@dataclass
class SearchResult:
    Chunk: Chunk
    Score: np.float32
    Summary: Summary


def _ids_to_bytes(ids: Iterable[uuid.UUID]) -> np.ndarray:
    ids = list(ids)
    out = np.empty((len(ids), 16), dtype=np.uint8)
    for i, u in enumerate(ids):
        out[i] = np.frombuffer(u.bytes, dtype=np.uint8)
    return out


def _ptr(a: np.ndarray):
    return a.ctypes.data if a.size else None


class GpuStore:
    """`STORE_PROVIDER=gpu`: device-resident embeddings, exact (or IVFFlat) cosine top-k."""

    def __init__(self, dim: int = 3072, capacity_rows: int = 1 << 16, devices: Sequence[int] = (0,),
                 min_similarity: float = L.DA_DEFAULT_MIN_SIMILARITY, world_size: int = 1, rank: int = 0,
                 nccl_id: bytes | None = None, batch_window_us: int = 0, batch_max: int = 0,
                 split16: bool = False, q8_sidecar: bool = False):
        lib = L.load()
        cfg = L.da_config()
        cfg.struct_size = C.sizeof(L.da_config)
        cfg.dim = dim
        cfg.n_devices = len(devices)
        for i, d in enumerate(devices):
            cfg.device_ids[i] = d
        cfg.capacity_rows = capacity_rows
        cfg.min_similarity = min_similarity
        # the floor as the float8 it is in the reference's SQL (0.7 is not a float32): compared as (double)score >= floor
        cfg.min_similarity_f64 = float(min_similarity) if math.isfinite(min_similarity) else 0.0
        cfg.world_size = world_size
        cfg.rank = rank
        cfg.batch_window_us = batch_window_us
        cfg.flags = (L.DA_STORE_SPLIT16 if split16 else 0) | (L.DA_STORE_Q8_SIDECAR if q8_sidecar else 0)
        cfg.batch_max = batch_max
        if nccl_id is not None:
            assert len(nccl_id) == L.DA_NCCL_ID_BYTES
            C.memmove(cfg.nccl_id, nccl_id, L.DA_NCCL_ID_BYTES)
        self._h = C.c_void_p()
        self.dim = dim
        self.min_similarity = min_similarity
        self._lib = lib
        _check(lib.da_store_create(C.byref(cfg), C.byref(self._h)))
        # stand-in for the relational tables the reference joins against (postgres.go:226-239)
        self._chunks: dict[uuid.UUID, Chunk] = {}
        self._summaries: dict[uuid.UUID, Summary] = {}

    # -- one store, two processes (writer = analysis agent, reader = query agent) -------
    def export(self, shm_name: str) -> None:
        """Publish this store under a POSIX shared-memory name so that another process can attach to it."""
        _check(self._lib.da_store_export(self._h, shm_name.encode()))

    @classmethod
    def attach(cls, shm_name: str) -> "GpuStore":
        """Read-only view, in THIS process, of the store another process exported: same HBM, kept current."""
        self = cls.__new__(cls)
        self._lib = L.load()
        self._h = C.c_void_p()
        _check(self._lib.da_store_attach(shm_name.encode(), C.byref(self._h)))
        s = self.stats()
        self.dim = s.dim
        self.min_similarity = L.DA_DEFAULT_MIN_SIMILARITY
        self._chunks, self._summaries = {}, {}
        return self

    # -- lifecycle ---------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None) and self._h.value:
            self._lib.da_store_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @staticmethod
    def nccl_unique_id() -> bytes:
        buf = (C.c_uint8 * L.DA_NCCL_ID_BYTES)()
        _check(L.load().da_nccl_unique_id(buf))
        return bytes(buf)

    # -- Store.SaveChunks (store.go:61; postgres.go:142-164) -------------------------------
    def SaveChunks(self, docID: uuid.UUID, chunks: Sequence[Chunk]) -> list[Chunk]:
        out = []
        for i, c in enumerate(chunks):
            cid = c.ID if c.ID.int != 0 else uuid.uuid4()      # ids are assigned here (postgres.go:150-157)
            out.append(Chunk(cid, docID, c.Index if c.Index else i, c.Text, c.TokenCount))
        ids = _ids_to_bytes(c.ID for c in out)
        doc = np.frombuffer(docID.bytes, dtype=np.uint8).copy()
        _check(self._lib.da_store_save_chunks(self._h, doc.ctypes.data, _ptr(ids), len(out)))
        for c in out:
            self._chunks[c.ID] = c
        return out

    def SaveSummary(self, docID: uuid.UUID, summary: Summary) -> None:
        self._summaries[docID] = Summary(docID, summary.Summary, list(summary.KeyPoints))

    # -- Store.SaveEmbeddings (store.go:64; postgres.go:176-201) ---------------------------
    def SaveEmbeddings(self, embs: Sequence[Embedding]) -> None:
        if len(embs) == 0:
            return
        dims = {int(np.asarray(e.Vector).shape[0]) for e in embs}
        if len(dims) != 1:
            raise StoreError(L.DA_ERR_DIM, f"expected {self.dim} dimensions, got mixed lengths {sorted(dims)}")
        dim = dims.pop()
        vecs = np.ascontiguousarray(np.stack([np.asarray(e.Vector, dtype=np.float32) for e in embs]))
        ids = _ids_to_bytes(e.ChunkID for e in embs)
        _check(self._lib.da_store_save_embeddings(self._h, _ptr(ids), _ptr(vecs), len(embs), dim))

    def save_embeddings_raw(self, chunk_ids: np.ndarray, vecs: np.ndarray) -> None:
        """Bulk form of SaveEmbeddings on raw arrays (n x 16 uint8, n x dim float32)."""
        chunk_ids = np.ascontiguousarray(chunk_ids, dtype=np.uint8).reshape(-1, 16)
        vecs = np.ascontiguousarray(vecs, dtype=np.float32)
        n = chunk_ids.shape[0]
        dim = vecs.shape[1] if vecs.ndim == 2 else self.dim
        _check(self._lib.da_store_save_embeddings(self._h, _ptr(chunk_ids), _ptr(vecs), n, dim))

    def bulk_load(self, chunk_ids: np.ndarray, doc_ids: np.ndarray, vecs: np.ndarray) -> None:
        """Cold start: rows of `embeddings JOIN chunks` (chunk id, document id, vector) in one call."""
        chunk_ids = np.ascontiguousarray(chunk_ids, dtype=np.uint8).reshape(-1, 16)
        doc_ids = np.ascontiguousarray(doc_ids, dtype=np.uint8).reshape(-1, 16)
        vecs = np.ascontiguousarray(vecs, dtype=np.float32)
        assert chunk_ids.shape[0] == doc_ids.shape[0] == vecs.shape[0]
        _check(self._lib.da_store_bulk_load(self._h, _ptr(chunk_ids), _ptr(doc_ids), _ptr(vecs), chunk_ids.shape[0],
                                            vecs.shape[1] if vecs.ndim == 2 else self.dim))

    def save_chunks_raw(self, doc_id: np.ndarray, chunk_ids: np.ndarray) -> None:
        doc_id = np.ascontiguousarray(doc_id, dtype=np.uint8).reshape(16)
        chunk_ids = np.ascontiguousarray(chunk_ids, dtype=np.uint8).reshape(-1, 16)
        _check(self._lib.da_store_save_chunks(self._h, doc_id.ctypes.data, _ptr(chunk_ids), chunk_ids.shape[0]))

    # -- Store.TopK (store.go:66; postgres.go:218-285) -------------------------------------
    def TopK(self, docIDs: Sequence[uuid.UUID], vector, k: int):
        """Returns list[SearchResult] best first, or None when nothing qualifies
        (the reference returns a nil slice: postgres.go:250,284)."""
        q = np.ascontiguousarray(vector, dtype=np.float32)
        ids = _ids_to_bytes(docIDs)
        kk = max(int(k), 1)
        out_ids = np.zeros((kk, 16), dtype=np.uint8)
        out_scores = np.zeros(kk, dtype=np.float32)
        n = C.c_int32(0)
        _check(self._lib.da_store_topk(self._h, _ptr(ids), ids.shape[0], _ptr(q), q.shape[0], int(k),
                                       out_ids.ctypes.data, out_scores.ctypes.data, C.byref(n)))
        if n.value == 0:
            return None
        results = []
        for i in range(n.value):
            cid = uuid.UUID(bytes=out_ids[i].tobytes())
            chunk = self._chunks.get(cid, Chunk(cid))
            summ = self._summaries.get(chunk.DocumentID, Summary(chunk.DocumentID, "", []))
            results.append(SearchResult(chunk, np.float32(out_scores[i]), summ))
        return results

    def doc_rows(self, doc_ids) -> np.ndarray:
        """Resident embeddings per document (0 = never given to this store): what a reader process checks before it
        searches documents that another process embedded."""
        d = np.ascontiguousarray(doc_ids, dtype=np.uint8).reshape(-1, 16) if not isinstance(doc_ids, (list, tuple)) \
            else _ids_to_bytes(doc_ids)
        out = np.zeros(d.shape[0], dtype=np.uint32)
        _check(self._lib.da_store_doc_rows(self._h, _ptr(d), d.shape[0], out.ctypes.data if d.shape[0] else None))
        return out

    # -- everything-exposed search on raw arrays -----------------------------------------
    def search(self, queries, k: int, doc_ids: np.ndarray | None = None, *, all_docs: bool = False,
               min_similarity: float | None = None, read_all_rows: bool = False, ivf: bool = False,
               probes: int = 0, no_tensor: bool = False, no_q8: bool = False,
               no_compact: bool = False, float32_floor: bool = False):
        """queries: dim or batch x dim.  doc_ids: one n x 16 array for the whole batch, or a list with one
        array per query.  Returns (chunk_ids[b][n,16], scores[b][n], rows[b][n])."""
        out_ids, out_scores, out_rows, out_n = self.search_arrays(
            queries, k, doc_ids, all_docs=all_docs, min_similarity=min_similarity, read_all_rows=read_all_rows,
            ivf=ivf, probes=probes, no_tensor=no_tensor, no_q8=no_q8, no_compact=no_compact, float32_floor=float32_floor)
        b = out_n.shape[0]
        return ([out_ids[i, :out_n[i]] for i in range(b)],
                [out_scores[i, :out_n[i]] for i in range(b)],
                [out_rows[i, :out_n[i]] for i in range(b)])

    def search_arrays(self, queries, k: int, doc_ids=None, *, all_docs: bool = False,
                      min_similarity: float | None = None, read_all_rows: bool = False, ivf: bool = False,
                      probes: int = 0, no_tensor: bool = False, no_q8: bool = False,
                      no_compact: bool = False, float32_floor: bool = False):
        """The C-ABI call as it is: (chunk_ids[b,k,16], scores[b,k], rows[b,k], n[b]); query i has n[i] hits,
        the rest of its k slots are zero.  No per-query Python objects (a batch of 4096 would spend longer
        building lists than searching)."""
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim == 1:
            q = q[None, :]
        p, keep = self._params(q, k, doc_ids, all_docs, min_similarity, read_all_rows, ivf, probes, no_tensor, no_q8, no_compact,
                               float32_floor)
        b, kk = q.shape[0], max(int(k), 1)
        out_ids = np.zeros((b, kk, 16), dtype=np.uint8)
        out_scores = np.zeros((b, kk), dtype=np.float32)
        out_rows = np.zeros((b, kk), dtype=np.uint32)
        out_n = np.zeros(b, dtype=np.int32)
        _check(self._lib.da_store_search(self._h, C.byref(p), out_ids.ctypes.data, out_scores.ctypes.data,
                                         out_rows.ctypes.data, out_n.ctypes.data))
        del keep
        return out_ids, out_scores, out_rows, out_n

    def _params(self, q, k, doc_ids, all_docs, min_similarity, read_all_rows, ivf, probes, no_tensor=False, no_q8=False,
                no_compact=False, float32_floor=False):
        p = L.da_search_params()
        p.struct_size = C.sizeof(L.da_search_params)
        p.flags = (L.DA_SEARCH_ALL_DOCS if all_docs else 0) | (L.DA_SEARCH_READ_ALL_ROWS if read_all_rows else 0) \
            | (L.DA_SEARCH_IVF if ivf else 0) | (L.DA_SEARCH_NO_TENSOR if no_tensor else 0) \
            | (L.DA_SEARCH_NO_Q8 if no_q8 else 0) | (L.DA_SEARCH_NO_COMPACT if no_compact else 0)
        d = offs = None
        if isinstance(doc_ids, (list, tuple)):                      # one document set per query
            assert len(doc_ids) == q.shape[0]
            parts = [np.ascontiguousarray(x, dtype=np.uint8).reshape(-1, 16) for x in doc_ids]
            offs = np.zeros(len(parts) + 1, dtype=np.uint64)
            offs[1:] = np.cumsum([x.shape[0] for x in parts])
            d = np.concatenate(parts) if parts else np.zeros((0, 16), np.uint8)
            d = np.ascontiguousarray(np.concatenate([d, np.zeros((1, 16), np.uint8)]))   # never a NULL pointer
            p.doc_ids = d.ctypes.data
            p.n_docs = int(offs[-1])
            p.doc_offsets = offs.ctypes.data
        elif doc_ids is not None:
            d = np.ascontiguousarray(doc_ids, dtype=np.uint8).reshape(-1, 16)
            p.doc_ids = _ptr(d)
            p.n_docs = d.shape[0]
        p.queries = q.ctypes.data
        p.batch = q.shape[0]
        p.dim = q.shape[1]
        p.k = int(k)
        ms = self.min_similarity if min_similarity is None else min_similarity
        p.min_similarity = -math.inf if ms is None else ms
        p.min_similarity_f64 = float(ms) if (ms is not None and math.isfinite(ms) and not float32_floor) else 0.0
        p.ivf_probes = probes
        return p, (q, d, offs)

    def bench_resident(self, queries, k: int, iters: int, doc_ids=None, *, all_docs=False,
                       min_similarity=None, read_all_rows=False, ivf=False, probes=0, no_tensor=False,
                       no_q8=False, no_compact=False) -> float:
        """Total CUDA-event milliseconds of `iters` searches with inputs resident in HBM."""
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim == 1:
            q = q[None, :]
        p, keep = self._params(q, k, doc_ids, all_docs, min_similarity, read_all_rows, ivf, probes, no_tensor, no_q8, no_compact)
        ms = C.c_float(0)
        _check(self._lib.da_store_bench_resident(self._h, C.byref(p), iters, C.byref(ms)))
        del keep
        return float(ms.value)

    def generate_synthetic(self, synth, row_lo: int, n_rows: int) -> None:
        """Device-side fill with the synthetic corpus (`synth` has the oracle.Synth fields)."""
        p = L.da_synth_params()
        p.struct_size = C.sizeof(L.da_synth_params)
        p.seed_data, p.seed_centroid, p.seed_ids = synth.seed_data, synth.seed_centroid, synth.seed_ids
        p.chunks_per_doc, p.n_clusters = synth.chunks_per_doc, synth.n_clusters
        p.cmul, p.nmul = synth.cmul, synth.nmul
        p.smul, p.super_group = getattr(synth, "smul", 0), getattr(synth, "super_group", 0)
        p.row_lo, p.n_rows = row_lo, n_rows
        _check(self._lib.da_store_generate_synthetic(self._h, C.byref(p)))

    def build_ivf(self, lists: int = 100, iters: int = 10, sample_rows: int = 0, seed: int = 1,
                  default_probes: int = 1, simt_assign: bool = False, split: bool = True) -> None:
        p = L.da_ivf_params()
        p.struct_size = C.sizeof(L.da_ivf_params)
        p.lists, p.iters, p.sample_rows, p.seed, p.default_probes = lists, iters, sample_rows, seed, default_probes
        p.flags = (L.DA_IVF_SIMT_ASSIGN if simt_assign else 0) | (0 if split else L.DA_IVF_NO_SPLIT)
        _check(self._lib.da_store_build_ivf(self._h, C.byref(p)))

    def stats(self) -> L.da_stats:
        s = L.da_stats()
        _check(self._lib.da_store_stats(self._h, C.byref(s)))
        return s

    def debug_rows(self, device_index: int, row_lo: int, n: int):
        vecs = np.empty((n, self.dim), dtype=np.float32)
        inv = np.empty(n, dtype=np.float32)
        slot = np.empty(n, dtype=np.uint32)
        ids = np.empty((n, 16), dtype=np.uint8)
        _check(self._lib.da_store_debug_read_rows(self._h, device_index, row_lo, n, vecs.ctypes.data,
                                                  inv.ctypes.data, slot.ctypes.data, ids.ctypes.data))
        return vecs, inv, slot, ids


    def debug_ivf(self, device_index: int = 0):
        """(list offsets, row numbers grouped by list, centroids) of the IVFFlat index on one shard."""
        st = self.stats()
        L, n = st.ivf_lists, int(st.rows_per_device[device_index])
        off = np.empty(L + 1, dtype=np.uint32)
        rows = np.empty(n, dtype=np.uint32)
        cent = np.empty((L, self.dim), dtype=np.float32)
        _check(self._lib.da_store_debug_read_ivf(self._h, device_index, off.ctypes.data, rows.ctypes.data,
                                                 cent.ctypes.data))
        return off, rows, cent


    def debug_probes(self, batch: int, probes: int, device_index: int = 0) -> np.ndarray:
        """The lists each query of the last IVFFlat search probed on one shard (batch x probes, -1 = none)."""
        out = np.empty((batch, probes), dtype=np.uint32)
        _check(self._lib.da_store_debug_read_probes(self._h, device_index, batch, probes, out.ctypes.data))
        return out.astype(np.int64) - (out == 0xFFFFFFFF) * (1 << 32)


def calculate_avg_similarity(results) -> np.float32:
    """cmd/query/main.go:161-170: float32 running sum over the results, divided by n."""
    if not results:
        return np.float32(0.0)
    total = np.float32(0.0)
    for r in results:
        total = np.float32(total + np.float32(r.Score))
    return np.float32(total / np.float32(len(results)))
