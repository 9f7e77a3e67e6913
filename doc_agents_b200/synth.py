"""The synthetic corpus of BASELINE.json's configs, as the product library states it (device generator
`da_store_generate_synthetic`, host queries/ids `da_synth_query` / `da_synth_ids`).  Used by bench.py and
tools/ so that they need nothing from oracle/; tests check that it equals the oracle's independent statement
of the same spec bit for bit."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib as L


@dataclass(frozen=True)
class Synth:
    seed_data: int = 0xD0C5EED
    seed_centroid: int = 0xC0FFEE
    seed_query: int = 0xBEEF
    seed_ids: int = 0x1D5EED
    chunks_per_doc: int = 200
    n_clusters: int = 1024
    cmul: int = 2
    nmul: int = 1
    smul: int = 0
    super_group: int = 0

    def c(self) -> L.da_synth_params:
        p = L.da_synth_params()
        p.struct_size = C.sizeof(L.da_synth_params)
        p.seed_data, p.seed_centroid, p.seed_ids = self.seed_data, self.seed_centroid, self.seed_ids
        p.chunks_per_doc, p.n_clusters, p.cmul, p.nmul = self.chunks_per_doc, self.n_clusters, self.cmul, self.nmul
        p.smul, p.super_group = self.smul, self.super_group
        return p


def _check(rc: int) -> None:
    if rc != L.DA_OK:
        raise RuntimeError(L.last_error())


def gen_query(s: Synth, qi: int, cluster: int, dim: int) -> np.ndarray:
    out = np.empty(dim, dtype=np.float32)
    p = s.c()
    _check(L.load().da_synth_query(C.byref(p), s.seed_query, qi, cluster, dim, out.ctypes.data))
    return out


def _ids(s: Synth, kind: int, first: int, n: int) -> np.ndarray:
    out = np.empty((n, 16), dtype=np.uint8)
    p = s.c()
    _check(L.load().da_synth_ids(C.byref(p), kind, first, n, out.ctypes.data if n else None))
    return out


def chunk_ids(s: Synth, first_row: int, n: int) -> np.ndarray:
    return _ids(s, 0, first_row, n)


def doc_ids(s: Synth, first_doc: int, n: int) -> np.ndarray:
    return _ids(s, 1, first_doc, n)
