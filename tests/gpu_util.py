"""Helpers shared by the GPU parity tests (all calls go through the C ABI via ctypes)."""
from __future__ import annotations

import ctypes as C

import numpy as np

import oracle
from doc_agents_b200 import GpuStore

S0 = oracle.Synth()


def device_count() -> int:
    try:
        cudart = C.CDLL("libcudart.so.12")
        n = C.c_int(0)
        if cudart.cudaGetDeviceCount(C.byref(n)) != 0:
            return 0
        return n.value
    except OSError:
        return 0


def ids_for(n_rows: int, n_docs: int):
    """Deterministic chunk ids (by row) and document ids (by document index, +200 spare unknown ones)."""
    return oracle.chunk_ids(S0, 0, n_rows), oracle.doc_ids(S0, 0, n_docs + 200)


def load_case(st: GpuStore, case):
    """SaveChunks per document, then ONE SaveEmbeddings in case row order, so that the store's
    insertion row i is the case's row i."""
    n = case.mat.shape[0]
    chunk, docs = ids_for(n, case.n_docs)
    rd = np.asarray(case.row_doc)
    for d in range(case.n_docs):
        st.save_chunks_raw(docs[d], chunk[rd == d])
    st.save_embeddings_raw(chunk, case.mat)
    return chunk, docs


def doc_filter(case, docs):
    if case.docs is None:
        return docs[:case.n_docs]
    return docs[np.asarray(case.docs, dtype=np.int64)] if len(case.docs) else docs[:0]
