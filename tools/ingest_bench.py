#!/usr/bin/env python
"""Write side of the path (Store.SaveEmbeddings, one call per document as the analysis agent makes them):
rows/s and GB/s of SaveChunks + SaveEmbeddings through the C ABI with host buffers, then a re-save of the
same chunks (the ON CONFLICT DO UPDATE path).     python tools/ingest_bench.py [--docs 2000] [--dim 3072]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from doc_agents_b200 import GpuStore, synth


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--docs", type=int, default=2000)
    ap.add_argument("--chunks", type=int, default=200)
    ap.add_argument("--dim", type=int, default=3072)
    a = ap.parse_args()
    S = synth.Synth()
    n = a.docs * a.chunks
    rng = np.random.default_rng(0)
    block = rng.standard_normal((a.chunks, a.dim)).astype(np.float32)
    block /= np.linalg.norm(block, axis=1, keepdims=True)
    chunk = synth.chunk_ids(S, 0, n)
    docs = synth.doc_ids(S, 0, a.docs)
    out = {"workload": f"{a.docs} documents x {a.chunks} chunks x {a.dim} fp32, one SaveChunks + one SaveEmbeddings per document"}
    with GpuStore(dim=a.dim, capacity_rows=n) as st:
        for phase in ("insert", "upsert_same_chunks"):
            t0 = time.perf_counter()
            for d in range(a.docs):
                ids = chunk[d * a.chunks:(d + 1) * a.chunks]
                if phase == "insert":
                    st.save_chunks_raw(docs[d], ids)
                st.save_embeddings_raw(ids, block)
            dt = time.perf_counter() - t0
            out[phase] = {"seconds": dt, "rows_per_s": n / dt, "GB_per_s": n * a.dim * 4 / dt / 1e9,
                          "ms_per_document": dt * 1e3 / a.docs}
        assert st.stats().rows == n
        big = np.tile(block, (100, 1))
        ids = synth.chunk_ids(S, 10_000_000, big.shape[0])
    with GpuStore(dim=a.dim, capacity_rows=big.shape[0]) as st:
        st.save_chunks_raw(docs[0], ids)
        t0 = time.perf_counter()
        st.save_embeddings_raw(ids, big)
        dt = time.perf_counter() - t0
        out["one_call_20000_rows"] = {"seconds": dt, "rows_per_s": big.shape[0] / dt, "GB_per_s": big.nbytes / dt / 1e9}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
