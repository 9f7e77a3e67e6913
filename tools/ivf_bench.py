#!/usr/bin/env python
"""BASELINE configs[4]: IVFFlat k-means build (lists=4096) on N x 3072 with tcgen05 assignment, then
probes=32 query scans; reports build times, assignment TFLOP/s, recall@10 vs the exact scan, query latency.

    python tools/ivf_bench.py [--rows 10000000] [--dim 3072] [--lists 4096] [--probes 32] [--queries 200] [--simt]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from doc_agents_b200 import GpuStore, synth as oracle   # corpus parameters and queries from the product library


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=3072)
    ap.add_argument("--lists", type=int, default=4096)
    ap.add_argument("--probes", type=int, default=32)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--queries", type=int, default=200)
    ap.add_argument("--clusters", type=int, default=1024)
    ap.add_argument("--simt", action="store_true")
    ap.add_argument("--devices", type=int, default=1)
    a = ap.parse_args()
    S = oracle.Synth(n_clusters=a.clusters)
    st = GpuStore(dim=a.dim, capacity_rows=-(-a.rows // a.devices) + 200, devices=tuple(range(a.devices)))
    st.generate_synthetic(S, 0, a.rows)
    t0 = time.perf_counter()
    st.build_ivf(lists=a.lists, iters=a.iters, seed=1, default_probes=a.probes, simt_assign=a.simt)
    wall = time.perf_counter() - t0
    s = st.stats()
    off, _, _ = st.debug_ivf()
    sizes = np.diff(off.astype(np.int64))
    hits = total = 0
    lat_ivf, lat_exact, rows_read = [], [], []
    for i in range(a.queries):
        q = oracle.gen_query(S, i, (i * 37) % S.n_clusters, a.dim)
        ex = st.search(q, 10, all_docs=True)[2][0]
        lat_exact.append(st.stats().last_search_ms)
        ap_ = st.search(q, 10, all_docs=True, ivf=True)[2][0]
        ss = st.stats()
        lat_ivf.append(ss.last_search_ms)
        rows_read.append(ss.last_rows_read)
        hits += len(np.intersect1d(ex, ap_))
        total += len(ex)
    out = {
        "workload": f"IVFFlat lists={a.lists} on {a.rows}x{a.dim}, k-means {a.iters} iters, probes={a.probes}, {a.queries} queries, {a.devices} GPU(s) in one process",
        "assign": "simt fp32" if a.simt else "tcgen05 tf32",
        "build_ms": s.ivf_build_ms, "kmeans_ms": s.ivf_kmeans_ms, "assign_all_rows_ms": s.ivf_assign_ms,
        "assign_tflops": s.ivf_assign_tflops, "build_wall_s": wall,
        "list_size": {"min": int(sizes.min()), "median": float(np.median(sizes)), "max": int(sizes.max()),
                      "empty": int((sizes == 0).sum())},
        "recall_at_10": hits / max(total, 1),
        "ivf_query_ms_median": float(np.median(lat_ivf)), "exact_query_ms_median": float(np.median(lat_exact)),
        "rows_read_mean": float(np.mean(rows_read)), "rows_read_frac": float(np.mean(rows_read)) / a.rows,
    }
    print(json.dumps(out))
    st.close()


if __name__ == "__main__":
    main()
