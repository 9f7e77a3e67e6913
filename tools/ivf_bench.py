#!/usr/bin/env python
"""BASELINE configs[4]: IVFFlat k-means build (lists=4096) on N x 3072 with tcgen05 assignment, then probed scans.
Reports build times, assignment TFLOP/s, list balance, recall@10 vs the exact scan at several probe counts, the latency
of single queries, and a batch of queries (one tensor-core ranking GEMM + one probe-mode scan) against the same
queries asked one at a time.

    python tools/ivf_bench.py [--rows 10000000] [--dim 3072] [--lists 4096] [--probes 32] [--queries 200]
                              [--preset baseline|two_level] [--simt] [--no-split]

--preset two_level: 64 clusters in groups of 8 that share a direction (same-cluster cosine ~0.67, same-group ~0.33,
no similarity floor) -- a corpus on which an approximate index CAN lose recall (on the baseline corpus, 1 024
near-orthogonal clusters, almost any index scores 1.0).
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
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--clusters", type=int, default=0)
    ap.add_argument("--preset", default="baseline", choices=["baseline", "two_level"])
    ap.add_argument("--simt", action="store_true")
    ap.add_argument("--no-split", action="store_true")
    ap.add_argument("--devices", type=int, default=1)
    a = ap.parse_args()
    if a.preset == "two_level":
        S = oracle.Synth(n_clusters=a.clusters or 64, cmul=1, nmul=1, smul=1, super_group=8)
        floor = float("-inf")
    else:
        S = oracle.Synth(n_clusters=a.clusters or 1024)
        floor = 0.7
    st = GpuStore(dim=a.dim, capacity_rows=-(-a.rows // a.devices) + 200, devices=tuple(range(a.devices)), min_similarity=floor)
    st.generate_synthetic(S, 0, a.rows)
    t0 = time.perf_counter()
    st.build_ivf(lists=a.lists, iters=a.iters, seed=1, default_probes=a.probes, simt_assign=a.simt, split=not a.no_split)
    wall = time.perf_counter() - t0
    s = st.stats()
    off, _, _ = st.debug_ivf()
    sizes = np.diff(off.astype(np.int64))
    qs = np.stack([oracle.gen_query(S, i, (i * 37) % S.n_clusters, a.dim) for i in range(a.queries)])
    exact, lat_exact = [], []
    for i in range(a.queries):
        exact.append(st.search(qs[i], 10, all_docs=True)[2][0])
        lat_exact.append(st.stats().last_search_ms)
    recall, lat, rows_read, bytes_q = {}, {}, {}, {}
    for probes in sorted({1, 4, 8, a.probes}):
        hits = total = 0
        l, r, b = [], [], []
        for i in range(a.queries):
            t1 = time.perf_counter()
            got = st.search_arrays(qs[i], 10, all_docs=True, ivf=True, probes=probes)
            wall_ms = (time.perf_counter() - t1) * 1e3
            ss = st.stats()
            l.append((ss.last_search_ms, wall_ms))
            r.append(ss.last_rows_read)
            b.append(ss.last_bytes_algorithmic)
            hits += len(np.intersect1d(exact[i], got[2][0, :got[3][0]]))
            total += len(exact[i])
        recall[probes] = hits / max(total, 1)
        lat[probes] = (float(np.median([x[0] for x in l])), float(np.median([x[1] for x in l])))
        rows_read[probes] = float(np.mean(r))
        bytes_q[probes] = float(np.mean(b))
    # a batch through the tensor-core ranking + ONE probe-mode scan, against the same queries one at a time
    nb = min(a.batch, a.queries)
    st.search_arrays(qs[:nb], 10, all_docs=True, ivf=True, probes=a.probes)
    batch_ms = min(st.bench_resident(qs[:nb], 10, 3, all_docs=True, ivf=True, probes=a.probes) / 3 for _ in range(3))
    loop_ms = sum(st.bench_resident(qs[i], 10, 1, all_docs=True, ivf=True, probes=a.probes) for i in range(nb))
    t1 = time.perf_counter()
    for i in range(nb):
        st.search_arrays(qs[i], 10, all_docs=True, ivf=True, probes=a.probes)
    loop_e2e_ms = (time.perf_counter() - t1) * 1e3
    t1 = time.perf_counter()
    bgot = st.search_arrays(qs[:nb], 10, all_docs=True, ivf=True, probes=a.probes)
    batch_e2e_ms = (time.perf_counter() - t1) * 1e3
    bhits = sum(len(np.intersect1d(exact[i], bgot[2][i, :bgot[3][i]])) for i in range(nb))
    single_ms, single_wall = lat[a.probes]
    out = {
        "workload": f"IVFFlat lists={a.lists} on {a.rows}x{a.dim} ({a.preset} corpus, {S.n_clusters} clusters), k-means {a.iters} iters"
                    f"{'' if a.no_split else ' + balancing rounds'}, {a.queries} queries, top-10, {a.devices} GPU(s) in one process",
        "assign": "simt fp32" if a.simt else "tcgen05 tf32",
        "build_ms": s.ivf_build_ms, "kmeans_ms": s.ivf_kmeans_ms, "assign_all_rows_ms": s.ivf_assign_ms,
        "assign_tflops": s.ivf_assign_tflops, "build_wall_s": wall,
        "list_size": {"min": int(sizes.min()), "median": float(np.median(sizes)), "max": int(sizes.max()),
                      "empty": int((sizes == 0).sum()), "empty_frac": float((sizes == 0).mean())},
        "recall_at_10": {str(p): recall[p] for p in recall},
        "rows_read_mean": {str(p): rows_read[p] for p in rows_read},
        "ivf_query_ms_median": single_ms, "ivf_query_wall_ms_median": single_wall,
        "ivf_query_gbs": bytes_q[a.probes] / (single_ms * 1e-3) / 1e9 if single_ms > 0 else None,
        "ivf_query_bytes": bytes_q[a.probes],
        "query_ms_median_by_probes": {str(p): lat[p][0] for p in lat},
        "exact_query_ms_median": float(np.median(lat_exact)),
        "batch": {"queries": nb, "probes": a.probes, "resident_ms": batch_ms, "one_at_a_time_resident_ms": loop_ms,
                  "speedup_resident": loop_ms / batch_ms if batch_ms > 0 else None,
                  "e2e_ms": batch_e2e_ms, "one_at_a_time_e2e_ms": loop_e2e_ms,
                  "recall_at_10": bhits / max(1, sum(len(exact[i]) for i in range(nb)))},
    }
    print(json.dumps(out))
    st.close()


if __name__ == "__main__":
    main()
