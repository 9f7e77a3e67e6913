#!/usr/bin/env python
"""One process driving G GPUs (the Go query agent's deployment): queries/s of the exact scan over the
row-sharded 10M x 3072 corpus with the peer-gather path (scan epilogue stores into shard 0 over NVLink)
and with the NCCL all-gather path.      python tools/inproc_bench.py --devices 8 [--rows 10000000]"""
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
    ap.add_argument("--devices", type=int, default=2)
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=3072)
    ap.add_argument("--steps", type=int, default=50)
    a = ap.parse_args()
    S = oracle.Synth()
    out = {"workload": f"exact scan, {a.rows}x{a.dim}, batch=1, top_k=10, one process, {a.devices} GPUs"}
    for mode in ("peer_gather", "nccl_allgather"):
        if mode == "nccl_allgather":
            os.environ["DA_DISABLE_P2P"] = "1"
        st = GpuStore(dim=a.dim, capacity_rows=-(-a.rows // a.devices) + 200, devices=tuple(range(a.devices)))
        os.environ.pop("DA_DISABLE_P2P", None)
        st.generate_synthetic(S, 0, a.rows)
        qs = [oracle.gen_query(S, i, (i * 37) % S.n_clusters, a.dim) for i in range(a.steps + 3)]
        for i in range(3):
            st.search(qs[i], 10, all_docs=True)
        st.bench_resident(qs[0], 10, 5, all_docs=True)
        ms = st.bench_resident(qs[1], 10, a.steps, all_docs=True)
        t0 = time.perf_counter()
        for i in range(a.steps):
            st.search(qs[3 + i], 10, all_docs=True)
        e2e = time.perf_counter() - t0
        s = st.stats()
        out[mode] = {"peer_gather": s.peer_gather, "resident_ms_per_query": ms / a.steps, "resident_qps": a.steps / (ms / 1e3),
                     "e2e_ms_per_query": e2e * 1e3 / a.steps, "e2e_qps": a.steps / e2e, "scan_ms_shard0": s.last_scan_ms}
        st.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
