"""Phase times of the scan kernel's last CTA (prologue, row loop, CTA merge + ticket, last-CTA merge + ids, fence), printed by the
kernel itself when the library is built with DA_NVCC_EXTRA=-DDA_SCAN_TIMING (clock64 stamps, da_topk.cuh): IVFFlat probe scans,
a 16-document small search and full scans at 10M x 3072.  Measurement only."""
import sys, os, numpy as np
sys.path.insert(0, "/root/repo")
from doc_agents_b200 import GpuStore, synth
S = synth.Synth()
n = int(os.environ.get("ROWS", 10_000_000))
st = GpuStore(dim=3072, capacity_rows=n)
st.generate_synthetic(S, 0, n)
st.build_ivf(lists=4096, iters=10, seed=1, default_probes=32)
docs = synth.doc_ids(S, 0, 50000)
for p in (32, 1):
    print("== ivf probes", p, flush=True)
    for i in range(4):
        q = synth.gen_query(S, i, (i * 37) % 1024, 3072)
        st.search_arrays(q, 10, all_docs=True, ivf=True, probes=p)
print("== 16 documents (cfg1-like), small-search path", flush=True)
for i in range(4):
    q = synth.gen_query(S, i, 5, 3072)
    sel = docs[[5 + 1024 * j for j in range(16)]]
    st.search_arrays(q, 10, sel)
print("== exact scan, all rows", flush=True)
for i in range(3):
    st.search_arrays(synth.gen_query(S, i, 5, 3072), 10, all_docs=True)
