import sys, os, numpy as np
sys.path.insert(0, "/root/repo")
import doc_agents_b200._lib as _L
_L.LIB_PATH = os.environ.get("DA_LIB", _L.LIB_PATH)      # A/B against another build of the library
from doc_agents_b200 import GpuStore, synth
S = synth.Synth()
n = int(os.environ.get("ROWS", 10_000_000))
st = GpuStore(dim=3072, capacity_rows=n)
st.generate_synthetic(S, 0, n)
st.build_ivf(lists=4096, iters=10, seed=1, default_probes=32)
for i in range(6):
    q = synth.gen_query(S, i, (i * 37) % 1024, 3072)
    st.search_arrays(q, 10, all_docs=True, ivf=True)
for rep in range(2):
    for p in (1, 8, 32):
        ms = st.bench_resident(synth.gen_query(S, 3, 111, 3072), 10, 200, all_docs=True, ivf=True, probes=p)
        print("probes", p, "resident ms per query %.4f" % (ms / 200), "rows", st.stats().last_rows_read)
