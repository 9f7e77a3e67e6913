"""One batched IVFFlat search (64 queries, lists 4096, probes 32) on a 2M x 3072 store: the workload of the ncu
captures of tc_scores_kernel, select_probes_kernel and the probe-mode scan (profiles/r2_ivf_batch_*)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from doc_agents_b200 import GpuStore, synth

S = synth.Synth()
n = int(os.environ.get("ROWS", 2_000_000))
st = GpuStore(dim=3072, capacity_rows=n)
st.generate_synthetic(S, 0, n)
st.build_ivf(lists=4096, iters=4, seed=1, default_probes=32)
qs = np.stack([synth.gen_query(S, i, (i * 37) % 1024, 3072) for i in range(64)])
for _ in range(3):
    out = st.search_arrays(qs, 10, all_docs=True, ivf=True)
print("hits", int(out[3].sum()), "rows read", int(st.stats().last_rows_read))
