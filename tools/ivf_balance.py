import sys, os, numpy as np, json
sys.path.insert(0, "/root/repo")
from doc_agents_b200 import GpuStore, synth
preset = os.environ.get("PRESET", "baseline")
S = synth.Synth() if preset == "baseline" else synth.Synth(n_clusters=64, cmul=1, nmul=1, smul=1, super_group=8)
n = int(os.environ.get("ROWS", 10_000_000))
st = GpuStore(dim=3072, capacity_rows=n, min_similarity=0.7 if preset == "baseline" else float("-inf"))
st.generate_synthetic(S, 0, n)
st.build_ivf(lists=4096, iters=10, seed=1, default_probes=32)
off, _, _ = st.debug_ivf()
sz = np.diff(off.astype(np.int64))
qs = np.stack([synth.gen_query(S, i, (i * 37) % S.n_clusters, 3072) for i in range(100)])
hits = tot = 0; rr = []
for i in range(100):
    ex = st.search(qs[i], 10, all_docs=True)[2][0]
    ap = st.search(qs[i], 10, all_docs=True, ivf=True)[2][0]
    rr.append(st.stats().last_rows_read)
    hits += len(np.intersect1d(ex, ap)); tot += len(ex)
print(json.dumps({"env": {k: v for k, v in os.environ.items() if k.startswith("DA_IVF")}, "preset": preset, "empty": int((sz == 0).sum()), "min": int(sz.min()),
                  "p1": float(np.percentile(sz, 1)), "median": float(np.median(sz)), "p99": float(np.percentile(sz, 99)), "max": int(sz.max()),
                  "recall32": hits / tot, "rows_read_mean": float(np.mean(rr)), "rows_read_max": int(np.max(rr)), "build_ms": st.stats().ivf_build_ms}))
