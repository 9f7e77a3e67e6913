"""The READER process of tests/test_gpu_ipc.py: plays the query agent.  It attaches to the store another process
exported (da_store_attach) and answers TopK requests read from stdin, one JSON object per line:
    {"docs": [document indexes], "qi": query number, "cluster": c, "k": k}   ->   {"rows": [...], "scores": [...], ...}
    {"quit": true}
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import oracle
    from doc_agents_b200 import GpuStore, StoreError

    name, dim = sys.argv[1], int(sys.argv[2])
    S = oracle.Synth(n_clusters=8)
    st = GpuStore.attach(name)
    print(json.dumps({"attached": True, "rows": int(st.stats().rows), "docs": int(st.stats().n_docs)}), flush=True)
    for line in sys.stdin:
        req = json.loads(line)
        if req.get("quit"):
            break
        if req.get("write"):                       # an attached store is read-only
            try:
                st.save_chunks_raw(oracle.doc_ids(S, 999, 1)[0], oracle.chunk_ids(S, 999_000, 2))
                print(json.dumps({"refused": False}), flush=True)
            except StoreError as e:
                print(json.dumps({"refused": True, "code": e.code}), flush=True)
            continue
        q = oracle.gen_query(S, req["qi"], req["cluster"], dim)
        docs = oracle.doc_ids(S, 0, 4096)[np.asarray(req["docs"], dtype=np.int64)]
        kw = dict(no_compact=True) if req.get("general_path") else {}
        ids, sc, rows = st.search(q, req["k"], docs, **kw)
        s = st.stats()
        print(json.dumps({"rows": [int(r) for r in rows[0]], "scores": [float(x) for x in sc[0]],
                          "ids": [bytes(i).hex() for i in ids[0]], "store_rows": int(s.rows), "store_docs": int(s.n_docs),
                          "doc_rows": [int(x) for x in st.doc_rows(docs)], "small": int(s.small_searches)}), flush=True)
    st.close()


if __name__ == "__main__":
    main()
