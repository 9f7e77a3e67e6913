"""Two processes, one device store.  In the reference the analysis agent saves embeddings and the
query agent searches them from another process (docker-compose.yml:87-128, cmd/analysis/main.go:106,
cmd/query/main.go:101).  Here this process is the WRITER (da_store_export) and tests/ipc_reader.py, a separate
process, the READER (da_store_attach): what the writer saves must be found by the reader's next TopK, with the oracle
as the checker."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
from doc_agents_b200 import GpuStore, StoreError
from doc_agents_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    return built_lib


pytestmark = pytest.mark.gpu


class Reader:
    def __init__(self, name, dim):
        self.p = subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "ipc_reader.py"), name, str(dim)],
                                  stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True)
        self.hello = self._read()

    def _read(self):
        line = self.p.stdout.readline()
        assert line, "the reader process died"
        return json.loads(line)

    def ask(self, **req):
        self.p.stdin.write(json.dumps(req) + "\n")
        self.p.stdin.flush()
        return self._read()

    def close(self):
        try:
            self.p.stdin.write(json.dumps({"quit": True}) + "\n")
            self.p.stdin.flush()
            self.p.wait(timeout=60)
        finally:
            if self.p.poll() is None:
                self.p.kill()


def test_reader_process_finds_what_the_writer_process_saves():
    S = oracle.Synth(n_clusters=8)
    dim, k, per = 1024, 6, 200
    name = f"/docagents-test-{os.getpid()}"
    n0 = 10 * per                                            # documents 0..9 exist before the export
    with GpuStore(dim=dim, capacity_rows=64 * per) as st:
        mat = oracle.gen_rows(S, 0, 40 * per, dim)
        chunk = oracle.chunk_ids(S, 0, 40 * per)
        docs = oracle.doc_ids(S, 0, 64)
        rd = oracle.row_docs(S, 0, 40 * per)

        def save(d):
            st.save_chunks_raw(docs[d], chunk[d * per:(d + 1) * per])
            st.save_embeddings_raw(chunk[d * per:(d + 1) * per], mat[d * per:(d + 1) * per])

        for d in range(10):
            save(d)
        with pytest.raises(StoreError):
            GpuStore.attach(name)                            # nothing exported yet
        st.export(name)
        rdr = Reader(name, dim)
        try:
            assert rdr.hello == {"attached": True, "rows": n0, "docs": 10}

            def check(ans, wanted, qi, cluster, upto_doc):
                n = upto_doc * per
                allowed = np.zeros(upto_doc, np.uint8)
                allowed[[d for d in wanted if d < upto_doc]] = 1
                o_rows, o_sc, _ = oracle.topk(mat[:n], rd[:n], allowed, oracle.gen_query(S, qi, cluster, dim), k)
                assert ans["rows"] == [int(r) for r in o_rows], (ans["rows"], o_rows)
                assert np.allclose(ans["scores"], o_sc, rtol=1e-5)
                assert ans["ids"] == [bytes(chunk[r]).hex() for r in o_rows]

            a = rdr.ask(docs=[0, 8, 3], qi=1, cluster=0, k=k)
            check(a, [0, 8, 3], 1, 0, 10)
            assert a["doc_rows"] == [per, per, per] and a["small"] >= 1
            # a document the writer has not saved yet: unknown to the reader, nothing found
            a = rdr.ask(docs=[16], qi=2, cluster=0, k=k)
            assert a["rows"] == [] and a["doc_rows"] == [0]
            # the writer saves documents 10..19 (one SaveEmbeddings each, as the analysis agent does) ...
            for d in range(10, 20):
                save(d)
            # ... and the reader's NEXT search sees them
            a = rdr.ask(docs=[16], qi=2, cluster=0, k=k)
            assert a["store_rows"] == 20 * per and a["store_docs"] == 20 and a["doc_rows"] == [per]
            check(a, [16], 2, 0, 20)
            a = rdr.ask(docs=[0, 8, 16, 11], qi=3, cluster=3, k=k)
            check(a, [0, 8, 16, 11], 3, 3, 20)
            a = rdr.ask(docs=[0, 8, 16, 11], qi=3, cluster=3, k=k, general_path=True)      # bitmap path on the shared arrays
            check(a, [0, 8, 16, 11], 3, 3, 20)
            # an upsert that overwrites rows in place is seen too: document 3 gets document 24's vectors (cluster 0)
            st.save_embeddings_raw(chunk[3 * per:4 * per], mat[24 * per:25 * per])
            mat[3 * per:4 * per] = mat[24 * per:25 * per]
            a = rdr.ask(docs=[3, 1], qi=4, cluster=0, k=k)
            check(a, [3, 1], 4, 0, 20)
            # the writer answers the same as the reader
            w = st.search(oracle.gen_query(S, 4, 0, dim), k, docs[[3, 1]])
            assert [int(r) for r in w[2][0]] == a["rows"]
            # the reader may not write
            assert rdr.ask(write=True) == {"refused": True, "code": L.DA_ERR_STATE}
        finally:
            rdr.close()
    assert not os.path.exists("/dev/shm" + name)             # the exporter removed the name
