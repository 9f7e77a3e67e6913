"""Narrow document filters (document_id = ANY of a few ids, postgres.go:237-239): the requested rows are compacted
into a list and scanned through it.  Same answers as the row loop with the bitmap, and as the oracle."""
import numpy as np
import pytest

import oracle
from cases import REL_TOL
from doc_agents_b200 import GpuStore

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    return built_lib


def same(a, b, i, ctx=None):
    assert np.array_equal(a[2][i], b[2][i]), ctx
    assert np.array_equal(a[1][i].view(np.uint32), b[1][i].view(np.uint32)), ctx
    assert np.array_equal(a[0][i], b[0][i]), ctx


def test_compacted_list_equals_bitmap_row_loop_and_oracle():
    S = oracle.Synth(n_clusters=16)
    n, dim = 80_000, 1536                                     # 400 documents of 200 chunks
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        docs = oracle.doc_ids(S, 0, n // 200 + 3)             # the last three are unknown to the store
        qs = np.stack([oracle.gen_query(S, i, i % 16, dim) for i in range(12)])
        filters = [docs[0:1], docs[[0, 16, 32, 399]], docs[[5, 5, 5]], docs[-3:], docs[[7, 401]],
                   docs[np.arange(0, 400, 17)]]               # 1, 4, duplicates, unknown only, mixed, 24 of 400 (= 1/16)
        for f in filters:
            for batch, k, floor in ((1, 10, 0.7), (1, 128, -np.inf), (5, 3, 0.0), (12, 10, 0.7)):
                a = st.search(qs[:batch], k, f, min_similarity=floor)
                b = st.search(qs[:batch], k, f, min_similarity=floor, no_compact=True)
                for i in range(batch):
                    same(a, b, i, (len(f), batch, k, floor, i))
        # the filter pass visits every row's slot once, the scan reads the requested rows only
        st.search(qs[0], 10, docs[[0, 16]])
        assert st.stats().last_rows_read == 400
        # against the oracle
        f = [0, 16, 32, 399]
        rows = np.concatenate([np.arange(d * 200, d * 200 + 200) for d in f])
        sub = np.concatenate([oracle.gen_rows(S, d * 200, 200, dim) for d in f])
        a = st.search(qs[0], 10, docs[f], min_similarity=-np.inf)
        o_rows, o_sc, _ = oracle.topk(sub, None, None, qs[0], 10, min_sim=-np.inf)
        assert np.array_equal(a[2][0], rows[o_rows])
        assert np.allclose(a[1][0], o_sc, rtol=REL_TOL, atol=5e-7)
        # one document set per query: one list per query
        sets = [docs[[i, i + 50]] for i in range(6)]
        a = st.search(qs[:6], 5, sets, min_similarity=-np.inf)
        b = st.search(qs[:6], 5, sets, min_similarity=-np.inf, no_compact=True)
        for i in range(6):
            same(a, b, i)
            assert set(np.asarray(a[2][i]) // 200) <= {i, i + 50}


def test_compaction_follows_upserts():
    rng = np.random.default_rng(3)
    dim, per, n_docs = 128, 50, 40
    n = per * n_docs
    mat = rng.standard_normal((n, dim)).astype(np.float32)
    chunk = rng.integers(0, 256, (n, 16), dtype=np.uint8)
    chunk[:, :4] = np.arange(n, dtype=np.uint32).view(np.uint8).reshape(n, 4)
    docs = rng.integers(0, 256, (n_docs, 16), dtype=np.uint8)
    docs[:, 0] = np.arange(n_docs)
    q = mat[per * 3 + 7] + 0.05 * rng.standard_normal(dim).astype(np.float32)
    with GpuStore(dim=dim, capacity_rows=n) as st:
        for d in range(n_docs):
            st.save_chunks_raw(docs[d], chunk[d * per:(d + 1) * per])
        # documents arrive interleaved: a document's rows are NOT contiguous
        order = rng.permutation(n)
        st.save_embeddings_raw(chunk[order], mat[order])
        a = st.search(q, 5, docs[[3]], min_similarity=-np.inf)
        b = st.search(q, 5, docs[[3]], min_similarity=-np.inf, no_compact=True)
        same(a, b, 0)
        want = np.array([chunk[per * 3 + 7]])
        assert np.array_equal(a[0][0][:1], want)
        sub = mat[per * 3:per * 4]
        o_rows, o_sc, _ = oracle.topk(sub, None, None, q, 5, min_sim=-np.inf)
        assert np.allclose(a[1][0], o_sc, rtol=REL_TOL, atol=5e-7)
