"""GPU suite for the batched path: tcgen05 TF32 candidate generation + exact fp32 re-score must
return exactly what the fp32 SIMT scan returns (same rows, same order, bit-identical scores), and
what the CPU oracle returns within the parity tolerance."""
import numpy as np
import pytest

import oracle
from cases import REL_TOL, same_rows_up_to_ties
from doc_agents_b200 import GpuStore


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    """Compile the CUDA library once if it is stale (the GPU box normally receives it pre-built)."""
    return built_lib


pytestmark = pytest.mark.gpu


def _compare(st, qs, k, **kw):
    tc = st.search(qs, k, **kw)
    assert st.stats().last_rows_read == st.stats().rows               # the contraction visits every row
    simt = st.search(qs, k, no_tensor=True, **kw)
    for i in range(len(qs)):
        assert np.array_equal(tc[2][i], simt[2][i]), (i, tc[2][i], simt[2][i])
        assert np.array_equal(tc[1][i].view(np.uint32), simt[1][i].view(np.uint32)), i
        assert np.array_equal(tc[0][i], simt[0][i]), i
    return tc


@pytest.mark.parametrize("batch", [16, 64, 130, 300])
@pytest.mark.parametrize("dim", [1536, 3072])
def test_tensor_batch_equals_simt_scan_and_oracle(batch, dim):
    S = oracle.Synth(n_clusters=16)
    n, k = 30_000, 10
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        qs = np.stack([oracle.gen_query(S, i, i % 16, dim) for i in range(batch)])
        tc = _compare(st, qs, k, all_docs=True)
        mat = oracle.gen_rows(S, 0, n, dim)
        for i in (0, batch // 2, batch - 1):
            o_rows, o_sc, _ = oracle.topk(mat, None, None, qs[i], k)
            assert same_rows_up_to_ties(tc[2][i], o_rows, o_sc)
            assert np.allclose(tc[1][i], o_sc, rtol=REL_TOL)
        # a wide document filter (2 of 3 documents) stays on the tensor path; a narrow one does not
        docs = oracle.doc_ids(S, 0, n // 200)
        wide = docs[[d for d in range(n // 200) if d % 3 != 1]]
        _compare(st, qs, k, doc_ids=wide)
        narrow = st.search(qs, k, doc_ids=docs[:5])
        assert st.stats().last_rows_read == 5 * 200 * ((batch + 7) // 8)   # compacted list of 1000 rows, 8 queries per pass
        assert all(len(r) <= k for r in narrow[2])
        loop = st.search(qs, k, doc_ids=docs[:5], no_compact=True)
        assert st.stats().last_rows_read == 5 * 200 * ((batch + 7) // 8)   # SIMT row loop, rows of other documents skipped
        for i in range(batch):
            assert np.array_equal(narrow[2][i], loop[2][i]) and np.array_equal(narrow[1][i], loop[1][i])


def test_tensor_batch_no_floor_top100_exercises_pruning():
    """No similarity floor, k = 100, ~2700 rows per CTA: every candidate list fills and is pruned several times."""
    S = oracle.Synth(n_clusters=64)
    n, dim, k, batch = 400_000, 256, 100, 40
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        qs = np.stack([oracle.gen_query(S, i, (5 * i) % 64, dim) for i in range(batch)])
        tc = _compare(st, qs, k, all_docs=True, min_similarity=-np.inf)
        assert all(len(r) == k for r in tc[2])
        mat = oracle.gen_rows(S, 0, n, dim)
        for i in (0, 17, 39):
            o_rows, o_sc, _ = oracle.topk(mat, None, None, qs[i], k, -np.inf)
            assert np.allclose(tc[1][i], o_sc, rtol=REL_TOL, atol=5e-7)
            assert len(np.intersect1d(tc[2][i], o_rows)) >= k - 1         # a tie at the k-th place may swap one row


def test_tensor_batch_small_corpus_and_padding_queries():
    S = oracle.Synth(n_clusters=2)
    n, dim = 700, 100                                                    # fewer rows than one 256-row tile per CTA, ragged K
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        qs = np.stack([oracle.gen_query(S, i, i % 2, dim) for i in range(17)])
        _compare(st, qs, 5, all_docs=True)
        _compare(st, qs, 128, all_docs=True, min_similarity=-np.inf)


@pytest.mark.parametrize("batch", [3, 9, 24, 150])
def test_per_query_document_sets_equal_single_searches(batch):
    """Each query of a batch names its own documents (what merged /api/query requests need): SIMT passes
    (batch < 16) and the tensor path (lazy per-query mask in the epilogue) against one search per query."""
    S = oracle.Synth(n_clusters=8)
    n, dim, k = 24_000, 1536, 10
    n_docs = n // 200
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        docs = oracle.doc_ids(S, 0, n_docs)
        rng = np.random.default_rng(batch)
        qs = np.stack([oracle.gen_query(S, i, i % 8, dim) for i in range(batch)])
        sets = []
        for i in range(batch):
            if i % 7 == 3:
                pick = np.zeros(0, dtype=np.int64)                             # an empty set: nil for that query
            elif i % 5 == 0:
                pick = np.arange(i % 8, n_docs, 8)                            # exactly the query's cluster
            else:
                pick = np.sort(rng.choice(n_docs, size=int(rng.integers(n_docs // 3, n_docs)), replace=False))
            sets.append(docs[pick])
        got = st.search(qs, k, sets)
        path_rows = st.stats().last_rows_read
        for i in range(batch):
            one = st.search(qs[i], k, sets[i])
            assert np.array_equal(got[2][i], one[2][0]), (i, got[2][i], one[2][0])
            assert np.array_equal(got[1][i].view(np.uint32), one[1][0].view(np.uint32)), i
            assert np.array_equal(got[0][i], one[0][0]), i
        if batch >= 16:
            assert path_rows == n                                              # union is wide: tensor path
            simt = st.search(qs, k, sets, no_tensor=True)
            for i in range(batch):
                assert np.array_equal(got[2][i], simt[2][i])
