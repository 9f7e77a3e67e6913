"""GPU tests of the small-search path (one query on a single-shard store: one copy, one kernel that walks the
requested documents' row ranges, result picked up from mapped memory): the reference's own regime, TopK with 1..16
document ids (cmd/query/main.go:20-24, internal/store/postgres.go:237-239).  Checked against the oracle, against the
general path (bitmap scan) for bit-equality, with interleaved and fragmented documents, and from 16 threads at once."""
import threading

import numpy as np
import pytest

import oracle
from cases import topk_matches_extended
from doc_agents_b200 import GpuStore
from gpu_util import S0


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    return built_lib


pytestmark = pytest.mark.gpu


def _oracle_rows(mat, row_doc, n_docs, wanted, q, k, floor=0.7):
    allowed = np.zeros(n_docs, np.uint8)
    allowed[[d for d in wanted if 0 <= d < n_docs]] = 1
    r, s, _ = oracle.topk(mat, row_doc, allowed, q, k, floor)
    return r, s


def test_range_scan_equals_oracle_and_the_bitmap_scan():
    S = oracle.Synth(n_clusters=8)
    n, dim, k = 12_000, 3072, 10
    n_docs = n // 200
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        mat = oracle.gen_rows(S, 0, n, dim)
        rd = oracle.row_docs(S, 0, n)
        docs = oracle.doc_ids(S, 0, n_docs + 50)
        for qi, wanted in enumerate(([0], [3, 11, 19, 27], list(range(0, n_docs, 2)), [5, 5, 5, 13], [n_docs + 7], [],
                                     [1, n_docs + 3, 9, 17, 25, 33, 41, 49, 57, 2, 10, 18, 26, 34, 42, 50])):
            q = oracle.gen_query(S, qi, qi % 8, dim)
            sel = docs[np.asarray(wanted, dtype=np.int64)] if wanted else docs[:0]
            before = st.stats().small_searches
            ids, sc, rows = st.search(q, k, sel)
            assert st.stats().small_searches == before + 1
            o_rows, o_sc = _oracle_rows(mat, rd, n_docs, wanted, q, k)
            assert np.array_equal(rows[0], o_rows), (wanted, rows[0], o_rows)
            assert np.allclose(sc[0], o_sc, rtol=1e-5)
            assert np.array_equal(ids[0], oracle.chunk_ids(S, 0, n)[o_rows])
            assert st.stats().last_rows_read == 200 * len({d for d in wanted if d < n_docs})
            # the general path (bitmap per document slot) gives the very same bits
            g = st.search(q, k, sel, no_compact=True)
            assert st.stats().small_searches == before + 1
            assert np.array_equal(g[2][0], rows[0]) and np.array_equal(g[1][0], sc[0]) and np.array_equal(g[0][0], ids[0])
        # every document requested, and k = 0, through the same path
        q = oracle.gen_query(S, 40, 3, dim)
        a = st.search(q, k, all_docs=True)
        o_rows, o_sc, _ = oracle.topk(mat, None, None, q, k)
        assert np.array_equal(a[2][0], o_rows) and np.allclose(a[1][0], o_sc, rtol=1e-5)
        assert st.stats().last_rows_read == n
        assert len(st.search(q, 0, docs[:3])[2][0]) == 0
        assert st.TopK([], q, 5) is None


def test_documents_saved_in_pieces_and_moved_chunks():
    """Ranges are learned from SaveEmbeddings: a document saved in interleaved pieces has several ranges, one saved in
    more than eight pieces (or with a chunk moved to another document) is served by the per-row document slots."""
    dim, k = 512, 8
    rng = np.random.default_rng(3)
    q = oracle.normalize(rng.standard_normal(dim).astype(np.float32))
    n_docs, per = 6, 40
    vecs = np.stack([oracle.normalize((0.9 * q + rng.standard_normal(dim).astype(np.float32) * 0.02).astype(np.float32))
                     for _ in range(n_docs * per)])
    chunk = oracle.chunk_ids(S0, 0, n_docs * per)
    docs = oracle.doc_ids(S0, 0, n_docs)
    doc_of = np.repeat(np.arange(n_docs), per)
    with GpuStore(dim=dim, capacity_rows=1024) as st:
        for d in range(n_docs):
            st.save_chunks_raw(docs[d], chunk[doc_of == d])
        # documents 0 and 1 alternate in pieces of 10 (4 ranges each); 2 and 3 in pieces of 4 (10 pieces: fragmented);
        # 4 and 5 whole
        order = []
        for piece in range(4):
            for d in (0, 1):
                order += [d * per + piece * 10 + i for i in range(10)]
        for piece in range(10):
            for d in (2, 3):
                order += [d * per + piece * 4 + i for i in range(4)]
        order += list(range(4 * per, 6 * per))
        pos = 0
        while pos < len(order):           # one SaveEmbeddings per run of one document
            end = pos
            while end < len(order) and doc_of[order[end]] == doc_of[order[pos]]:
                end += 1
            st.save_embeddings_raw(chunk[order[pos:end]], vecs[order[pos:end]])
            pos = end
        stored = vecs[order]                          # insertion row i holds vecs[order[i]]
        stored_doc = doc_of[order].astype(np.uint32)
        stored_chunk = chunk[order]
        for wanted, small in (([0], True), ([1, 4], True), ([2], False), ([0, 3], False), ([4, 5], True)):
            before = st.stats().small_searches
            ids, sc, rows = st.search(q, k, docs[wanted], min_similarity=0.5)
            assert (st.stats().small_searches == before + 1) == small, wanted
            o_rows, o_sc = _oracle_rows(stored, stored_doc, n_docs, wanted, q, k, 0.5)
            assert np.array_equal(rows[0], o_rows) and np.allclose(sc[0], o_sc, rtol=1e-5)
            assert np.array_equal(ids[0], stored_chunk[o_rows])
        # a chunk of document 4 moves to document 5: both leave the range path, answers stay right
        moved = order.index(4 * per + 3)
        st.save_chunks_raw(docs[5], chunk[[4 * per + 3]])
        stored_doc[moved] = 5
        assert list(st.doc_rows(docs[[4, 5]])) == [per - 1, per + 1]
        for wanted in ([4], [5]):
            before = st.stats().small_searches
            ids, sc, rows = st.search(q, k, docs[wanted], min_similarity=0.5)
            assert st.stats().small_searches == before
            o_rows, o_sc = _oracle_rows(stored, stored_doc, n_docs, wanted, q, k, 0.5)
            assert np.array_equal(rows[0], o_rows) and np.allclose(sc[0], o_sc, rtol=1e-5)


def test_sixteen_threads_call_topk_at_once():
    S = oracle.Synth(n_clusters=8)
    n, dim, k = 40_000, 1536, 5
    n_docs = n // 200
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        mat = oracle.gen_rows(S, 0, n, dim)
        rd = oracle.row_docs(S, 0, n)
        docs = oracle.doc_ids(S, 0, n_docs)
        chunk = oracle.chunk_ids(S, 0, n)
        row_of = {chunk[i].tobytes(): i for i in range(n)}
        import uuid
        errors = []

        def worker(t):
            try:
                for it in range(25):
                    cl = (t + it) % 8
                    wanted = [(cl + 8 * j) % n_docs for j in range(1 + (t + it) % 16)]
                    q = oracle.gen_query(S, 1000 + 37 * t + it, cl, dim)
                    res = st.TopK([uuid.UUID(bytes=docs[d].tobytes()) for d in wanted], q, k)
                    o_rows, o_sc = _oracle_rows(mat, rd, n_docs, wanted, q, k + 8)     # past k: a tie across the cut shows
                    got = [] if res is None else res
                    got_rows = [row_of[r.Chunk.ID.bytes] for r in got]
                    why = topk_matches_extended(got_rows, [r.Score for r in got], o_rows, o_sc, k)
                    assert why == "", (t, it, why)
            except Exception as e:       # noqa: BLE001
                errors.append(repr(e))

        ts = [threading.Thread(target=worker, args=(t,)) for t in range(16)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        assert not errors, errors[:3]
        s = st.stats()
        assert s.small_searches >= 16 * 25 and s.topk_calls == 16 * 25
