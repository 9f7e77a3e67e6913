"""GPU parity suite (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle on
the same seeded inputs, against the committed golden fixtures, and -- at BASELINE.json's full
sizes -- through checks that stay cheap for the oracle (a document filter or the similarity floor
restricts the rows that can qualify to a few thousand)."""
import json
import os
import uuid

import numpy as np
import pytest

import oracle
from cases import ABS_TOL, REL_TOL, bits_f32, build_cases, check_against_oracle, same_rows_up_to_ties, topk_matches_extended
from doc_agents_b200 import Chunk, Embedding, GpuStore, StoreError, Summary, calculate_avg_similarity
from doc_agents_b200 import _lib as L
from gpu_util import S0, device_count, doc_filter, ids_for, load_case


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    """Compile the CUDA library once if it is stale (the GPU box normally receives it pre-built)."""
    return built_lib


pytestmark = pytest.mark.gpu

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "topk_cases.json")))
CASES = build_cases()


@pytest.mark.parametrize("case", CASES, ids=lambda c: c.name)
def test_case_matches_oracle_and_golden(case):
    n, dim = case.mat.shape
    with GpuStore(dim=dim, capacity_rows=n + 8) as st:
        chunk, docs = load_case(st, case)
        ids, scores, rows = st.search(case.q, case.k, doc_filter(case, docs), min_similarity=case.min_sim)
        check_against_oracle(case, rows[0], scores[0])
        assert np.array_equal(ids[0], chunk[rows[0]])                   # chunk ids gathered on the device
        g = GOLDEN["cases"][case.name]
        assert len(rows[0]) == len(g["f32"]["rows"])
        assert np.allclose(scores[0], bits_f32(g["f32"]["score_bits"]), rtol=REL_TOL, atol=ABS_TOL)
        assert np.allclose(scores[0], bits_f32(g["f64"]["score_bits"]), rtol=REL_TOL, atol=ABS_TOL)
        assert same_rows_up_to_ties(rows[0], g["f64"]["rows"], bits_f32(g["f64"]["score_bits"]))
        assert same_rows_up_to_ties(rows[0], g["f32"]["rows"], bits_f32(g["f32"]["score_bits"]))
        # the mask-as-predicate scan (reads every row) returns the same thing
        ids2, scores2, rows2 = st.search(case.q, case.k, doc_filter(case, docs), min_similarity=case.min_sim,
                                         read_all_rows=True)
        assert np.array_equal(rows2[0], rows[0]) and np.array_equal(scores2[0], scores[0])
        assert st.stats().last_rows_read == n


def test_store_interface_mirrors_the_reference_contract():
    """Replays what the reference's tests pin about TopK/SaveEmbeddings (cmd/query/main_test.go:72-79,
    :226-254; cmd/analysis/main_test.go:76-78,:231) against the real C ABI."""
    case = CASES[0]
    n, dim = case.mat.shape
    doc = uuid.UUID(bytes=bytes(oracle.doc_ids(S0, 0, 1)[0]))
    other = uuid.uuid4()
    with GpuStore(dim=dim, capacity_rows=512) as st:
        saved = st.SaveChunks(doc, [Chunk(uuid.UUID(int=0), Text=f"chunk {i}", TokenCount=3) for i in range(n)])
        assert len(saved) == n and all(c.DocumentID == doc for c in saved) and len({c.ID for c in saved}) == n
        st.SaveSummary(doc, Summary(doc, "a summary", ["p1", "p2"]))
        st.SaveEmbeddings([])                                            # empty slice: no-op, no error
        assert st.TopK([doc], case.q, 5) is None                         # nothing stored yet -> nil slice
        st.SaveEmbeddings([Embedding(c.ID, case.mat[i], "text-embedding-3-large") for i, c in enumerate(saved)])
        res = st.TopK([doc], case.q, 5)                                  # top_k default of the handler is 5
        o_rows, o_scores, _ = oracle.topk(case.mat, case.row_doc, case.allowed(), case.q, 5)
        assert [r.Chunk.ID for r in res] == [saved[i].ID for i in o_rows]
        assert np.allclose([r.Score for r in res], o_scores, rtol=REL_TOL)
        assert all(isinstance(r.Score, np.float32) for r in res)         # Score float32 (store.go:52)
        assert res[0].Chunk.Text == f"chunk {o_rows[0]}" and res[0].Chunk.DocumentID == doc
        assert res[0].Summary.Summary == "a summary" and res[0].Summary.KeyPoints == ["p1", "p2"]
        assert st.TopK([other], case.q, 5) is None                       # unknown document: nil, no error
        assert st.TopK([], case.q, 5) is None                            # ANY('{}')
        assert st.TopK([doc], case.q, 0) is None                         # LIMIT 0
        assert len(st.TopK([other, doc, doc], case.q, 20)) == 20         # API cap 20, duplicates harmless
        avg = calculate_avg_similarity(res)
        assert avg == np.float32(oracle.avg_similarity([r.Score for r in res]))
        assert calculate_avg_similarity(None) == 0.0
        with pytest.raises(StoreError) as e:                             # dimension mismatch is an error (-> HTTP 500)
            st.TopK([doc], case.q[:1536], 5)
        assert e.value.code == L.DA_ERR_DIM
        with pytest.raises(StoreError) as e:
            st.TopK([doc], case.q, -1)
        assert e.value.code == L.DA_ERR_INVALID


def test_save_embeddings_upsert_semantics():
    S = oracle.Synth(n_clusters=1)
    dim, n = 1536, 300
    mat = oracle.gen_rows(S, 0, n, dim)
    q = oracle.gen_query(S, 0, 0, dim)
    chunk, docs = ids_for(n, 2)
    with GpuStore(dim=dim, capacity_rows=n) as st:
        with pytest.raises(StoreError) as e:                             # FK: chunk never announced
            st.save_embeddings_raw(chunk[:3], mat[:3])
        assert e.value.code == L.DA_ERR_UNKNOWN_CHUNK
        st.save_chunks_raw(docs[0], chunk[:200])
        st.save_chunks_raw(docs[1], chunk[200:])
        st.save_embeddings_raw(chunk[:200], mat[:200])
        st.save_embeddings_raw(chunk[200:], mat[200:])
        assert st.stats().rows == n
        base_rows = st.search(q, 10, docs[:2])[2][0]
        # overwrite 5 rows with copies of the best row's vector scaled: ON CONFLICT DO UPDATE keeps the row count
        tgt = np.array([7, 250, 3, 299, 100])
        newv = oracle.gen_rows(S, 5000, 5, dim)
        st.save_embeddings_raw(chunk[tgt], newv)
        assert st.stats().rows == n
        mat2 = mat.copy()
        mat2[tgt] = newv
        got_v, got_inv, got_slot, got_ids = st.debug_rows(0, 0, n)
        assert np.array_equal(got_v.view(np.uint32), mat2.view(np.uint32))      # lossless, like vectorToString
        assert np.array_equal(got_ids, chunk)
        assert np.array_equal(got_slot, (np.arange(n) >= 200).astype(np.uint32))
        assert np.allclose(got_inv, 1.0 / np.linalg.norm(mat2.astype(np.float64), axis=1), rtol=1e-6)
        rows = st.search(q, 10, docs[:2])[2][0]
        o_rows, _, _ = oracle.topk(mat2, (np.arange(n) >= 200).astype(np.uint32), None, q, 10)
        assert np.array_equal(rows, o_rows)
        assert not np.array_equal(rows, base_rows) or not np.isin(tgt, base_rows).any()
        # statement-level atomicity: a bad element / repeated id / wrong dim writes nothing
        bad = newv.copy()
        bad[2, 5] = np.nan
        for ids_, vecs_, code in ((chunk[tgt], bad, L.DA_ERR_NONFINITE),
                                  (chunk[[1, 2, 1]], newv[:3], L.DA_ERR_INVALID),
                                  (chunk[:2], newv[:2, :512], L.DA_ERR_DIM)):
            with pytest.raises(StoreError) as e:
                st.save_embeddings_raw(ids_, vecs_)
            assert e.value.code == code
        assert np.array_equal(st.debug_rows(0, 0, n)[0].view(np.uint32), mat2.view(np.uint32))
        # moving a chunk to another document changes which filter finds it
        best = int(rows[0])
        st.save_chunks_raw(docs[1 - int(best >= 200)], chunk[best:best + 1])
        rows_moved = st.search(q, 10, docs[int(best >= 200):int(best >= 200) + 1])[2][0]
        assert best not in rows_moved
    with GpuStore(dim=dim, capacity_rows=4) as st:                       # full store: loud error, no fallback
        st.save_chunks_raw(docs[0], chunk[:5])
        with pytest.raises(StoreError) as e:
            st.save_embeddings_raw(chunk[:5], mat[:5])
        assert e.value.code == L.DA_ERR_CAPACITY
        assert st.stats().rows == 0


def test_large_upsert_streams_in_chunks_and_stays_atomic(monkeypatch):
    """A call larger than the staging buffer is streamed twice (check pass, commit pass); a bad element anywhere
    still leaves the store untouched.  The chunk size is shrunk through the library's test knob."""
    monkeypatch.setenv("DA_UPSERT_CHUNK_ROWS", "37")
    S = oracle.Synth(n_clusters=2)
    dim, n = 256, 500
    mat = oracle.gen_rows(S, 0, n, dim)
    chunk, docs = ids_for(n, 3)
    with GpuStore(dim=dim, capacity_rows=n) as st:
        for d in range(3):
            st.save_chunks_raw(docs[d], chunk[d::3])
        order = np.random.default_rng(1).permutation(n)             # scattered input order: many short runs
        bad = mat[order].copy()
        bad[411, 7] = np.inf
        with pytest.raises(StoreError) as e:
            st.save_embeddings_raw(chunk[order], bad)
        assert e.value.code == L.DA_ERR_NONFINITE and "embedding 411" in e.value.message and "element 7" in e.value.message
        assert st.stats().rows == 0
        st.save_embeddings_raw(chunk[order], mat[order])
        assert st.stats().rows == n
        got_v, got_inv, got_slot, got_ids = st.debug_rows(0, 0, n)
        assert np.array_equal(got_v.view(np.uint32), mat[order].view(np.uint32))
        assert np.array_equal(got_ids, chunk[order])
        assert np.array_equal(got_slot, (order % 3).astype(np.uint32))
        assert np.allclose(got_inv, 1.0 / np.linalg.norm(mat[order].astype(np.float64), axis=1), rtol=1e-6)
        st.save_embeddings_raw(chunk[:200], mat[:200] * np.float32(2.0))     # chunked overwrite in place
        q = oracle.gen_query(S, 0, 0, dim)
        rows = st.search(q, 10, all_docs=True)[2][0]
        ref = mat[order].copy()
        pos = {int(o): i for i, o in enumerate(order)}
        o_rows, _, _ = oracle.topk(ref, None, None, q, 10)
        assert np.array_equal(rows, o_rows)


def test_bulk_load_equals_per_document_saves():
    """Cold start: one call with (chunk id, document id, vector) rows in arbitrary order."""
    case = CASES[1]                                                   # 10 documents x 200 rows, 3072-d
    n, dim = case.mat.shape
    chunk, docs = ids_for(n, case.n_docs)
    order = np.random.default_rng(5).permutation(n)
    with GpuStore(dim=dim, capacity_rows=n) as a, GpuStore(dim=dim, capacity_rows=n) as b:
        load_case(a, case)
        b.bulk_load(chunk[order], docs[np.asarray(case.row_doc)[order]], case.mat[order])
        assert b.stats().rows == n and b.stats().n_docs == case.n_docs
        for filt in (docs[[0, 8]], docs[:case.n_docs], docs[[3]]):
            ra = a.search(case.q, 10, filt)
            rb = b.search(case.q, 10, filt)
            assert np.array_equal(ra[0][0], rb[0][0])                 # same chunks (rows differ: other insertion order)
            assert np.allclose(ra[1][0], rb[1][0], rtol=1e-6)


@pytest.mark.parametrize("dim,row_lo,n", [(3072, 0, 400), (1536, 123400, 257), (100, 5, 9)])
def test_device_generator_is_bit_identical_to_the_oracle(dim, row_lo, n):
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S0, row_lo, n)
        vecs, inv, slot, ids = st.debug_rows(0, 0, n)
        ref = oracle.gen_rows(S0, row_lo, n, dim)
        assert np.array_equal(vecs.view(np.uint32), ref.view(np.uint32))
        assert np.array_equal(ids, oracle.chunk_ids(S0, row_lo, n))
        assert np.array_equal(slot, oracle.row_docs(S0, row_lo, n) - row_lo // S0.chunks_per_doc)
        assert np.allclose(inv, 1.0 / np.linalg.norm(ref.astype(np.float64), axis=1), rtol=1e-6)
        assert st.stats().n_docs == (row_lo + n - 1) // 200 - row_lo // 200 + 1


@pytest.mark.parametrize("batch", [1, 2, 3, 5, 8, 9, 17])
def test_batched_queries_equal_single_queries(batch):
    S = oracle.Synth(n_clusters=4)
    n, dim, k = 3000, 1536, 10
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        docs = oracle.doc_ids(S, 0, 15)
        qs = np.stack([oracle.gen_query(S, i, i % 4, dim) for i in range(batch)])
        ids_b, sc_b, rows_b = st.search(qs, k, docs[[0, 1, 2, 3, 5, 6, 8, 13]])
        mat = oracle.gen_rows(S, 0, n, dim)
        allowed = np.zeros(15, np.uint8)
        allowed[[0, 1, 2, 3, 5, 6, 8, 13]] = 1
        for i in range(batch):
            ids_1, sc_1, rows_1 = st.search(qs[i], k, docs[[0, 1, 2, 3, 5, 6, 8, 13]])
            assert np.array_equal(rows_b[i], rows_1[0]) and np.array_equal(ids_b[i], ids_1[0])
            assert np.allclose(sc_b[i], sc_1[0], rtol=1e-6)
            o_rows, o_sc, _ = oracle.topk(mat, oracle.row_docs(S, 0, n), allowed, qs[i], k)
            assert np.array_equal(rows_b[i], o_rows)
            assert np.allclose(sc_b[i], o_sc, rtol=REL_TOL)


@pytest.mark.parametrize("k", [1, 2, 7, 16, 17, 20, 21, 33])
def test_top_k_on_both_sides_of_the_register_merge(k):
    """k <= 20 (the reference's API range, cmd/query/main.go:23) is merged by warp tournaments in registers (CTA level, last CTA, cross-GPU), larger k by CTA-wide sorts
    (da_topk.cuh): the same searches on both sides of the switch, through the bitmap scan, the range path, the SIMT batch
    (2 and 8 queries per pass) and with more qualifying rows than k in every CTA -- all against the oracle."""
    S = oracle.Synth(n_clusters=2)
    n, dim = 40_000, 256                   # 148 CTAs x 16 warps each see a few rows
    n_docs = n // S.chunks_per_doc
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        mat = oracle.gen_rows(S, 0, n, dim)
        rd = oracle.row_docs(S, 0, n)
        docs = oracle.doc_ids(S, 0, n_docs)
        qs = np.stack([oracle.gen_query(S, 50 + i, i % 2, dim) for i in range(8)])
        for floor in (0.7, -1.0):
            for wanted in (None, list(range(0, n_docs, 3)), [4, 5, 11]):
                allowed = None
                kw = dict(all_docs=True) if wanted is None else dict(doc_ids=docs[wanted])
                if wanted is not None:
                    allowed = np.zeros(n_docs, np.uint8)
                    allowed[wanted] = 1
                for batch in (1, 2, 8):
                    ids, sc, rows = st.search(qs[:batch], k, min_similarity=floor, no_tensor=True, **kw)
                    for i in range(batch):
                        o_rows, o_sc, _ = oracle.topk(mat, rd if wanted is not None else None, allowed, qs[i], k + 8, floor)
                        why = topk_matches_extended(rows[i], sc[i], o_rows, o_sc, k)
                        assert why == "", (k, floor, wanted and len(wanted), batch, i, why)
                        assert np.array_equal(ids[i], oracle.chunk_ids(S, 0, n)[rows[i]])


def test_pgvectors_published_cosine_distance_answers_on_the_device():
    """The eight answers of pgvector's own regression suite for cosine_distance (see tests/test_oracle.py for the provenance),
    through the CUDA path: one stored row (padded to dim 4 with zeros, which changes no cosine), one query, no floor."""
    from test_oracle import PGVECTOR_PUBLISHED
    chunk = oracle.chunk_ids(S0, 0, 1)
    docs = oracle.doc_ids(S0, 0, 1)
    for a, b, want in PGVECTOR_PUBLISHED:
        row = np.zeros((1, 4), np.float32); row[0, :len(a)] = a
        q = np.zeros(4, np.float32); q[:len(b)] = b
        with GpuStore(dim=4, capacity_rows=8) as st:
            st.save_chunks_raw(docs[0], chunk)
            st.save_embeddings_raw(chunk, row)
            ids, sc, rows = st.search(q, 5, all_docs=True, min_similarity=-10.0)
            if np.isnan(want) and abs(a[0]) < 1e30:
                assert len(rows[0]) == 0                 # zero vector: NaN in pgvector, never a match here (deviation 1)
            elif np.isnan(want):
                # [3e38] . [3e38]: pgvector's FLOAT accumulators overflow (inf / inf = NaN).  The device keeps norms in double and
                # answers the true cosine, clamped: deviation 5 of DESIGN.md section 2 (norms outside float range: 19 orders of
                # magnitude from any embedding, which openai.go:146-158 normalises to length 1).
                assert len(rows[0]) == 1 and abs(float(sc[0][0]) - 1.0) <= REL_TOL
            else:
                assert len(rows[0]) == 1 and rows[0][0] == 0
                # (the device multiplies by two rounded reciprocal norms: [1,1].[1,1] comes out as 0.99999994, inside the tolerance)
                assert abs(float(sc[0][0]) - (1.0 - want)) <= REL_TOL * abs(1.0 - want) + ABS_TOL, (a, b, sc[0][0], want)
                assert np.array_equal(ids[0][0], chunk[0])


def test_pgvectors_ivfflat_cosine_regression_on_the_device():
    """test/sql/ivfflat_cosine.sql of pgvector (see tests/test_oracle.py): rows [0,0,0], [1,2,3], [1,1,1]; ORDER BY val <=> '[3,3,3]'
    gives [1,1,1], [1,2,3] -- the zero vector is not in a cosine index.  Exact scan and IVFFlat (lists = 1) on the device."""
    from test_oracle import PGVECTOR_IVFFLAT_COSINE as c
    mat = np.zeros((3, 4), np.float32); mat[:, :3] = c["rows"]
    q = np.zeros(4, np.float32); q[:3] = c["query"]
    chunk = oracle.chunk_ids(S0, 0, 3)
    docs = oracle.doc_ids(S0, 0, 1)
    with GpuStore(dim=4, capacity_rows=8) as st:
        st.save_chunks_raw(docs[0], chunk)
        st.save_embeddings_raw(chunk, mat)
        ids, sc, rows = st.search(q, 10, all_docs=True, min_similarity=-10.0)
        assert [int(r) for r in rows[0]] == c["order"]
        assert abs(float(sc[0][0]) - 1.0) <= REL_TOL and abs(float(sc[0][1]) - 6.0 / np.sqrt(42.0)) <= REL_TOL
        assert np.array_equal(ids[0], chunk[c["order"]])
        st.build_ivf(lists=1, iters=2, seed=1, default_probes=1)
        ids2, sc2, rows2 = st.search(q, 10, all_docs=True, min_similarity=-10.0, ivf=True)
        assert [int(r) for r in rows2[0]] == c["order"] and np.array_equal(sc2[0], sc[0])


@pytest.mark.parametrize("sidecar", [False, True])
def test_a_zero_query_matches_nothing_on_any_path_whatever_the_floor(sidecar):
    """pgvector answers NaN for a zero vector; here such a query (like a zero row) matches nothing (DESIGN.md section 2,
    deviation 1) -- also below a floor of 0, alone or inside a batch, on the SIMT scan, the tensor path and the int8 sidecar."""
    S = oracle.Synth(n_clusters=2)
    n, dim, k = 2000, 64, 7
    with GpuStore(dim=dim, capacity_rows=n, q8_sidecar=sidecar) as st:
        st.generate_synthetic(S, 0, n)
        mat = oracle.gen_rows(S, 0, n, dim)
        qs = np.stack([oracle.gen_query(S, i, i % 2, dim) for i in range(5)])
        qs[2] = 0.0
        for floor in (0.7, 0.5, -1.0):
            for kw in (dict(), dict(no_tensor=True), dict(no_tensor=True, no_q8=True)):
                ids, sc, rows = st.search(qs, k, all_docs=True, min_similarity=floor, **kw)
                assert len(rows[2]) == 0, (floor, kw, rows[2], sc[2])
                for i in (0, 1, 3, 4):
                    o_rows, o_sc, _ = oracle.topk(mat, None, None, qs[i], k + 8, floor)
                    assert topk_matches_extended(rows[i], sc[i], o_rows, o_sc, k) == "", (floor, kw, i)
            alone = st.search(qs[2], k, all_docs=True, min_similarity=floor)
            assert len(alone[2][0]) == 0
            assert len(oracle.topk(mat, None, None, qs[2], k, floor)[0]) == 0


def _cluster_oracle(S, n_rows, dim, q, cluster, k, min_sim=0.7):
    """Oracle top-k over the documents of one cluster (the only rows that can rank: same-cluster cosine ~0.8 against
    ~0 across clusters).  Returns global rows and scores."""
    n_docs = n_rows // S.chunks_per_doc
    cl_docs = list(range(cluster, n_docs, S.n_clusters))
    sub_rows = np.concatenate([np.arange(d * 200, d * 200 + 200) for d in cl_docs])
    sub = np.concatenate([oracle.gen_rows(S, d * 200, 200, dim) for d in cl_docs])
    o_rows, o_sc, _ = oracle.topk(sub, None, None, q, k + 8, min_sim)     # a few places past k: ties across the cut show
    return sub_rows[o_rows], o_sc


def _full_size_check(n_rows, dim, k, n_filter_docs, big_batch=0):
    """Parity at a BASELINE size: the document filter (or the 0.7 floor plus the cluster
    structure) leaves only a few thousand rows that can qualify; the oracle scores exactly those."""
    S = S0
    n_docs = n_rows // S.chunks_per_doc
    with GpuStore(dim=dim, capacity_rows=n_rows) as st:
        st.generate_synthetic(S, 0, n_rows)
        assert st.stats().rows == n_rows
        for qi, cluster in ((0, 3), (1, 1000)):
            q = oracle.gen_query(S, qi, cluster, dim)
            # (i) filtered to n_filter_docs documents, half of them in the query's cluster
            in_cl = [d for d in range(cluster, n_docs, S.n_clusters)][: n_filter_docs // 2]
            others = [d for d in range(7, n_docs, max(1, n_docs // n_filter_docs)) if d % S.n_clusters != cluster]
            chosen = sorted(in_cl + others[: n_filter_docs - len(in_cl)])
            docs = oracle.doc_ids(S, 0, n_docs)[chosen] if n_docs <= 60000 else \
                np.stack([oracle.doc_ids(S, d, 1)[0] for d in chosen])
            ids, sc, rows = st.search(q, k, docs)
            assert st.stats().last_rows_read == len(chosen) * S.chunks_per_doc      # rows of other docs skipped
            sub_rows = np.concatenate([np.arange(d * 200, d * 200 + 200) for d in chosen])
            sub = np.concatenate([oracle.gen_rows(S, d * 200, 200, dim) for d in chosen])
            o_rows, o_sc, _ = oracle.topk(sub, None, None, q, k)
            assert np.array_equal(rows[0], sub_rows[o_rows])
            assert np.allclose(sc[0], o_sc, rtol=REL_TOL)
            assert np.array_equal(ids[0], np.stack([oracle.chunk_ids(S, int(r), 1)[0] for r in rows[0]]))
            # (ii) every document requested: only the query's cluster can clear the floor
            ids_a, sc_a, rows_a = st.search(q, k, all_docs=True)
            assert st.stats().last_rows_read == n_rows
            cl_docs = list(range(cluster, n_docs, S.n_clusters))
            sub_rows = np.concatenate([np.arange(d * 200, d * 200 + 200) for d in cl_docs])
            sub = np.concatenate([oracle.gen_rows(S, d * 200, 200, dim) for d in cl_docs])
            o_rows, o_sc, _ = oracle.topk(sub, None, None, q, k)
            assert np.array_equal(rows_a[0], sub_rows[o_rows])
            assert np.allclose(sc_a[0], o_sc, rtol=REL_TOL)
            # (iii) size-independent properties: idempotence, LIMIT prefix, batch == single
            again = st.search(q, k, all_docs=True)
            assert np.array_equal(again[2][0], rows_a[0]) and np.array_equal(again[1][0], sc_a[0])
            assert np.array_equal(st.search(q, k // 2, all_docs=True)[2][0], rows_a[0][: k // 2])
        qs = np.stack([oracle.gen_query(S, i, (37 * i) % S.n_clusters, dim) for i in range(4)])
        b = st.search(qs, k, all_docs=True)
        for i in range(4):
            one = st.search(qs[i], k, all_docs=True)
            assert np.array_equal(b[2][i], one[2][0]) and np.allclose(b[1][i], one[1][0], rtol=1e-6)
        if big_batch:
            # configs[2] at batch = 64: the tcgen05 candidate path + exact re-score, against the oracle on sampled queries
            qs = np.stack([oracle.gen_query(S, 500 + i, (37 * i) % S.n_clusters, dim) for i in range(big_batch)])
            l0 = st.stats().kernel_launches
            b = st.search(qs, k, all_docs=True)
            assert st.stats().kernel_launches - l0 <= 6                      # one contraction, not 8 SIMT passes
            assert st.stats().tensor_overflows == 0
            for i in range(0, big_batch, max(1, big_batch // 8)):
                o_rows, o_sc = _cluster_oracle(S, n_rows, dim, qs[i], (37 * i) % S.n_clusters, k)
                assert topk_matches_extended(b[2][i], b[1][i], o_rows, o_sc, k) == "", i
                assert np.array_equal(b[0][i], np.stack([oracle.chunk_ids(S, int(r), 1)[0] for r in b[2][i]]))


def test_config2_exact_scan_1m_x_3072_filtered_to_16_docs():
    _full_size_check(1_000_000, 3072, 10, 16)


def test_config3_exact_scan_10m_x_3072_single_gpu_batch_1_and_64():
    _full_size_check(10_000_000, 3072, 10, 16, big_batch=64)


def test_config4_one_shard_4096_queries_x_1p25m_x_1536_top100_no_floor():
    """One GPU's share of BASELINE configs[3] (4 096 questions x 10M x 1536 over 8 GPUs, k = 100, no floor): the
    tensor-core score GEMM fused with the per-query candidate select, then the exact re-score; sampled queries
    against the oracle, every query for size-independent properties."""
    S = S0
    n_rows, dim, k, batch = 1_250_000, 1536, 100, 4096
    with GpuStore(dim=dim, capacity_rows=n_rows) as st:
        st.generate_synthetic(S, 0, n_rows)
        qs = np.stack([oracle.gen_query(S, 9000 + i, (37 * i) % S.n_clusters, dim) for i in range(batch)])
        ids, sc, rows, cnt = st.search_arrays(qs, k, all_docs=True, min_similarity=-np.inf)
        assert st.stats().tensor_overflows == 0
        assert np.all(cnt == k)
        assert np.all(np.diff(sc, axis=1) <= 0)                                  # best first, every query
        assert all(len(np.unique(rows[i])) == k for i in range(0, batch, 97))    # no row twice
        for i in (0, 1, 511, 1024, 2047, 3000, 4000, 4095):
            o_rows, o_sc = _cluster_oracle(S, n_rows, dim, qs[i], (37 * i) % S.n_clusters, k, -np.inf)
            assert o_sc[-1] > 0.5                                                # all k come from the query's own cluster
            assert topk_matches_extended(rows[i], sc[i], o_rows, o_sc, k) == "", i
        one = st.search(qs[1024], k, all_docs=True, min_similarity=-np.inf)      # the SIMT scan answers the same
        assert np.array_equal(one[2][0], rows[1024]) and np.array_equal(one[1][0], sc[1024])


def test_dim_1536_store():
    _full_size_check(400_000, 1536, 20, 8)


@pytest.mark.skipif(device_count() < 2, reason="needs 2 GPUs in one process")
def test_two_shards_in_one_process_equal_one_shard():
    S = oracle.Synth(n_clusters=8)
    n, dim, k = 40_000, 3072, 10
    with GpuStore(dim=dim, capacity_rows=n, devices=(0,)) as one, \
            GpuStore(dim=dim, capacity_rows=n, devices=(0, 1)) as two:
        one.generate_synthetic(S, 0, n)
        two.generate_synthetic(S, 0, n)
        assert list(two.stats().rows_per_device[:2]) == [20_000, 20_000]
        docs = oracle.doc_ids(S, 0, 200)
        qs = np.stack([oracle.gen_query(S, i, i % 8, dim) for i in range(5)])
        for kw in (dict(all_docs=True), dict(doc_ids=docs[::3])):
            a = one.search(qs, k, **kw)
            b = two.search(qs, k, **kw)
            for i in range(5):
                assert np.array_equal(a[2][i], b[2][i]) and np.array_equal(a[0][i], b[0][i])
                assert np.array_equal(a[1][i], b[1][i])


def test_empty_store_answers_every_kind_of_search_with_nothing():
    S = oracle.Synth()
    dim = 1536
    with GpuStore(dim=dim, capacity_rows=64) as st:
        q1 = oracle.gen_query(S, 0, 0, dim)
        q40 = np.stack([oracle.gen_query(S, i, 0, dim) for i in range(40)])
        docs = oracle.doc_ids(S, 0, 3)
        for qs, kw in ((q1, dict(all_docs=True)), (q1, dict(doc_ids=docs)), (q40, dict(all_docs=True)),
                       (q40, dict(all_docs=True, min_similarity=-np.inf)), (q40[:5], dict(doc_ids=[docs] * 5))):
            ids, sc, rows = st.search(qs, 10, **kw)
            assert all(len(r) == 0 for r in rows)
        assert st.TopK([], q1, 5) is None


def test_floor_is_compared_as_the_float8_it_is_in_the_reference():
    """`const minSimilarity = 0.7` reaches Postgres as a float8 and is compared with a float8 expression
    (postgres.go:223,240); 0.7 is not a float32: (float)0.7 = 0.69999999.  A row whose fp32 score is exactly that value is
    OUT in the SQL.  The floor therefore travels as a double and rows qualify iff (double)score >= floor; the legacy
    float field keeps the round-1 behaviour."""
    dim = 4
    below, above = np.float32(0.69999999), np.float32(0.70000005)
    assert float(below) < 0.7 < float(above) and np.nextafter(below, np.float32(1)) == above
    rows = np.zeros((3, dim), np.float32)
    for i, a in enumerate((below, above, np.float32(0.9))):
        rows[i, 0] = a
        rows[i, 1] = np.float32(np.sqrt(1.0 - float(a) ** 2))
    q = np.array([1, 0, 0, 0], np.float32)
    chunk, docs = ids_for(3, 1)
    with GpuStore(dim=dim, capacity_rows=16) as st:                      # store floor: the default 0.7, as a double
        st.save_chunks_raw(docs[0], chunk)
        st.save_embeddings_raw(chunk, rows)
        _, inv, _, _ = st.debug_rows(0, 0, 3)
        ids, sc, got = st.search(q, 5, docs[:1])
        if np.all(inv == np.float32(1.0)):                               # the scores are then exactly the planted values
            assert list(sc[0]) == [np.float32(0.9), above] and list(got[0]) == [2, 1]
            legacy = st.search(q, 5, docs[:1], float32_floor=True)
            assert list(legacy[2][0]) == [2, 1, 0]                       # (float)0.7 lets 0.69999999 in
        o_rows, o_sc, _ = oracle.topk(rows, np.zeros(3, np.uint32), np.ones(1, np.uint8), q, 5, 0.7)
        assert np.array_equal(got[0], o_rows) and np.allclose(sc[0], o_sc, rtol=REL_TOL)
        res = st.TopK([uuid.UUID(bytes=docs[0].tobytes())], q, 5)        # the reference's call, the store's own floor
        assert [r.Chunk.ID.bytes for r in res] == [chunk[r].tobytes() for r in o_rows]
