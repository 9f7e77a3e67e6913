"""GPU suite for the int8 sidecar (DA_STORE_Q8_SIDECAR): the scan reads one byte per element
to find the candidates, the fp32 re-score makes the answers bit-identical to the default store and the oracle."""
import numpy as np
import pytest

import oracle
from cases import REL_TOL, build_cases, check_against_oracle
from doc_agents_b200 import GpuStore, StoreError
from doc_agents_b200 import _lib as L
from gpu_util import doc_filter, load_case


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    return built_lib


pytestmark = pytest.mark.gpu

CASES = [c for c in build_cases() if c.mat.shape[1] % 32 == 0]


def same(a, b, i, ctx=None):
    assert np.array_equal(a[2][i], b[2][i]), ctx
    assert np.array_equal(a[1][i].view(np.uint32), b[1][i].view(np.uint32)), ctx      # bit-identical scores
    assert np.array_equal(a[0][i], b[0][i]), ctx


@pytest.mark.parametrize("case", CASES, ids=lambda c: c.name)
def test_case_matches_oracle_and_fp32_scan(case):
    n, dim = case.mat.shape
    with GpuStore(dim=dim, capacity_rows=n + 8, q8_sidecar=True) as st:
        _, docs = load_case(st, case)
        a = st.search(case.q, case.k, doc_filter(case, docs), min_similarity=case.min_sim)
        b = st.search(case.q, case.k, doc_filter(case, docs), min_similarity=case.min_sim, no_q8=True)
        check_against_oracle(case, a[2][0], a[1][0])
        same(a, b, 0)


def test_sidecar_equals_fp32_scan_on_synthetic_corpus():
    S = oracle.Synth(n_clusters=16)
    n, dim = 60_000, 3072
    with GpuStore(dim=dim, capacity_rows=n, q8_sidecar=True) as st, GpuStore(dim=dim, capacity_rows=n) as fp:
        st.generate_synthetic(S, 0, n)
        fp.generate_synthetic(S, 0, n)
        docs = oracle.doc_ids(S, 0, n // 200)
        qs = np.stack([oracle.gen_query(S, i, i % 16, dim) for i in range(20)])
        for kw in (dict(all_docs=True), dict(doc_ids=docs[::3]), dict(all_docs=True, min_similarity=-np.inf),
                   dict(doc_ids=docs[5:9], min_similarity=0.0), dict(all_docs=True, read_all_rows=True)):
            for batch, k in ((1, 10), (2, 1), (5, 3), (9, 128), (15, 10)):
                a = st.search(qs[:batch], k, **kw)
                b = fp.search(qs[:batch], k, **kw)
                for i in range(batch):
                    same(a, b, i, (kw, batch, k, i))
        # per-query document sets
        sets = [docs[i::5] for i in range(7)]
        a = st.search(qs[:7], 10, sets)
        b = fp.search(qs[:7], 10, sets)
        for i in range(7):
            same(a, b, i)
        # the scan read one byte per element: about a quarter of the fp32 scan's bytes
        st.search(qs[0], 10, all_docs=True)
        fp.search(qs[0], 10, all_docs=True)
        assert st.stats().last_bytes_algorithmic < 0.27 * fp.stats().last_bytes_algorithmic
        assert st.stats().last_rows_read == n
        # batches of 16+ still take the tensor path over the fp32 rows, IVFFlat still works
        a = st.search(qs[:20], 10, all_docs=True)
        b = fp.search(qs[:20], 10, all_docs=True, no_tensor=True)
        for i in range(20):
            same(a, b, i)
        st.build_ivf(lists=16, iters=3)
        a = st.search(qs[0], 10, all_docs=True, ivf=True, probes=16)
        same(a, st.search(qs[0], 10, all_docs=True), 0)
    with pytest.raises(StoreError) as e:
        GpuStore(dim=100, capacity_rows=10, q8_sidecar=True)
    assert e.value.code == L.DA_ERR_DIM


def test_sidecar_follows_upserts_and_survives_hostile_rows():
    """Rows rewritten by SaveEmbeddings are re-quantised; rows that quantise badly (one huge element, tiny values,
    zero rows) only widen their own interval; a mass of near-ties falls back to the fp32 scan."""
    rng = np.random.default_rng(7)
    n, dim, k = 3000, 256, 10
    mat = rng.standard_normal((n, dim)).astype(np.float32)
    mat[5] = 0.0                                             # zero row: never returned
    mat[6] *= 1e-20
    mat[7, 3] = 1e6                                          # everything else quantises to 0
    mat[100:400] = mat[99] * (1 + 1e-7 * rng.standard_normal((300, 1)).astype(np.float32))    # near-ties
    q = (mat[99] + 0.01 * rng.standard_normal(dim)).astype(np.float32)
    chunk = rng.integers(0, 256, (n, 16), dtype=np.uint8)
    doc = rng.integers(0, 256, (1, 16), dtype=np.uint8)
    with GpuStore(dim=dim, capacity_rows=n, q8_sidecar=True) as st:
        st.save_chunks_raw(doc[0], chunk)
        st.save_embeddings_raw(chunk, mat)
        for floor in (-np.inf, 0.0, 0.7):
            a = st.search(q, k, all_docs=True, min_similarity=floor)
            b = st.search(q, k, all_docs=True, min_similarity=floor, no_q8=True)
            same(a, b, 0, floor)
            o_rows, o_sc, _ = oracle.topk(mat, None, None, q, k, min_sim=floor)
            assert np.allclose(a[1][0], o_sc, rtol=REL_TOL, atol=5e-7)
        # overwrite the best rows with something unrelated: the sidecar must forget them
        mat2 = mat.copy()
        mat2[99:400] = rng.standard_normal((301, dim)).astype(np.float32)
        st.save_embeddings_raw(chunk[99:400], mat2[99:400])
        a = st.search(q, k, all_docs=True, min_similarity=-np.inf)
        b = st.search(q, k, all_docs=True, min_similarity=-np.inf, no_q8=True)
        same(a, b, 0)
        o_rows, o_sc, _ = oracle.topk(mat2, None, None, q, k, min_sim=-np.inf)
        assert np.array_equal(a[2][0], o_rows)


def test_sidecar_full_size_10m():
    S = oracle.Synth()
    n, dim, k = 10_000_000, 3072, 10
    with GpuStore(dim=dim, capacity_rows=n, q8_sidecar=True) as st:
        st.generate_synthetic(S, 0, n)
        n_docs = n // 200
        for qi, cluster in ((0, 3), (1, 1000)):
            q = oracle.gen_query(S, qi, cluster, dim)
            a = st.search(q, k, all_docs=True)
            assert st.stats().last_rows_read == n
            b = st.search(q, k, all_docs=True, no_q8=True)
            same(a, b, 0)
            cl_docs = list(range(cluster, n_docs, S.n_clusters))
            sub_rows = np.concatenate([np.arange(d * 200, d * 200 + 200) for d in cl_docs])
            sub = np.concatenate([oracle.gen_rows(S, d * 200, 200, dim) for d in cl_docs])
            o_rows, o_sc, _ = oracle.topk(sub, None, None, q, k)
            assert np.array_equal(a[2][0], sub_rows[o_rows])
            assert np.allclose(a[1][0], o_sc, rtol=REL_TOL)


def test_sidecar_overflow_falls_back_to_the_fp32_scan():
    """More rows inside the interval slack than the survivor list holds (here: 20 000 copies of one row): the
    overflow flag sends the search to the fp32 scan, and the answer is still the fp32 scan's."""
    rng = np.random.default_rng(11)
    n, dim, k = 20_000, 64, 10
    base = rng.standard_normal(dim).astype(np.float32)
    mat = np.tile(base, (n, 1))
    mat[::7] *= np.float32(1.5)                               # same direction, other norms
    chunk = rng.integers(0, 256, (n, 16), dtype=np.uint8)
    chunk[:, :4] = np.arange(n, dtype=np.uint32).view(np.uint8).reshape(n, 4)      # unique ids
    doc = rng.integers(0, 256, 16, dtype=np.uint8)
    with GpuStore(dim=dim, capacity_rows=n, q8_sidecar=True) as st:
        st.save_chunks_raw(doc, chunk)
        st.save_embeddings_raw(chunk, mat)
        before = st.stats().tensor_overflows
        a = st.search(base, k, all_docs=True, min_similarity=-np.inf)
        assert st.stats().tensor_overflows == before + 1
        b = st.search(base, k, all_docs=True, min_similarity=-np.inf, no_q8=True)
        same(a, b, 0)
        assert len(a[2][0]) == k


@pytest.mark.parametrize("batch,k,floor", [(3, 5, 0.7), (16, 10, 0.7), (64, 10, 0.7), (300, 128, 0.55), (150, 100, -np.inf),
                                           (300, 128, 0.0)])
def test_sidecar_int8_tensor_batches_equal_fp32_scan(batch, k, floor):
    """Batches on a sidecar store: tcgen05 kind::i8 over the int8 rows and queries generates the candidates (slack =
    the rows' largest error bound combined with each query's own), the exact re-score from the fp32 rows makes the
    answer bit-identical to the fp32 scan.  Searches without a floor of at least 0.5 keep the TF32 contraction."""
    S = oracle.Synth(n_clusters=16)
    n, dim = 50_000, 1536
    with GpuStore(dim=dim, capacity_rows=n, q8_sidecar=True) as st, GpuStore(dim=dim, capacity_rows=n) as fp:
        st.generate_synthetic(S, 0, n)
        fp.generate_synthetic(S, 0, n)
        docs = oracle.doc_ids(S, 0, n // 200)
        qs = np.stack([oracle.gen_query(S, i, i % 16, dim) for i in range(batch)])
        for kw in (dict(all_docs=True), dict(doc_ids=docs[::2]), dict(doc_ids=[docs[i % 3::3] for i in range(batch)])):
            a = st.search(qs, k, min_similarity=floor, **kw)
            assert st.stats().last_rows_read == n                        # the contraction path ran
            if floor >= 0.5:
                assert st.stats().last_bytes_algorithmic < n * (dim + 9)  # ... over the int8 plane
            else:                                                         # no floor to bound the int8 slack: TF32 over fp32 rows
                assert st.stats().last_bytes_algorithmic >= n * dim * 4
            b = fp.search(qs, k, min_similarity=floor, no_tensor=True, **kw)
            for i in range(batch):
                same(a, b, i, (list(kw.keys()), i))
        # the same store through the TF32 path over its fp32 rows
        c = st.search(qs, k, min_similarity=floor, all_docs=True, no_q8=True)
        d = st.search(qs, k, min_similarity=floor, all_docs=True)
        for i in range(batch):
            same(c, d, i)


def test_one_badly_quantised_row_keeps_batches_on_tf32():
    """The int8 contraction has one slack for all rows; a row whose own bound is large (one huge element) must not
    widen it for every batch: such a store answers batches through TF32, single searches still through the int8 scan."""
    S = oracle.Synth(n_clusters=16)
    n, dim, batch, k = 20_000, 1024, 32, 10
    mat = np.array(oracle.gen_rows(S, 0, n, dim))
    mat[123, 7] = np.float32(2.0)        # ~64x a typical element: the others keep only ~2 levels, e_x ~ 0.065
    chunk = oracle.chunk_ids(S, 0, n)
    docs = oracle.doc_ids(S, 0, n // 200)
    qs = np.stack([oracle.gen_query(S, i, i % 16, dim) for i in range(batch)])
    with GpuStore(dim=dim, capacity_rows=n, q8_sidecar=True) as st:
        for d in range(n // 200):
            st.save_chunks_raw(docs[d], chunk[d * 200:(d + 1) * 200])
        st.save_embeddings_raw(chunk, mat)
        a = st.search(qs, k, all_docs=True)
        assert st.stats().last_bytes_algorithmic >= n * dim * 4          # fp32 rows, not the int8 plane
        assert st.stats().tensor_overflows == 0
        b = st.search(qs, k, all_docs=True, no_tensor=True, no_q8=True)
        for i in range(batch):
            same(a, b, i)
        st.search(qs[0], k, all_docs=True)
        assert st.stats().last_bytes_algorithmic < n * (dim + 20)
