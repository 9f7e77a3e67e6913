"""GPU suite for the split-plane resident format (DA_STORE_SPLIT16): lossless storage,
the scan reads half the bytes, answers are bit-identical to the default fp32 layout and match the oracle."""
import numpy as np
import pytest

import oracle
from cases import ABS_TOL, REL_TOL, build_cases, check_against_oracle
from doc_agents_b200 import GpuStore, StoreError
from doc_agents_b200 import _lib as L
from gpu_util import doc_filter, load_case


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    return built_lib


pytestmark = pytest.mark.gpu

CASES = [c for c in build_cases() if c.mat.shape[1] % 16 == 0]


@pytest.mark.parametrize("case", CASES, ids=lambda c: c.name)
def test_case_matches_oracle_and_fp32_layout(case):
    n, dim = case.mat.shape
    with GpuStore(dim=dim, capacity_rows=n + 8, split16=True) as sp, GpuStore(dim=dim, capacity_rows=n + 8) as fp:
        chunk, docs = load_case(sp, case)
        load_case(fp, case)
        a = sp.search(case.q, case.k, doc_filter(case, docs), min_similarity=case.min_sim)
        b = fp.search(case.q, case.k, doc_filter(case, docs), min_similarity=case.min_sim)
        check_against_oracle(case, a[2][0], a[1][0])
        assert np.array_equal(a[2][0], b[2][0])
        assert np.array_equal(a[1][0].view(np.uint32), b[1][0].view(np.uint32))      # bit-identical scores
        assert np.array_equal(a[0][0], b[0][0])
        got_v, got_inv, _, got_ids = sp.debug_rows(0, 0, n)
        assert np.array_equal(got_v.view(np.uint32), case.mat.view(np.uint32))       # lossless
        assert np.array_equal(got_ids, chunk)


def test_split_store_equals_fp32_store_on_synthetic_corpus():
    S = oracle.Synth(n_clusters=16)
    n, dim = 60_000, 3072
    with GpuStore(dim=dim, capacity_rows=n, split16=True) as sp, GpuStore(dim=dim, capacity_rows=n) as fp:
        sp.generate_synthetic(S, 0, n)
        fp.generate_synthetic(S, 0, n)
        v_sp, inv_sp, slot_sp, ids_sp = sp.debug_rows(0, 1000, 300)
        v_fp, inv_fp, slot_fp, ids_fp = fp.debug_rows(0, 1000, 300)
        assert np.array_equal(v_sp.view(np.uint32), v_fp.view(np.uint32)) and np.array_equal(inv_sp, inv_fp)
        assert np.array_equal(slot_sp, slot_fp) and np.array_equal(ids_sp, ids_fp)
        docs = oracle.doc_ids(S, 0, n // 200)
        qs = np.stack([oracle.gen_query(S, i, i % 16, dim) for i in range(20)])
        for kw in (dict(all_docs=True), dict(doc_ids=docs[::3]), dict(all_docs=True, min_similarity=-np.inf),
                   dict(doc_ids=docs[5:9], min_similarity=0.0), dict(all_docs=True, read_all_rows=True)):
            for batch, k in ((1, 10), (5, 3), (9, 128), (20, 10)):
                a = sp.search(qs[:batch], k, **kw)
                b = fp.search(qs[:batch], k, no_tensor=True, **kw)
                for i in range(batch):
                    assert np.array_equal(a[2][i], b[2][i]), (kw, batch, k, i)
                    assert np.array_equal(a[1][i].view(np.uint32), b[1][i].view(np.uint32))
                    assert np.array_equal(a[0][i], b[0][i])
        # per-query document sets
        sets = [docs[i::5] for i in range(7)]
        a = sp.search(qs[:7], 10, sets)
        b = fp.search(qs[:7], 10, sets)
        for i in range(7):
            assert np.array_equal(a[2][i], b[2][i])
        # the scan read the upper plane only: half the vector bytes of the fp32 layout
        sp.search(qs[0], 10, all_docs=True)
        fp.search(qs[0], 10, all_docs=True)
        assert sp.stats().last_bytes_algorithmic < 0.52 * fp.stats().last_bytes_algorithmic
        # the paths that need fp32 rows say so
        with pytest.raises(StoreError) as e:
            sp.build_ivf(lists=16)
        assert e.value.code == L.DA_ERR_STATE
    with pytest.raises(StoreError) as e:
        GpuStore(dim=100, capacity_rows=10, split16=True)
    assert e.value.code == L.DA_ERR_DIM


@pytest.mark.parametrize("batch,k,floor", [(16, 10, 0.7), (64, 10, 0.7), (150, 100, -np.inf), (40, 128, 0.0)])
def test_split_store_bf16_tensor_batches_equal_fp32_store(batch, k, floor):
    """Batches of 16+ queries on a split-plane store: tcgen05 kind::f16 over the upper planes (bf16) generates the
    candidates, the exact re-score from both planes makes the answer bit-identical to the fp32 layout."""
    S = oracle.Synth(n_clusters=16)
    n, dim = 50_000, 1536
    with GpuStore(dim=dim, capacity_rows=n, split16=True) as sp, GpuStore(dim=dim, capacity_rows=n) as fp:
        sp.generate_synthetic(S, 0, n)
        fp.generate_synthetic(S, 0, n)
        docs = oracle.doc_ids(S, 0, n // 200)
        qs = np.stack([oracle.gen_query(S, i, i % 16, dim) for i in range(batch)])
        for kw in (dict(all_docs=True), dict(doc_ids=docs[::2]), dict(doc_ids=[docs[i % 3::3] for i in range(batch)])):
            a = sp.search(qs, k, min_similarity=floor, **kw)
            assert sp.stats().last_rows_read == n                        # the contraction path ran
            b = fp.search(qs, k, min_similarity=floor, no_tensor=True, **kw)
            for i in range(batch):
                assert np.array_equal(a[2][i], b[2][i]), (kw.keys(), i)
                assert np.array_equal(a[1][i].view(np.uint32), b[1][i].view(np.uint32))
                assert np.array_equal(a[0][i], b[0][i])


def test_split_store_full_size_10m():
    S = oracle.Synth()
    n, dim, k = 10_000_000, 3072, 10
    with GpuStore(dim=dim, capacity_rows=n, split16=True) as st:
        st.generate_synthetic(S, 0, n)
        n_docs = n // 200
        for qi, cluster in ((0, 3), (1, 1000)):
            q = oracle.gen_query(S, qi, cluster, dim)
            ids, sc, rows = st.search(q, k, all_docs=True)
            cl_docs = list(range(cluster, n_docs, S.n_clusters))
            sub_rows = np.concatenate([np.arange(d * 200, d * 200 + 200) for d in cl_docs])
            sub = np.concatenate([oracle.gen_rows(S, d * 200, 200, dim) for d in cl_docs])
            o_rows, o_sc, _ = oracle.topk(sub, None, None, q, k)
            assert np.array_equal(rows[0], sub_rows[o_rows])
            assert np.allclose(sc[0], o_sc, rtol=REL_TOL)
            assert st.stats().last_rows_read == n
