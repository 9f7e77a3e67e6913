"""GPU suite for the IVFFlat path (the index of internal/store/postgres.go:95-99).  k-means need not
match any CPU run bit for bit; what is checked is structural (lists partition the rows, every row
sits in the list of its nearest centroid), exactness of the probed scan (probing every list equals
the exact scan; probing some lists equals the exact scan restricted to those lists) and recall."""
import numpy as np
import pytest

import oracle
from doc_agents_b200 import GpuStore, StoreError
from doc_agents_b200 import _lib as L


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    """Compile the CUDA library once if it is stale (the GPU box normally receives it pre-built)."""
    return built_lib


pytestmark = pytest.mark.gpu


def _store(n, dim, n_clusters):
    S = oracle.Synth(n_clusters=n_clusters)
    st = GpuStore(dim=dim, capacity_rows=n + 1000)
    st.generate_synthetic(S, 0, n)
    return S, st


def test_ivf_search_before_build_is_an_error():
    S, st = _store(2000, 256, 4)
    with st:
        with pytest.raises(StoreError) as e:
            st.search(oracle.gen_query(S, 0, 0, 256), 5, all_docs=True, ivf=True)
        assert e.value.code == L.DA_ERR_STATE
        with pytest.raises(StoreError):
            st.build_ivf(lists=0)
        with pytest.raises(StoreError):
            st.build_ivf(lists=5000)             # more lists than rows


@pytest.mark.parametrize("n,dim,lists,clusters", [(6000, 256, 16, 8), (20000, 1536, 64, 32), (3000, 100, 7, 5)])
def test_ivf_structure_and_exactness(n, dim, lists, clusters):
    S, st = _store(n, dim, clusters)
    with st:
        st.build_ivf(lists=lists, iters=8, seed=3)
        assert st.stats().ivf_lists == lists
        off, rows, cent = st.debug_ivf()
        assert off[0] == 0 and off[-1] == n and np.all(np.diff(off.astype(np.int64)) >= 0)
        assert np.array_equal(np.sort(rows), np.arange(n, dtype=np.uint32))          # a partition of the rows
        assert np.allclose(np.linalg.norm(cent.astype(np.float64), axis=1), 1.0, atol=1e-4)   # spherical
        mat = oracle.gen_rows(S, 0, n, dim)
        dots = mat.astype(np.float64) @ cent.astype(np.float64).T
        list_of_row = np.empty(n, dtype=np.int64)
        for l in range(lists):
            list_of_row[rows[off[l]:off[l + 1]]] = l
        best = dots.max(axis=1)
        got = dots[np.arange(n), list_of_row]
        # tensor-core (TF32) assignment may pick a centroid that is nearest only up to TF32 rounding
        assert np.all(best - got <= 2e-3), "a row is not in the list of a nearest centroid"
        assert np.mean(best - got <= 1e-6) > 0.99

        docs = oracle.doc_ids(S, 0, n // 200)
        rd = oracle.row_docs(S, 0, n)
        for qi in range(6):
            q = oracle.gen_query(S, qi, qi % clusters, dim)
            exact = st.search(q, 10, all_docs=True)
            # probing every list is the exact scan
            full = st.search(q, 10, all_docs=True, ivf=True, probes=lists)
            assert np.array_equal(full[2][0], exact[2][0]) and np.allclose(full[1][0], exact[1][0], rtol=1e-6)
            assert np.array_equal(full[0][0], exact[0][0])
            # probing p lists equals the oracle restricted to the rows of the p nearest lists
            for probes in (1, 3):
                qd = cent.astype(np.float64) @ q.astype(np.float64)
                order = np.lexsort((np.arange(lists), -qd))[:probes]
                if probes < lists and abs(qd[np.argsort(-qd)[probes - 1]] - qd[np.argsort(-qd)[probes]]) < 1e-5:
                    continue                                                          # list ranking itself ties
                member = np.isin(list_of_row, order)
                got = st.search(q, 10, all_docs=True, ivf=True, probes=probes)
                o_rows, o_sc, _ = oracle.topk(mat[member], None, None, q, 10)
                assert np.array_equal(got[2][0], np.nonzero(member)[0][o_rows])
                assert np.allclose(got[1][0], o_sc, rtol=1e-5)
                assert st.stats().last_rows_read == int(member.sum())
            # the document filter is applied to the probed rows (pgvector filters after the index scan)
            pick = list(range(0, n // 200, 2))
            allowed = np.zeros(n // 200, np.uint8)
            allowed[pick] = 1
            f_ivf = st.search(q, 10, docs[pick], ivf=True, probes=lists)
            o_rows, o_sc, _ = oracle.topk(mat, rd, allowed, q, 10)
            assert np.array_equal(f_ivf[2][0], o_rows)
        # batch of queries through the IVF path
        qs = np.stack([oracle.gen_query(S, 20 + i, i % clusters, dim) for i in range(5)])
        b = st.search(qs, 10, all_docs=True, ivf=True, probes=2)
        for i in range(5):
            one = st.search(qs[i], 10, all_docs=True, ivf=True, probes=2)
            assert np.array_equal(b[2][i], one[2][0])


def test_ivf_recall_and_rows_read():
    n, dim, lists, clusters = 100_000, 1536, 256, 128
    S, st = _store(n, dim, clusters)
    with st:
        st.build_ivf(lists=lists, iters=10, seed=1, default_probes=8)
        hits = total = 0
        read = []
        for qi in range(40):
            q = oracle.gen_query(S, qi, (qi * 7) % clusters, dim)
            exact = st.search(q, 10, all_docs=True)[2][0]
            approx = st.search(q, 10, all_docs=True, ivf=True)[2][0]        # default probes = 8
            read.append(st.stats().last_rows_read)
            hits += len(np.intersect1d(exact, approx))
            total += len(exact)
        recall = hits / max(total, 1)
        assert recall >= 0.95, recall
        assert np.mean(read) < 0.2 * n                                       # far fewer rows than the exact scan


def test_ivf_stays_current_after_upserts():
    n, dim = 4000, 256
    S, st = _store(n, dim, 4)
    with st:
        st.build_ivf(lists=8, iters=5)
        q = oracle.gen_query(S, 1, 1, dim)
        chunk = oracle.chunk_ids(oracle.Synth(seed_ids=99), 0, 3)
        doc = oracle.doc_ids(oracle.Synth(seed_ids=99), 0, 1)
        st.save_chunks_raw(doc[0], chunk)
        # three copies of the query itself: similarity 1.0, must come back first through the index
        st.save_embeddings_raw(chunk, np.stack([q, q, q]))
        off, rows, _ = st.debug_ivf()
        assert off[-1] == n + 3 and np.array_equal(np.sort(rows), np.arange(n + 3, dtype=np.uint32))
        got = st.search(q, 5, all_docs=True, ivf=True, probes=1)
        assert list(got[2][0][:3]) == [n, n + 1, n + 2] and np.allclose(got[1][0][:3], 1.0, atol=1e-6)
        assert np.array_equal(got[0][0][:3], chunk)


@pytest.mark.parametrize("n,dim,lists", [(5000, 3072, 300), (4100, 1536, 64), (777, 100, 9)])
def test_tensor_core_assignment_agrees_with_fp32_simt(n, dim, lists):
    """The tcgen05 TF32 GEMM+argmax against the fp32 SIMT GEMM+argmax on the same centroids
    (iters=0 -> k-means keeps the seeded centroids, so both builds assign against identical ones)."""
    S = oracle.Synth(n_clusters=32)
    results = {}
    for simt in (True, False):
        with GpuStore(dim=dim, capacity_rows=n) as st:
            st.generate_synthetic(S, 0, n)
            st.build_ivf(lists=lists, iters=1, seed=5, simt_assign=simt)
            off, rows, cent = st.debug_ivf()
            lor = np.empty(n, dtype=np.int64)
            for l in range(lists):
                lor[rows[off[l]:off[l + 1]]] = l
            results[simt] = (lor, cent)
    (l_simt, c_simt), (l_tc, c_tc) = results[True], results[False]
    mat = oracle.gen_rows(S, 0, n, dim).astype(np.float64)
    # after one k-means step the two centroid sets differ only where TF32 flipped a near-tie
    d_tc = mat @ c_tc.astype(np.float64).T
    best = d_tc.max(axis=1)
    got = d_tc[np.arange(n), l_tc]
    assert np.all(best - got <= 2e-3)
    assert np.mean(best - got <= 1e-6) > 0.99
    assert np.mean(l_simt == l_tc) > 0.97
