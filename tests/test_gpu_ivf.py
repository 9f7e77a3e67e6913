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
        # (sibling lists made by splitting a heavy list sit closer together than TF32 resolves for a few rows in a hundred)
        assert np.mean(best - got <= 1e-6) > 0.9

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
            st.build_ivf(lists=lists, iters=1, seed=5, simt_assign=simt, split=False)   # Lloyd step only: same inputs both ways
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


def _list_of_row(off, rows, n, lists):
    lor = np.empty(n, dtype=np.int64)
    for l in range(lists):
        lor[rows[off[l]:off[l + 1]]] = l
    return lor


@pytest.mark.parametrize("n,dim,lists,clusters,batch,probes", [(20000, 1536, 64, 32, 8, 3), (30000, 768, 300, 16, 40, 5),
                                                             (6000, 100, 37, 8, 130, 37), (12000, 256, 16, 4, 9, 1)])
def test_batched_ivf_ranks_centroids_on_tensor_cores(n, dim, lists, clusters, batch, probes):
    """Batches of 8+ rank the centroids with ONE tcgen05 GEMM and scan with ONE probe-mode launch.  For every query:
    (i) the answer equals the oracle restricted to the lists the batch actually probed; (ii) those lists are the
    fp32-nearest ones up to TF32 rounding of the ranking; (iii) the batch agrees with single searches wherever the
    fp32 ranking has no near-tie at the cut."""
    S, st = _store(n, dim, clusters)
    with st:
        st.build_ivf(lists=lists, iters=6, seed=2)
        off, rows, cent = st.debug_ivf()
        lor = _list_of_row(off, rows, n, lists)
        mat = oracle.gen_rows(S, 0, n, dim)
        qs = np.stack([oracle.gen_query(S, 50 + i, i % clusters, dim) for i in range(batch)])
        l0 = st.stats().kernel_launches
        got = st.search(qs, 10, all_docs=True, ivf=True, probes=probes)
        assert st.stats().kernel_launches - l0 == 3              # score GEMM, selection, ONE scan (which computes 1/|q| itself)
        probed = st.debug_probes(batch, probes)
        read_by_batch = st.stats().last_rows_read
        rows_read = 0
        for i in range(batch):
            mine = probed[i][probed[i] >= 0]
            assert len(mine) == min(probes, lists) and len(set(mine.tolist())) == len(mine)
            member = np.isin(lor, mine)
            rows_read += int(member.sum())
            o_rows, o_sc, _ = oracle.topk(mat[member], None, None, qs[i], 10)
            assert np.array_equal(got[2][i], np.nonzero(member)[0][o_rows]), i
            assert np.allclose(got[1][i], o_sc, rtol=1e-5)
            qd = (cent.astype(np.float64) @ qs[i].astype(np.float64)) / np.linalg.norm(qs[i].astype(np.float64))   # lists rank by inner product
            kth = np.sort(qd)[::-1][len(mine) - 1]
            assert np.all(qd[mine] >= kth - 2e-3), "a probed list is not among the nearest up to TF32 rounding"
            clear = len(mine) == lists or (kth - np.sort(qd)[::-1][len(mine)]) > 4e-3
            if clear:
                one = st.search(qs[i], 10, all_docs=True, ivf=True, probes=probes)
                one_probed = st.debug_probes(1, probes)[0]
                assert set(one_probed.tolist()) == set(mine.tolist()), (i, sorted(one_probed.tolist()), sorted(mine.tolist()),
                                                                        np.sort(qd)[::-1][:probes + 2].tolist(), qd[one_probed].tolist())
                assert np.array_equal(one[2][0], got[2][i]) and np.array_equal(one[1][0], got[1][i])
        assert read_by_batch == rows_read
        # the same batch kept off the tensor cores: fp32 centroid scores, 8 queries per pass, the same selection, ONE probe scan
        simt = st.search(qs, 10, all_docs=True, ivf=True, probes=probes, no_tensor=True)
        probed_simt = st.debug_probes(batch, probes)
        for i in range(batch):
            qd = cent.astype(np.float64) @ qs[i].astype(np.float64)
            srt = np.sort(qd)[::-1]
            if probes >= lists or srt[probes - 1] - srt[probes] > 1e-5:
                want = set(np.argsort(-qd)[:min(probes, lists)].tolist())
                assert set(probed_simt[i][probed_simt[i] >= 0].tolist()) == want
            if set(probed_simt[i].tolist()) == set(probed[i].tolist()):
                assert np.array_equal(simt[2][i], got[2][i]) and np.array_equal(simt[1][i], got[1][i])


def test_batched_ivf_with_per_query_document_sets():
    n, dim, lists, clusters, batch = 16000, 512, 32, 8, 12
    S, st = _store(n, dim, clusters)
    with st:
        st.build_ivf(lists=lists, iters=5, seed=4)
        off, rows, _ = st.debug_ivf()
        lor = _list_of_row(off, rows, n, lists)
        mat = oracle.gen_rows(S, 0, n, dim)
        rd = oracle.row_docs(S, 0, n)
        docs = oracle.doc_ids(S, 0, n // 200)
        qs = np.stack([oracle.gen_query(S, 7 + i, i % clusters, dim) for i in range(batch)])
        sets = [list(range(i % 3, n // 200, 2 + i % 3)) for i in range(batch)]
        got = st.search(qs, 10, [docs[s] for s in sets], ivf=True, probes=4)
        probed = st.debug_probes(batch, 4)
        for i in range(batch):
            allowed = np.zeros(n // 200, np.uint8)
            allowed[sets[i]] = 1
            member = np.isin(lor, probed[i])
            o_rows, o_sc, _ = oracle.topk(mat[member], rd[member], allowed, qs[i], 10)
            assert np.array_equal(got[2][i], np.nonzero(member)[0][o_rows]), i
            assert np.allclose(got[1][i], o_sc, rtol=1e-5)


def test_heavy_lists_split_into_empty_ones():
    """More lists than the corpus has clusters: plain spherical k-means leaves whole clusters in single lists and the
    other lists empty (the larger sibling always wins); the balancing rounds split the heavy lists."""
    n, dim, lists, clusters = 64000, 512, 128, 16           # 4 000 rows per cluster, 8 lists per cluster wanted
    S, st = _store(n, dim, clusters)
    with st:
        st.build_ivf(lists=lists, iters=8, seed=9, split=False)
        off_plain, _, _ = st.debug_ivf()
        sizes_plain = np.diff(off_plain.astype(np.int64))
        st.build_ivf(lists=lists, iters=8, seed=9)
        off, rows, cent = st.debug_ivf()
        sizes = np.diff(off.astype(np.int64))
        assert np.array_equal(np.sort(rows), np.arange(n, dtype=np.uint32))
        assert (sizes == 0).sum() <= lists // 20, (int((sizes == 0).sum()), int((sizes_plain == 0).sum()))
        assert sizes.max() <= 4 * n / lists, (int(sizes.max()), int(sizes_plain.max()))
        assert (sizes == 0).sum() <= (sizes_plain == 0).sum() and sizes.max() <= sizes_plain.max()
        # the index still answers like the exact scan when every list is probed, and recall with a few probes is whole
        lor = _list_of_row(off, rows, n, lists)
        hits = 0
        for qi in range(20):
            q = oracle.gen_query(S, qi, qi % clusters, dim)
            exact = st.search(q, 10, all_docs=True)
            full = st.search(q, 10, all_docs=True, ivf=True, probes=lists)
            assert np.array_equal(full[2][0], exact[2][0])
            some = st.search(q, 10, all_docs=True, ivf=True, probes=16)
            hits += len(np.intersect1d(some[2][0], exact[2][0]))
        assert hits >= 0.95 * 200, hits


def test_two_level_synthetic_corpus_equals_the_oracle_bit_for_bit():
    S = oracle.Synth(n_clusters=64, cmul=1, nmul=1, smul=1, super_group=8)
    n, dim = 3000, 384
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 400, n)
        vecs, _, _, _ = st.debug_rows(0, 0, n)
        ref = oracle.gen_rows(S, 400, n, dim)
        assert np.array_equal(vecs.view(np.uint32), ref.view(np.uint32))
    from doc_agents_b200 import synth as psynth
    PS = psynth.Synth(n_clusters=64, cmul=1, nmul=1, smul=1, super_group=8)
    for qi, cl in ((0, 0), (5, 63)):
        assert np.array_equal(psynth.gen_query(PS, qi, cl, dim).view(np.uint32), oracle.gen_query(S, qi, cl, dim).view(np.uint32))
    # clusters of one group overlap, clusters of different groups do not
    cl = (np.arange(400, 400 + n) // 200) % 64
    a, b = ref[cl == 2][:40].astype(np.float64), ref[cl == 3][:40].astype(np.float64)
    c = ref[cl == 13][:40].astype(np.float64)
    assert 0.2 < np.mean(a @ b.T) < 0.45 and abs(np.mean(a @ c.T)) < 0.1


def test_config4_ivfflat_lists4096_probes32_on_2m_rows():
    """BASELINE configs[4] at a size one test can afford: 2M x 3072, lists = 4096, probes = 32.  (i) the probed scan
    equals the oracle restricted to the probed lists; (ii) recall@10 against the exact scan; (iii) at most 1 % of
    the lists are empty."""
    n, dim, lists, probes, k = 2_000_000, 3072, 4096, 32, 10
    S = oracle.Synth()
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        st.build_ivf(lists=lists, iters=10, seed=1, default_probes=probes)
        off, rows, _ = st.debug_ivf()
        sizes = np.diff(off.astype(np.int64))
        assert off[-1] == n
        assert (sizes == 0).sum() <= lists // 100, int((sizes == 0).sum())
        hits = total = 0
        for qi in range(12):
            q = oracle.gen_query(S, qi, (qi * 37) % S.n_clusters, dim)
            exact = st.search(q, k, all_docs=True)
            got = st.search(q, k, all_docs=True, ivf=True)
            hits += len(np.intersect1d(got[2][0], exact[2][0]))
            total += len(exact[2][0])
            if qi < 4:
                mine = st.debug_probes(1, probes)[0]
                member = np.concatenate([rows[off[l]:off[l + 1]] for l in mine if l >= 0]).astype(np.int64)
                assert st.stats().last_rows_read == len(member)
                sub = oracle.gen_rows_at(S, member, dim)
                o_rows, o_sc, _ = oracle.topk(sub, None, None, q, k)
                # ties inside the oracle's subset are broken by position there, by global row here
                assert np.allclose(got[1][0], o_sc, rtol=1e-5)
                assert set(got[2][0].tolist()) == set(member[o_rows].tolist())
        assert hits / total >= 0.99, hits / total
        # a batch of 64 through the tensor-core ranking gives the same answers where the list ranking is clear
        qs = np.stack([oracle.gen_query(S, 100 + i, (i * 37) % S.n_clusters, dim) for i in range(64)])
        b = st.search(qs, k, all_docs=True, ivf=True)
        agree = 0
        for i in range(0, 64, 8):
            one = st.search(qs[i], k, all_docs=True, ivf=True)
            agree += int(np.array_equal(one[2][0], b[2][i]))
        assert agree >= 7
