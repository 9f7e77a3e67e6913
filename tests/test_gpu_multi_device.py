"""GPU suite for one process driving several GPUs (what the Go query agent does): the peer-gather
path (scan epilogue stores its records into shard 0 over NVLink) and the NCCL path must both equal
a single-GPU store; IVFFlat and the batched tensor path on two shards."""
import os

import numpy as np
import pytest

import oracle
from doc_agents_b200 import GpuStore
from gpu_util import device_count


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    """Compile the CUDA library once if it is stale (the GPU box normally receives it pre-built)."""
    return built_lib


pytestmark = [pytest.mark.gpu, pytest.mark.skipif(device_count() < 2, reason="needs 2 GPUs in one process")]

S = oracle.Synth(n_clusters=8)
N, DIM, K = 40_000, 1536, 10


def _two(p2p: bool):
    if p2p:
        os.environ.pop("DA_DISABLE_P2P", None)
    else:
        os.environ["DA_DISABLE_P2P"] = "1"
    try:
        st = GpuStore(dim=DIM, capacity_rows=N, devices=(0, 1))
    finally:
        os.environ.pop("DA_DISABLE_P2P", None)
    st.generate_synthetic(S, 0, N)
    assert st.stats().peer_gather == (1 if p2p else 0)
    return st


@pytest.mark.parametrize("p2p", [True, False], ids=["peer_gather", "nccl_allgather"])
def test_two_shards_equal_one_shard(p2p):
    docs = oracle.doc_ids(S, 0, N // 200)
    with GpuStore(dim=DIM, capacity_rows=N) as one, _two(p2p) as two:
        one.generate_synthetic(S, 0, N)
        qs = np.stack([oracle.gen_query(S, i, i % 8, DIM) for i in range(40)])
        for batch in (1, 5, 9, 40):                      # 1 pass, 1 pass of 5, 2 passes, tensor path
            for kw in (dict(all_docs=True), dict(doc_ids=docs[::3])):
                a = one.search(qs[:batch], K, **kw)
                b = two.search(qs[:batch], K, **kw)
                b2 = two.search(qs[:batch], K, **kw)     # back to back: gather blocks are reused
                for i in range(batch):
                    assert np.array_equal(a[2][i], b[2][i]) and np.array_equal(a[0][i], b[0][i])
                    assert np.array_equal(a[1][i].view(np.uint32), b[1][i].view(np.uint32))
                    assert np.array_equal(b[2][i], b2[2][i])
        ms = two.bench_resident(qs[0], K, 20, all_docs=True)      # 20 searches in flight without host syncs
        assert ms > 0
        again = two.search(qs[:5], K, all_docs=True)
        ref = one.search(qs[:5], K, all_docs=True)
        assert all(np.array_equal(again[2][i], ref[2][i]) for i in range(5))


def test_ivf_on_two_shards_shares_centroids_and_probes_exactly():
    with _two(True) as two:
        two.build_ivf(lists=32, iters=6, seed=2)
        off0, rows0, cent0 = two.debug_ivf(0)
        off1, rows1, cent1 = two.debug_ivf(1)
        assert np.array_equal(cent0.view(np.uint32), cent1.view(np.uint32))      # data-parallel k-means: same centroids
        assert off0[-1] == N // 2 and off1[-1] == N // 2
        for qi in range(8):
            q = oracle.gen_query(S, qi, qi % 8, DIM)
            exact = two.search(q, K, all_docs=True)
            full = two.search(q, K, all_docs=True, ivf=True, probes=32)
            assert np.array_equal(exact[2][0], full[2][0]) and np.array_equal(exact[0][0], full[0][0])
            # (balanced lists: a cluster holds about four of the 32 lists, sometimes five or six)
            some = two.search(q, K, all_docs=True, ivf=True, probes=8)
            assert len(np.intersect1d(some[2][0], exact[2][0])) >= K - 1


def test_split_plane_store_on_two_shards_equals_fp32_single_shard():
    docs = oracle.doc_ids(S, 0, N // 200)
    with GpuStore(dim=DIM, capacity_rows=N) as one, GpuStore(dim=DIM, capacity_rows=N, devices=(0, 1), split16=True) as two:
        one.generate_synthetic(S, 0, N)
        two.generate_synthetic(S, 0, N)
        qs = np.stack([oracle.gen_query(S, i, i % 8, DIM) for i in range(24)])
        for batch in (1, 7, 24):                          # SIMT upper-plane scan, two passes, bf16 tensor path
            for kw in (dict(all_docs=True), dict(doc_ids=docs[::3])):
                a = one.search(qs[:batch], K, no_tensor=True, **kw)
                b = two.search(qs[:batch], K, **kw)
                for i in range(batch):
                    assert np.array_equal(a[2][i], b[2][i]) and np.array_equal(a[0][i], b[0][i])
                    assert np.array_equal(a[1][i].view(np.uint32), b[1][i].view(np.uint32))


def test_int8_sidecar_store_on_two_shards_equals_fp32_single_shard():
    docs = oracle.doc_ids(S, 0, N // 200)
    with GpuStore(dim=DIM, capacity_rows=N) as one, GpuStore(dim=DIM, capacity_rows=N, devices=(0, 1), q8_sidecar=True) as two:
        one.generate_synthetic(S, 0, N)
        two.generate_synthetic(S, 0, N)
        qs = np.stack([oracle.gen_query(S, i, i % 8, DIM) for i in range(24)])
        for batch in (1, 7, 24):                          # int8 scan, two passes of it, tensor path over the fp32 rows
            for kw in (dict(all_docs=True), dict(doc_ids=docs[::3])):
                a = one.search(qs[:batch], K, no_tensor=True, **kw)
                b = two.search(qs[:batch], K, **kw)
                for i in range(batch):
                    assert np.array_equal(a[2][i], b[2][i]) and np.array_equal(a[0][i], b[0][i])
                    assert np.array_equal(a[1][i].view(np.uint32), b[1][i].view(np.uint32))


def test_upsert_loaded_shards_answer_like_one_shard_up_to_exact_ties():
    """A store filled by SaveEmbeddings places every new document on the emptiest shard, so a row's global number
    (shard x capacity + local row) depends on the device count.  Scores and the set of rows never do; the ORDER among
    rows whose scores are exactly equal follows the global number (ties are unordered in the
    reference's SQL too).  Duplicate vectors are planted in different documents to make such ties."""
    dim, k, per, n_docs = 256, 12, 50, 8
    rng = np.random.default_rng(21)
    q = oracle.normalize(rng.standard_normal(dim).astype(np.float32))
    vecs = np.stack([oracle.normalize((0.9 * q + 0.03 * rng.standard_normal(dim)).astype(np.float32)) for _ in range(n_docs * per)])
    for d in range(1, n_docs):                       # the best row of document 0 repeated in every other document
        vecs[d * per + 3] = vecs[3]
    chunk = oracle.chunk_ids(S, 0, n_docs * per)
    docs = oracle.doc_ids(S, 0, n_docs)
    answers = []
    for devices in ((0,), (0, 1)):
        with GpuStore(dim=dim, capacity_rows=1024, devices=devices, min_similarity=0.5) as st:
            for d in range(n_docs):
                st.save_chunks_raw(docs[d], chunk[d * per:(d + 1) * per])
                st.save_embeddings_raw(chunk[d * per:(d + 1) * per], vecs[d * per:(d + 1) * per])
            if len(devices) == 2:
                assert list(st.stats().rows_per_device[:2]) == [n_docs * per // 2, n_docs * per // 2]
            ids, sc, rows = st.search(q, k, docs)
            answers.append(([bytes(i) for i in ids[0]], sc[0].copy()))
    (ids1, sc1), (ids2, sc2) = answers
    assert np.array_equal(sc1, sc2)                                  # the same scores, position by position
    assert set(ids1) == set(ids2)                                    # the same rows
    i = 0
    while i < k:                                                     # the same order outside runs of exactly equal scores
        j = i + 1
        while j < k and sc1[j] == sc1[i]:
            j += 1
        assert set(ids1[i:j]) == set(ids2[i:j])
        i = j
    o_rows, o_sc, _ = oracle.topk(vecs, None, None, q, k, 0.5)
    assert np.allclose(sc1, o_sc, rtol=1e-5) and set(ids1) == {bytes(chunk[r]) for r in o_rows} or sc1[-1] == sc1[-2]
