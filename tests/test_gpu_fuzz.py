"""Randomised differential test: many small random configurations (rows, dimension, k, batch, document
filter, floor, upsert order) through the C ABI against the CPU oracle.  Seeds are fixed; every
configuration prints its parameters on failure."""
import numpy as np
import pytest

import oracle
from cases import ABS_TOL, REL_TOL, Case, check_against_oracle, topk_matches_extended
from doc_agents_b200 import GpuStore


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    """Compile the CUDA library once if it is stale (the GPU box normally receives it pre-built)."""
    return built_lib


pytestmark = pytest.mark.gpu


def _random_case(rng, idx):
    dim = int(rng.choice([4, 8, 12, 64, 100, 128, 256, 772, 1024, 1536, 3072]))
    n = int(rng.choice([1, 2, 3, 5, 31, 32, 33, 100, 255, 256, 257, 1000, 2049]))
    n_docs = int(rng.integers(1, min(n, 9) + 1))
    k = int(rng.choice([1, 2, 5, 10, 20, 33, 128]))
    centers = oracle.normalize(rng.standard_normal(dim).astype(np.float32))
    # rows: a mix of near-query rows and random rows, not normalised on purpose for some
    q = centers
    rows = []
    for i in range(n):
        mode = rng.integers(0, 4)
        noise = rng.standard_normal(dim).astype(np.float32)
        if rng.random() < 0.04:
            rows.append(np.zeros(dim, np.float32))                              # a zero row: NaN in pgvector, never a match here
            continue
        if mode == 0:
            v = q + np.float32(0.3 / np.sqrt(dim)) * noise * np.float32(rng.uniform(0.2, 3.0))
        elif mode == 1:
            v = noise
        elif mode == 2:
            v = (q * np.float32(rng.uniform(0.1, 5.0))).astype(np.float32)      # exact multiples: ties after normalisation
        else:
            v = -q + np.float32(0.5 / np.sqrt(dim)) * noise
        rows.append(v.astype(np.float32))
    mat = np.stack(rows)
    norm_mask = rng.random(n) < 0.7
    for i in np.nonzero(norm_mask)[0]:
        if np.any(mat[i]):
            mat[i] = oracle.normalize(mat[i])
    for i in np.nonzero(rng.random(n) < 0.1)[0]:                               # any length that keeps |row|^2 a normal float
        mat[i] = (mat[i] * np.float32(10.0 ** rng.uniform(-8, 8))).astype(np.float32)
    row_doc = rng.integers(0, n_docs, size=n).astype(np.uint32)
    docs = None if rng.random() < 0.3 else sorted(set(int(d) for d in rng.integers(0, n_docs + 2, size=rng.integers(0, n_docs + 2))))
    min_sim = float(rng.choice([0.7, 0.7, 0.0, -np.inf, 0.95, -0.5]))
    return Case(f"fuzz{idx}", mat, row_doc, n_docs, q * np.float32(rng.uniform(0.5, 2.0)), k, min_sim, docs)


@pytest.mark.parametrize("seed", range(12))
def test_random_configurations_match_the_oracle(seed):
    rng = np.random.default_rng(1000 + seed)
    S0 = oracle.Synth()
    for idx in range(12):
        c = _random_case(rng, seed * 100 + idx)
        n, dim = c.mat.shape
        chunk = oracle.chunk_ids(S0, 0, n)
        docs = oracle.doc_ids(S0, 0, c.n_docs + 8)
        info = dict(name=c.name, n=n, dim=dim, k=c.k, docs=c.docs, min_sim=c.min_sim, n_docs=c.n_docs)
        with GpuStore(dim=dim, capacity_rows=n + 4) as st:
            for d in range(c.n_docs):
                st.save_chunks_raw(docs[d], chunk[c.row_doc == d])
            # insert in two pieces, the second first half-overwritten later, to exercise the upsert path
            half = n // 2
            st.save_embeddings_raw(chunk[:half], c.mat[:half] * np.float32(0.5) if half else c.mat[:0])
            st.save_embeddings_raw(chunk[half:], c.mat[half:])
            if half:
                st.save_embeddings_raw(chunk[:half], c.mat[:half])
            filt = docs[: c.n_docs] if c.docs is None else (docs[np.asarray(c.docs, dtype=np.int64)] if len(c.docs) else docs[:0])
            try:
                ids, scores, rows = st.search(c.q, c.k, filt, min_similarity=c.min_sim)
                check_against_oracle(c, rows[0], scores[0])
                assert np.array_equal(ids[0], chunk[rows[0]])
                # a small batch of scaled copies of the query returns the same rows for each
                qs = np.stack([c.q, c.q * np.float32(2.0), c.q * np.float32(0.25)])
                b = st.search(qs, c.k, filt, min_similarity=c.min_sim)
                for i in range(3):
                    assert len(b[2][i]) == len(rows[0]), info
                    assert np.allclose(b[1][i], scores[0], rtol=REL_TOL, atol=ABS_TOL), info
            except AssertionError as e:
                raise AssertionError(f"{info}: {e}") from e


@pytest.mark.parametrize("seed", range(8))
def test_random_configurations_every_path_gives_the_same_answer(seed):
    """The same random cases on a store with the int8 sidecar: the default routing (int8 scan, int8/TF32 contraction,
    compacted lists) against the plain fp32 row loop, bit for bit, for one query and for batches of 3 and 20."""
    rng = np.random.default_rng(5000 + seed)
    S0 = oracle.Synth()
    done = 0
    while done < 10:
        c = _random_case(rng, seed * 100 + done)
        n, dim = c.mat.shape
        if dim % 32:
            continue
        done += 1
        chunk = oracle.chunk_ids(S0, 0, n)
        docs = oracle.doc_ids(S0, 0, c.n_docs + 8)
        info = dict(name=c.name, n=n, dim=dim, k=c.k, docs=c.docs, min_sim=c.min_sim, n_docs=c.n_docs)
        with GpuStore(dim=dim, capacity_rows=n + 4, q8_sidecar=True) as st:
            for d in range(c.n_docs):
                st.save_chunks_raw(docs[d], chunk[c.row_doc == d])
            st.save_embeddings_raw(chunk, c.mat)
            filt = docs[: c.n_docs] if c.docs is None else (docs[np.asarray(c.docs, dtype=np.int64)] if len(c.docs) else docs[:0])
            for batch in (1, 3, 20):
                qs = np.stack([c.q * np.float32(1 + 0.01 * i) + np.float32(0.02 * i / np.sqrt(dim)) * rng.standard_normal(dim).astype(np.float32)
                               for i in range(batch)]).astype(np.float32)
                a = st.search(qs, c.k, filt, min_similarity=c.min_sim)
                b = st.search(qs, c.k, filt, min_similarity=c.min_sim, no_q8=True, no_tensor=True, no_compact=True)
                for i in range(batch):
                    assert np.array_equal(a[2][i], b[2][i]), (info, batch, i)
                    assert np.array_equal(a[1][i].view(np.uint32), b[1][i].view(np.uint32)), (info, batch, i)
                    assert np.array_equal(a[0][i], b[0][i]), (info, batch, i)
            check_against_oracle(c, st.search(c.q, c.k, filt, min_similarity=c.min_sim)[2][0],
                                 st.search(c.q, c.k, filt, min_similarity=c.min_sim)[1][0])


@pytest.mark.parametrize("seed", range(6))
def test_random_write_sequences_match_a_model_of_the_tables(seed):
    """The write side (SaveChunks / SaveEmbeddings: postgres.go:142-215) under random sequences -- documents saved in pieces and
    interleaved, chunks re-embedded, chunks moved to another document, embeddings for a subset only -- against a Python model of
    the two tables; after every few writes a search (one query: the range path or the bitmap paths; a batch: per-row document
    slots) must equal the oracle over the model's current rows.  A row's number is the order of its chunk's first embedding."""
    rng = np.random.default_rng(9000 + seed)
    S0 = oracle.Synth()
    dim, n_chunks, n_docs = 64, 500, 9
    chunk = oracle.chunk_ids(S0, 0, n_chunks)
    docs = oracle.doc_ids(S0, 0, n_docs + 2)
    q = oracle.normalize(rng.standard_normal(dim).astype(np.float32))
    doc_of = {}            # chunk index -> document index        (table chunks)
    row_of = {}            # chunk index -> row                   (table embeddings, in order of first insert)
    vec_of = {}
    def new_vec():
        v = (q * np.float32(rng.uniform(0.5, 1.0)) + np.float32(rng.uniform(0.02, 0.4) / np.sqrt(dim)) * rng.standard_normal(dim)).astype(np.float32)
        return v if rng.random() < 0.5 else oracle.normalize(v)
    with GpuStore(dim=dim, capacity_rows=n_chunks) as st:
        for step in range(60):
            op = rng.random()
            if op < 0.35 or not doc_of:                                   # SaveChunks: a run of chunks for one document (new or moved)
                d = int(rng.integers(0, n_docs))
                start = int(rng.integers(0, n_chunks - 1))
                idx = list(range(start, min(n_chunks, start + int(rng.integers(1, 40)))))
                if rng.random() < 0.3:
                    idx = [int(i) for i in rng.choice(n_chunks, size=int(rng.integers(1, 12)), replace=False)]
                st.save_chunks_raw(docs[d], chunk[idx])
                for i in idx:
                    doc_of[i] = d
            elif op < 0.8:                                                # SaveEmbeddings for chunks that exist
                have = sorted(doc_of)
                take = [int(i) for i in rng.choice(have, size=min(len(have), int(rng.integers(1, 50))), replace=False)]
                if rng.random() < 0.5:
                    take.sort()
                vecs = np.stack([new_vec() for _ in take])
                st.save_embeddings_raw(chunk[take], vecs)
                for i, v in zip(take, vecs):
                    if i not in row_of:
                        row_of[i] = len(row_of)
                    vec_of[i] = v
            if step % 4 != 3 or not row_of:
                continue
            n = len(row_of)
            by_row = sorted(row_of, key=row_of.get)
            mat = np.stack([vec_of[i] for i in by_row])
            row_doc = np.asarray([doc_of[i] for i in by_row], dtype=np.uint32)
            assert st.stats().rows == n
            k = int(rng.choice([1, 5, 10, 20, 40]))
            floor = float(rng.choice([0.7, 0.9, -1.0]))
            wanted = sorted(set(int(d) for d in rng.integers(0, n_docs + 2, size=int(rng.integers(1, 6)))))
            allowed = np.zeros(n_docs + 2, np.uint8)
            allowed[wanted] = 1
            info = dict(seed=seed, step=step, n=n, k=k, floor=floor, wanted=wanted)
            o_rows, o_sc, _ = oracle.topk(mat, row_doc, allowed, q, k + 8, floor)
            for kw in (dict(), dict(no_compact=True), dict(read_all_rows=True)):
                ids, sc, rows = st.search(q, k, docs[wanted], min_similarity=floor, **kw)
                why = topk_matches_extended(rows[0], sc[0], o_rows, o_sc, k)
                assert why == "", (info, kw, why)
                assert np.array_equal(ids[0], chunk[[by_row[r] for r in rows[0]]]), (info, kw)
            qs = np.stack([q, q * np.float32(3.0)])
            b = st.search(qs, k, docs[wanted], min_similarity=floor)
            for i in range(2):
                assert topk_matches_extended(b[2][i], b[1][i], o_rows, o_sc, k) == "", (info, "batch", i)
            counts = st.doc_rows(docs[: n_docs])
            assert list(counts) == [int((row_doc == d).sum()) for d in range(n_docs)], info
