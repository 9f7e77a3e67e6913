"""numpy restatement of oracle.c -- a second, independent CPU statement of the
same path, used by tests/ to cross-check the C oracle (generator bit-exact,
cosine within float rounding).  TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED:
see the header of oracle.c for why no reference vector can pin either.
"""
from __future__ import annotations

import numpy as np

M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
GOLD = np.uint64(0x9E3779B97F4A7C15)
PAIR = np.uint64(0xD1B54A32D192ED03)


def mix64(z):
    z = np.asarray(z, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = z ^ (z >> np.uint64(30)); z = z * np.uint64(0xBF58476D1CE4E5B9)
        z = z ^ (z >> np.uint64(27)); z = z * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def stream_key(seed: int, index):
    with np.errstate(over="ignore"):
        return mix64(np.uint64(seed) ^ ((np.asarray(index, dtype=np.uint64) + np.uint64(1)) * GOLD))


def ih4(w):
    w = np.asarray(w, dtype=np.uint32)
    s = (w & np.uint32(255)).astype(np.int32) + ((w >> np.uint32(8)) & np.uint32(255)).astype(np.int32) \
        + ((w >> np.uint32(16)) & np.uint32(255)).astype(np.int32) + (w >> np.uint32(24)).astype(np.int32)
    return s - 510


def elems(key, dim: int):
    """int32 [..., dim] Irwin-Hall elements of the stream(s) `key`."""
    key = np.asarray(key, dtype=np.uint64)[..., None]
    pairs = (np.arange((dim + 1) // 2, dtype=np.uint64) + np.uint64(1)) * PAIR
    with np.errstate(over="ignore"):
        h = mix64(key + pairs)
    lo = (h & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    hi = (h >> np.uint64(32)).astype(np.uint32)
    w = np.stack([lo, hi], axis=-1).reshape(*h.shape[:-1], -1)[..., :dim]
    return ih4(w)


def gen_vecs(seed_noise: int, noise_index, seed_centroid: int, cluster, cmul: int, nmul: int, dim: int):
    u = cmul * elems(stream_key(seed_centroid, cluster), dim) + nmul * elems(stream_key(seed_noise, noise_index), dim)
    ss = (u.astype(np.int64) ** 2).sum(axis=-1)
    mag = np.sqrt(ss.astype(np.float64)).astype(np.float32)
    x = u.astype(np.float32)
    nz = ss != 0
    x[nz] = x[nz] / mag[nz][..., None]
    return x


def gen_rows(p, row_lo: int, n: int, dim: int):
    rows = np.arange(row_lo, row_lo + n, dtype=np.uint64)
    cl = (rows // np.uint64(p.chunks_per_doc)) % np.uint64(p.n_clusters)
    return gen_vecs(p.seed_data, rows, p.seed_centroid, cl, p.cmul, p.nmul, dim)


def gen_query(p, qi: int, cluster: int, dim: int):
    return gen_vecs(p.seed_query, np.array([qi]), p.seed_centroid, np.array([cluster]), p.cmul, p.nmul, dim)[0]


def normalize(v):
    v = np.asarray(v, dtype=np.float32).copy()
    ss = float((v.astype(np.float64) ** 2).sum())
    if ss == 0.0:
        return v
    return v / np.float32(np.sqrt(ss))


def cosine_distance(a, b, f64: bool = False):
    """pgvector `<=>`: float (or double) accumulators, double finish, clamp."""
    a = np.asarray(a, dtype=np.float32)
    b = np.asarray(b, dtype=np.float32)
    if f64:
        a64, b64 = a.astype(np.float64), b.astype(np.float64)
        dot, na, nb = (a64 * b64).sum(), (a64 * a64).sum(), (b64 * b64).sum()
    else:
        # numpy's pairwise float32 sum: a different (valid) summation order than
        # the C loop -- pgvector's own order is build-dependent (oracle.c header).
        dot = np.float32((a * b).sum(dtype=np.float32))
        na = np.float32((a * a).sum(dtype=np.float32))
        nb = np.float32((b * b).sum(dtype=np.float32))
    with np.errstate(invalid="ignore", divide="ignore"):
        sim = np.float64(dot) / np.sqrt(np.float64(na) * np.float64(nb))
    if sim > 1.0:
        sim = 1.0
    elif sim < -1.0:
        sim = -1.0
    return 1.0 - sim


def topk(mat, row_doc, doc_allowed, q, k: int, min_sim: float = 0.7, f64: bool = True):
    """Store.TopK restated with numpy (double accumulators by default)."""
    mat = np.asarray(mat, dtype=np.float32)
    q = np.asarray(q, dtype=np.float32)
    acc = np.float64 if f64 else np.float32
    m, qq = mat.astype(acc), q.astype(acc)
    dot = m @ qq
    na = (m * m).sum(axis=1)
    nb = (qq * qq).sum()
    with np.errstate(invalid="ignore", divide="ignore"):
        sim = dot.astype(np.float64) / np.sqrt(na.astype(np.float64) * np.float64(nb))
    sim = np.clip(sim, -1.0, 1.0)
    dist = 1.0 - sim
    ok = ~np.isnan(dist) & ((1.0 - dist) >= min_sim)
    if doc_allowed is not None:
        ok &= np.asarray(doc_allowed, dtype=bool)[np.asarray(row_doc)]
    idx = np.nonzero(ok)[0]
    order = np.lexsort((idx, dist[idx]))[:max(k, 0)]
    rows = idx[order].astype(np.uint32)
    return rows, (1.0 - dist[rows]).astype(np.float32), dist[rows]
