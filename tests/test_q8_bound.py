"""CPU check of the inequality the int8 sidecar rests on (doc_agents_b200/csrc/da_q8.cu):

    | cos(q, x) - s_q s_x (q^ . x^) / (|q| |x|) |  <=  e_x + e_q + e_x e_q,

with x^ = rint(x / s_x), s_x = max|x| / 127, e_x = |x - s_x x^| / |x| (same for q).  numpy in float64 restates what the
quantisation kernel computes; random, heavy-tailed and degenerate vectors.  No GPU, no library call."""
import numpy as np
import pytest


def quantise(v):
    v = v.astype(np.float64)
    amax = np.abs(v).max()
    if amax == 0:
        return np.zeros(v.shape, np.int64), 0.0, 0.0
    s = amax / 127.0
    q = np.clip(np.rint(v / s), -127, 127).astype(np.int64)
    e = np.linalg.norm(v - s * q) / np.linalg.norm(v)
    return q, s, e


def vectors(rng, dim):
    g = rng.standard_normal(dim)
    yield g
    yield g ** 3                                             # heavy tails
    h = g.copy(); h[rng.integers(dim)] = 50.0; yield h       # one dominant element
    yield np.sign(g) * 1e-20                                 # tiny, all the same magnitude
    yield np.where(rng.random(dim) < 0.02, g * 30, g * 0.01)  # sparse spikes
    yield np.arange(dim, dtype=np.float64) - dim / 2         # a ramp


@pytest.mark.parametrize("dim", [32, 256, 1536, 3072])
def test_interval_holds_the_true_cosine(dim):
    rng = np.random.default_rng(dim)
    worst = 0.0
    for _ in range(20):
        for x in vectors(rng, dim):
            for q in vectors(rng, dim):
                xq, sx, ex = quantise(x)
                qq, sq, eq = quantise(q)
                cos = float(x @ q / (np.linalg.norm(x) * np.linalg.norm(q)))
                dot = int(xq @ qq)
                assert abs(dot) <= 127 * 127 * dim < 2 ** 31           # exact in int32
                approx = sq * sx * dot / (np.linalg.norm(x) * np.linalg.norm(q))
                E = ex + eq + ex * eq
                assert abs(cos - approx) <= E + 1e-12, (dim, abs(cos - approx), E)
                worst = max(worst, (abs(cos - approx) - 1e-12) / max(E, 1e-9))
    assert worst <= 1.0


def test_typical_bound_for_embedding_like_vectors():
    """Unit vectors with Gaussian elements at dim 3072: the bound the scan works with is about 0.01 per side."""
    rng = np.random.default_rng(0)
    es = [quantise(rng.standard_normal(3072))[2] for _ in range(200)]
    assert 0.005 < float(np.mean(es)) < 0.015
