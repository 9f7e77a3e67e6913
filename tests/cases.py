"""Deterministic inputs of the parity cases.  Inputs are rebuilt from seeds everywhere
(CPU container and GPU box); tests/golden/*.json holds the oracle's OUTPUTS for them,
minted by tests/golden/make_golden.py.  Edge cases follow SURVEY.md section 8c."""
from __future__ import annotations

import hashlib
from dataclasses import dataclass, field

import numpy as np

import oracle

REL_TOL = 1e-5   # BASELINE.json north_star: scores within 1e-5 relative in fp32
# Absolute slack for similarities near zero, where a relative bound is ill-conditioned (the dot
# product cancels; fp32 accumulation alone moves it by ~1e-7).  Irrelevant above the 0.7 floor.
ABS_TOL = 5e-7


@dataclass
class Case:
    name: str
    mat: np.ndarray                 # n x dim float32
    row_doc: np.ndarray             # n document indexes
    n_docs: int
    q: np.ndarray
    k: int
    min_sim: float = 0.7
    docs: list | None = None        # requested document indexes; None = all; may include unknown (>= n_docs)
    note: str = ""
    extra: dict = field(default_factory=dict)

    def allowed(self) -> np.ndarray:
        a = np.zeros(self.n_docs, dtype=np.uint8)
        if self.docs is None:
            a[:] = 1
        else:
            for d in self.docs:
                if 0 <= d < self.n_docs:
                    a[d] = 1
        return a

    def digest(self) -> str:
        h = hashlib.sha256()
        h.update(np.ascontiguousarray(self.mat).tobytes())
        h.update(np.ascontiguousarray(self.row_doc, dtype=np.uint32).tobytes())
        h.update(np.ascontiguousarray(self.q).tobytes())
        return h.hexdigest()


def _unit(rng, dim):
    v = rng.standard_normal(dim).astype(np.float32)
    return oracle.normalize(v)


def _planted(q, r, cos):
    """unit vector with cosine `cos` to q, in span(q, r) (r need not be orthogonal to q)."""
    q64 = q.astype(np.float64)
    r64 = r.astype(np.float64)
    r64 = r64 - (r64 @ q64) * q64
    r64 /= np.linalg.norm(r64)
    v = cos * q64 + np.sqrt(max(0.0, 1 - cos * cos)) * r64
    return oracle.normalize(v.astype(np.float32))


def build_cases() -> list[Case]:
    cases: list[Case] = []
    S = oracle.Synth()

    # configs[0]: 1 question, top_k=5 over 200 synthetic 3072-d chunks of one document
    n, dim = 200, 3072
    cases.append(Case("cfg1_single_doc_top5", oracle.gen_rows(S, 0, n, dim), oracle.row_docs(S, 0, n), 1,
                      oracle.gen_query(S, 0, 0, dim), 5, docs=[0], note="BASELINE configs[0]"))

    # document filter: 10 docs x 200 rows, 4 clusters => docs 0,4,8 share the query's cluster
    S4 = oracle.Synth(n_clusters=4)
    n = 2000
    mat = oracle.gen_rows(S4, 0, n, dim)
    rd = oracle.row_docs(S4, 0, n)
    q = oracle.gen_query(S4, 1, 0, dim)
    cases.append(Case("filter_two_of_three_matching_docs", mat, rd, 10, q, 10, docs=[0, 8]))
    cases.append(Case("filter_all_docs", mat, rd, 10, q, 10, docs=None))
    cases.append(Case("filter_empty_list", mat, rd, 10, q, 10, docs=[], note="ANY('{}') matches nothing"))
    cases.append(Case("filter_unknown_id_only", mat, rd, 10, q, 10, docs=[77]))
    cases.append(Case("filter_duplicate_and_unknown_ids", mat, rd, 10, q, 10, docs=[4, 4, 99, 4]))
    cases.append(Case("filter_doc_of_other_cluster", mat, rd, 10, q, 10, docs=[1, 2, 3], note="all below the 0.7 floor"))
    cases.append(Case("k_exceeds_api_cap", mat, rd, 10, q, 100, docs=None, note="k=100 (configs[3])"))
    cases.append(Case("k_one", mat, rd, 10, q, 1, docs=None))
    cases.append(Case("no_floor_negative_scores", mat[200:400], rd[200:400] - 1, 1, q, 128, min_sim=-np.inf,
                      docs=None, note="cross-cluster rows: scores around 0, both signs; full ordering"))

    # similarity floor edge: rows planted at 0.7 +- 1e-3 and +- 1e-2 around the floor
    rng = np.random.default_rng(7)
    dim2 = 1536
    q2 = _unit(rng, dim2)
    targets = [0.7 + d for d in (-1e-2, -1e-3, 1e-3, 1e-2, 0.05, 0.2, -0.3)] * 3
    rows = np.stack([_planted(q2, _unit(rng, dim2), c) for c in targets])
    cases.append(Case("floor_edge_dim1536", rows, np.zeros(len(rows), np.uint32), 1, q2, 20, docs=[0],
                      note="only rows at or above 0.7 survive (postgres.go:240)"))
    cases.append(Case("k_larger_than_matches", rows, np.zeros(len(rows), np.uint32), 1, q2, 20, min_sim=0.85, docs=[0]))

    # ties: the same vector stored under several chunks => equal scores, lower row first
    base = np.stack([_planted(q2, _unit(rng, dim2), c) for c in (0.9, 0.85, 0.8, 0.75)])
    tie = np.concatenate([base, base[[1, 1, 3]], base[[0]]])
    cases.append(Case("exact_ties_duplicate_rows", tie, np.zeros(len(tie), np.uint32), 1, q2, 6, docs=[0]))

    # zero-norm row: NaN distance in pgvector; this repository never returns it (DESIGN.md Deviations)
    z = np.concatenate([base, np.zeros((1, dim2), np.float32), base[[2]] * np.float32(0.5)])
    cases.append(Case("zero_norm_row_and_unnormalised_row", z, np.zeros(len(z), np.uint32), 1, q2, 10, docs=[0],
                      note="row 5 has norm 0.5: cosine still divides by the norms"))

    # unnormalised query: `<=>` divides by the query norm too
    cases.append(Case("unnormalised_query", base, np.zeros(len(base), np.uint32), 1, q2 * np.float32(3.0), 4, docs=[0]))

    # ragged dimension (not a multiple of the 128-float tile) and a tiny one
    rng = np.random.default_rng(11)
    for d in (100, 4, 772):
        qd = _unit(rng, d)
        m = np.stack([_planted(qd, _unit(rng, d), c) if d > 4 else _unit(rng, d) for c in np.linspace(0.2, 0.95, 37)])
        cases.append(Case(f"ragged_dim{d}", m, (np.arange(37) % 3).astype(np.uint32), 3, qd, 7,
                          min_sim=0.5 if d > 4 else -1.0, docs=[0, 2]))
    return cases


def case_by_name(name: str) -> Case:
    for c in build_cases():
        if c.name == name:
            return c
    raise KeyError(name)


def f32_bits(a) -> list[int]:
    return [int(x) for x in np.asarray(a, dtype=np.float32).view(np.uint32)]


def bits_f32(b) -> np.ndarray:
    return np.asarray(b, dtype=np.uint32).view(np.float32)


def check_against_oracle(case: Case, got_rows, got_scores, tol: float = REL_TOL):
    """The acceptance rule of BASELINE.json: scores within `tol` relative, same rows in the same
    order except where the oracle's own scores tie within `tol`, same filtering."""
    got_rows = np.asarray(got_rows, dtype=np.int64)
    got_scores = np.asarray(got_scores, dtype=np.float32)
    o_rows, o_scores, _ = oracle.topk(case.mat, case.row_doc, case.allowed(), case.q, case.k, case.min_sim)
    full = oracle.all_scores(case.mat, case.q).astype(np.float64)
    allowed = case.allowed()[np.asarray(case.row_doc)].astype(bool)
    assert len(np.unique(got_rows)) == len(got_rows), "a row was returned twice"
    assert np.all(allowed[got_rows]), "a row of a document that was not requested came back"
    assert np.all(np.diff(got_scores) <= 0), f"scores not descending: {got_scores}"
    for r, s in zip(got_rows, got_scores):
        ref = full[r]
        assert abs(float(s) - ref) <= tol * abs(ref) + ABS_TOL, (case.name, r, s, ref)
        assert ref >= case.min_sim - tol * abs(case.min_sim if np.isfinite(case.min_sim) else 0.0), (case.name, r, ref)
    # count: may differ only through rows within tol of the floor
    if len(got_rows) != len(o_rows):
        near = np.sum(allowed & (np.abs(full - case.min_sim) <= tol * abs(case.min_sim)))
        assert near > 0 and abs(len(got_rows) - len(o_rows)) <= near, (case.name, len(got_rows), len(o_rows))
        return
    # positional: equal rows, or a tie (within tol) between what each side put there
    for i, (gr, orow) in enumerate(zip(got_rows, o_rows)):
        if gr == orow:
            continue
        a, b = full[gr], full[orow]
        assert abs(a - b) <= tol * max(abs(a), abs(b)) + ABS_TOL, (case.name, i, gr, orow, a, b)
    assert np.allclose(got_scores, o_scores, rtol=tol, atol=ABS_TOL), (case.name, got_scores, o_scores)


def same_rows_up_to_ties(got_rows, ref_rows, ref_scores, tol: float = REL_TOL) -> bool:
    """True if got_rows equals ref_rows after collapsing runs of reference scores that tie within tol
    (the acceptance rule: identical order except for ties within the tolerance)."""
    got_rows, ref_rows = list(map(int, got_rows)), list(map(int, ref_rows))
    if len(got_rows) != len(ref_rows):
        return False
    ref_scores = np.asarray(ref_scores, dtype=np.float64)
    i = 0
    while i < len(ref_rows):
        j = i + 1
        while j < len(ref_rows) and abs(ref_scores[j] - ref_scores[j - 1]) <= tol * abs(ref_scores[j - 1]) + ABS_TOL:
            j += 1
        if set(got_rows[i:j]) != set(ref_rows[i:j]):
            return False
        i = j
    return True


def topk_matches_extended(got_rows, got_scores, ref_rows_ext, ref_scores_ext, k: int, tol: float = REL_TOL) -> str:
    """Acceptance rule against a reference list that runs a few places PAST k (so that a tie across the cut is
    visible): scores within tol, and at every rank either the reference's row or a row whose reference score ties
    the reference's score at that rank within tol.  Returns "" when it holds, else what is wrong."""
    got_rows = [int(r) for r in got_rows]
    ref_rows_ext = [int(r) for r in ref_rows_ext]
    ref_scores_ext = np.asarray(ref_scores_ext, dtype=np.float64)
    n_ref = min(k, len(ref_rows_ext))
    if len(got_rows) != n_ref:
        return f"{len(got_rows)} rows, reference {n_ref}"
    if len(set(got_rows)) != len(got_rows):
        return "a row came back twice"
    where = {r: i for i, r in enumerate(ref_rows_ext)}
    for i, (r, s) in enumerate(zip(got_rows, np.asarray(got_scores, dtype=np.float64))):
        if r not in where:
            return f"rank {i}: row {r} is not among the reference's first {len(ref_rows_ext)}"
        if abs(ref_scores_ext[where[r]] - ref_scores_ext[i]) > tol * abs(ref_scores_ext[i]) + ABS_TOL:
            return f"rank {i}: row {r} (reference score {ref_scores_ext[where[r]]}) where the reference has {ref_rows_ext[i]} ({ref_scores_ext[i]})"
        if abs(s - ref_scores_ext[where[r]]) > tol * abs(ref_scores_ext[where[r]]) + ABS_TOL:
            return f"rank {i}: score {s} of row {r}, reference {ref_scores_ext[where[r]]}"
    return ""
