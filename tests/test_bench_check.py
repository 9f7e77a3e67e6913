"""CPU tests of bench.py's own checker: the oracle comparison every bench line carries (`result_check`), fed with
the oracle's answer (must pass), with a wrong row, a wrong score and a swapped near-tie."""
import numpy as np

import bench
import oracle


def _answer(mat, qs, k, floor):
    b = len(qs)
    ids = np.zeros((b, k, 16), np.uint8)
    sc = np.zeros((b, k), np.float32)
    rows = np.zeros((b, k), np.uint32)
    n = np.zeros(b, np.int32)
    for i in range(b):
        r, s, _ = oracle.topk(mat, None, None, qs[i], k, floor)
        n[i] = len(r)
        rows[i, :len(r)] = r
        sc[i, :len(r)] = s
    return ids, sc, rows, n


def test_result_check_accepts_the_oracle_and_rejects_wrong_answers(monkeypatch):
    monkeypatch.setattr(bench, "N_ROWS", 300_000)
    monkeypatch.setattr(bench, "DIM", 128)
    monkeypatch.setattr(bench, "TOP_K", 10)
    monkeypatch.setattr(bench, "MIN_SIM", 0.6)
    S = oracle.Synth()
    cl = lambda i: (i * 37) % S.n_clusters     # noqa: E731
    qs = np.stack([oracle.gen_query(S, i, cl(i), 128) for i in range(12)])
    mat = oracle.gen_rows(S, 0, 300_000, 128)
    good = _answer(mat, qs, 10, 0.6)
    res = bench.check_against_oracle(good, qs, cl, None)
    assert res["matches_oracle"] and res["queries_checked"] == 8 and res["rows_checked"] > 0
    bad = tuple(a.copy() for a in good)
    bad[2][0, 1] += 1                                       # a row the oracle does not rank there
    assert not bench.check_against_oracle(bad, qs, cl, None)["matches_oracle"]
    bad = tuple(a.copy() for a in good)
    bad[1][0, 0] *= np.float32(1.0001)                      # a score off by 1e-4 relative
    assert not bench.check_against_oracle(bad, qs, cl, None)["matches_oracle"]
    bad = tuple(a.copy() for a in good)
    bad[3][0] -= 1                                          # a hit missing
    assert not bench.check_against_oracle(bad, qs, cl, None)["matches_oracle"]
    # a filter: the oracle scores the requested documents only
    monkeypatch.setattr(bench, "FILTER_DOCS", 4)
    docs = [0, 1, cl(0), cl(0) + S.n_clusters]
    sub_rows = np.concatenate([np.arange(d * 200, d * 200 + 200) for d in sorted(set(docs))])
    f = _answer(mat[sub_rows], qs[:1], 10, 0.6)
    f[2][0, :f[3][0]] = sub_rows[f[2][0, :f[3][0]]]
    assert bench.check_against_oracle(f, qs[:1], cl, docs)["matches_oracle"]


def test_reference_arm_scans_what_fits(monkeypatch):
    monkeypatch.setattr(bench, "N_ROWS", 10_000_000)
    monkeypatch.setattr(bench, "DIM", 3072)
    monkeypatch.delenv("DA_CPU_SAMPLE_ROWS", raising=False)
    monkeypatch.setattr(bench, "host_mem_available_gb", lambda: 196.0)
    assert bench.cpu_rows_that_fit(full=True) == 10_000_000          # 122.9 GB x 1.15 fits in 196 GB: measured, not extrapolated
    assert bench.cpu_rows_that_fit(full=False) == 1_000_000          # the cpu_baseline leg stays a bounded sample
    monkeypatch.setattr(bench, "host_mem_available_gb", lambda: 62.0)
    assert bench.cpu_rows_that_fit(full=True) == 4_000_000           # the largest multiple of 1M rows that fits
    a, b = bench.bench_config(1, "cfg3_b1"), bench.bench_config(1, "cfg3_b1")
    assert a == b and a["rows"] == 10_000_000 and "l2" in a
