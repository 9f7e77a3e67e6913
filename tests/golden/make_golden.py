"""Mints tests/golden/topk_cases.json: the oracle's outputs (float accumulators = pgvector
restated, and double accumulators = truth) for every case of tests/cases.py, cross-checked
against the independent numpy restatement before anything is written.

    python tests/golden/make_golden.py

The reference holds no golden vector for this path and cannot be run here (Go + pgvector in
Postgres), so these fixtures pin the ORACLE against regressions and against the numpy
restatement -- they do not pin it to pgvector (PARITY UNPINNED, see oracle/oracle.c).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

import oracle                      # noqa: E402
from oracle import oracle_np      # noqa: E402
from cases import build_cases, f32_bits, REL_TOL, ABS_TOL  # noqa: E402


def main():
    out = {"_about": "oracle outputs per case; scores are float32 bit patterns; made by make_golden.py",
           "cases": {}}
    for c in build_cases():
        allowed = c.allowed()
        r32, s32, d32 = oracle.topk(c.mat, c.row_doc, allowed, c.q, c.k, c.min_sim, f64=False)
        r64, s64, d64 = oracle.topk(c.mat, c.row_doc, allowed, c.q, c.k, c.min_sim, f64=True)
        rn, sn, dn = oracle_np.topk(c.mat, c.row_doc, allowed, c.q, c.k, c.min_sim, f64=True)
        # cross-check 1: C double accumulators == numpy double accumulators (rows; scores to 1e-12)
        assert np.array_equal(r64, rn), (c.name, r64, rn)
        assert np.allclose(d64, dn, rtol=0, atol=1e-12), c.name
        # cross-check 2: float accumulators within tolerance of truth
        assert len(r32) == len(r64), c.name
        assert np.allclose(s32, s64, rtol=REL_TOL, atol=ABS_TOL), (c.name, s32, s64)
        out["cases"][c.name] = {
            "digest": c.digest(), "n": int(c.mat.shape[0]), "dim": int(c.mat.shape[1]), "k": int(c.k),
            "min_sim": None if not np.isfinite(c.min_sim) else float(c.min_sim),
            "docs": c.docs, "note": c.note,
            "f32": {"rows": [int(x) for x in r32], "score_bits": f32_bits(s32)},
            "f64": {"rows": [int(x) for x in r64], "score_bits": f32_bits(s64),
                    "dist": [float.hex(float(x)) for x in d64]},
        }
        print(f"{c.name:40s} n={c.mat.shape[0]:5d} dim={c.mat.shape[1]:5d} k={c.k:3d} -> {len(r32)} rows")
    # generator known-answers: first elements and a digest of the synthetic corpus
    S = oracle.Synth()
    rows = oracle.gen_rows(S, 0, 400, 3072)
    assert np.array_equal(rows.view(np.uint32), oracle_np.gen_rows(S, 0, 400, 3072).view(np.uint32))
    import hashlib
    out["synthetic"] = {
        "rows_0_400_dim3072_sha256": hashlib.sha256(rows.tobytes()).hexdigest(),
        "row0_first8_bits": f32_bits(rows[0, :8]),
        "row399_last4_bits": f32_bits(rows[399, -4:]),
        "query0_cluster0_first4_bits": f32_bits(oracle.gen_query(S, 0, 0, 3072)[:4]),
        "chunk_id_row0": oracle.chunk_ids(S, 0, 1)[0].tolist(),
        "chunk_id_row123456": oracle.chunk_ids(S, 123456, 1)[0].tolist(),
        "doc_id_0": oracle.doc_ids(S, 0, 1)[0].tolist(),
    }
    with open(os.path.join(HERE, "topk_cases.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote topk_cases.json")


if __name__ == "__main__":
    main()
