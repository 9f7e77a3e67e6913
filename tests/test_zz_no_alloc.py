"""GPU test of da_scan_topk_noalloc, the entry point for hosts that must not allocate
(a latency-critical serving loop, a process whose HBM is already fully committed, code running
where the CUDA allocator is off limits).  It keeps its scratch in a fixed __device__ array inside
the library instead of cudaMalloc'ed buffers.

The test plays such a host: it brings its own device buffers (made here with torch, which is only
the harness's way to obtain device memory), calls the entry point through the C ABI twice in a row
-- the second call finds the scratch array as the first one left it -- and checks both answers
against the allocating entry point (da_store_search) on the same rows and, so that the entry point
stands on its own, against the CPU oracle.  A later call with a dimension that selects another
kernel variant (128-bit loads, ragged tiles) and a non-zero row base follows.  The file name makes
pytest collect it last, so a failure here cannot hide the other tests under `-x`.
"""
import ctypes as C

import numpy as np
import pytest

import oracle
from doc_agents_b200 import GpuStore
from doc_agents_b200 import _lib as L


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    """Compile the CUDA library once if it is stale (the GPU box normally receives it pre-built)."""
    return built_lib


pytestmark = pytest.mark.gpu


def _dev(t):
    return C.c_void_p(t.data_ptr())


def test_no_alloc_scan_matches_store_search_twice_in_a_row():
    import torch

    S = oracle.Synth(n_clusters=4)
    n, dim, k = 5000, 1536, 10
    lib = L.load()
    with GpuStore(dim=dim, capacity_rows=n) as st:
        st.generate_synthetic(S, 0, n)
        vecs, inv, slot, ids = st.debug_rows(0, 0, n)
        docs = oracle.doc_ids(S, 0, 25)
        wanted = [0, 4, 5, 8, 12, 13, 16, 21]
        bitmap = np.zeros(2, dtype=np.uint32)
        for d in wanted:
            bitmap[d >> 5] |= np.uint32(1 << (d & 31))

        # the caller's own device buffers
        dev = torch.device("cuda", 0)
        d_mat = torch.from_numpy(vecs).to(dev)
        d_inv = torch.from_numpy(inv).to(dev)
        d_slot = torch.from_numpy(slot.astype(np.int32)).to(dev)
        d_ids = torch.from_numpy(ids).to(dev)
        d_bitmap = torch.from_numpy(bitmap.astype(np.int32)).to(dev)
        d_scores = torch.zeros(k, dtype=torch.float32, device=dev)
        d_rows = torch.zeros(k, dtype=torch.int32, device=dev)
        d_out_ids = torch.zeros(k * 16, dtype=torch.uint8, device=dev)
        d_n = torch.zeros(1, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()

        # two calls in a row: different query, different filter; the second reuses the scratch as left
        for qi, cluster, use_filter in ((0, 0, True), (1, 1, False)):
            q = oracle.gen_query(S, qi, cluster, dim)
            d_q = torch.from_numpy(q).to(dev)
            torch.cuda.synchronize()
            rc = lib.da_scan_topk_noalloc(0, _dev(d_mat), _dev(d_inv), _dev(d_slot) if use_filter else None,
                                          _dev(d_ids), _dev(d_bitmap) if use_filter else None, _dev(d_q), n, 0, dim,
                                          k, 0.7, None, _dev(d_scores), _dev(d_rows), _dev(d_out_ids), _dev(d_n))
            assert rc == L.DA_OK, L.last_error()
            torch.cuda.synchronize()
            cnt = int(d_n.cpu()[0])
            got_scores = d_scores.cpu().numpy()[:cnt]
            got_rows = d_rows.cpu().numpy().astype(np.uint32)[:cnt]
            got_ids = d_out_ids.cpu().numpy().reshape(k, 16)[:cnt]

            if use_filter:
                ref_ids, ref_scores, ref_rows = st.search(q, k, docs[wanted])
            else:
                ref_ids, ref_scores, ref_rows = st.search(q, k, all_docs=True)
            assert cnt == len(ref_rows[0]) == k
            assert np.array_equal(got_rows, ref_rows[0])
            assert np.allclose(got_scores, ref_scores[0], rtol=1e-5, atol=0)   # the tolerance of the parity suite
            assert np.array_equal(got_ids, ref_ids[0])
            # ... and against the oracle on the same 5 000 rows
            allowed = np.zeros(25, np.uint8)
            allowed[wanted if use_filter else slice(None)] = 1
            o_rows, o_scores, _ = oracle.topk(vecs, oracle.row_docs(S, 0, n), allowed, q, k)
            assert np.array_equal(got_rows, o_rows)
            assert np.allclose(got_scores, o_scores, rtol=1e-5, atol=0)
            assert np.array_equal(got_ids, oracle.chunk_ids(S, 0, n)[o_rows])

        # another kernel variant (dim % 8 != 0: 128-bit loads, ragged tiles) and a row base, after the first calls
        dim2, n2, base = 772, 1000, 7000
        rng = np.random.default_rng(5)
        q2 = oracle.normalize(rng.standard_normal(dim2).astype(np.float32))
        m2 = (0.6 * q2[None, :] + rng.standard_normal((n2, dim2)).astype(np.float32) / np.sqrt(dim2)).astype(np.float32)
        inv2 = (1.0 / np.linalg.norm(m2.astype(np.float64), axis=1)).astype(np.float32)
        ids2 = oracle.chunk_ids(S, 0, n2)
        t_m2, t_inv2, t_ids2, t_q2 = (torch.from_numpy(a).to(dev) for a in (m2, inv2, ids2, q2))
        prev = torch.cuda.current_device()
        rc = lib.da_scan_topk_noalloc(0, _dev(t_m2), _dev(t_inv2), None, _dev(t_ids2), None, _dev(t_q2), n2, base, dim2,
                                      k, 0.3, None, _dev(d_scores), _dev(d_rows), _dev(d_out_ids), _dev(d_n))
        assert rc == L.DA_OK, L.last_error()
        torch.cuda.synchronize()
        assert torch.cuda.current_device() == prev
        o_rows, o_scores, _ = oracle.topk(m2, None, None, q2, k, 0.3)
        cnt = int(d_n.cpu()[0])
        assert cnt == len(o_rows)
        assert np.array_equal(d_rows.cpu().numpy().astype(np.uint32)[:cnt], o_rows + base)
        assert np.allclose(d_scores.cpu().numpy()[:cnt], o_scores, rtol=1e-5, atol=0)
        assert np.array_equal(d_out_ids.cpu().numpy().reshape(k, 16)[:cnt], ids2[o_rows])
        rc = lib.da_scan_topk_noalloc(0, _dev(t_m2), _dev(t_inv2), None, _dev(t_ids2), None, _dev(t_q2), n2, 0xFFFFFF00, dim2,
                                      k, 0.3, None, _dev(d_scores), _dev(d_rows), _dev(d_out_ids), _dev(d_n))
        assert rc == L.DA_ERR_CAPACITY and "row_base" in L.last_error()

        # requests beyond the fixed scratch are refused with a clear error and launch nothing
        rc = lib.da_scan_topk_noalloc(0, _dev(d_mat), _dev(d_inv), None, _dev(d_ids), None, _dev(d_q), n, 0, dim,
                                      L.DA_NOALLOC_MAX_K + 1, 0.7, None, _dev(d_scores), _dev(d_rows), _dev(d_out_ids),
                                      _dev(d_n))
        assert rc == L.DA_ERR_CAPACITY and "DA_NOALLOC_MAX_K" in L.last_error()
        rc = lib.da_scan_topk_noalloc(0, None, _dev(d_inv), None, _dev(d_ids), None, _dev(d_q), n, 0, dim, k, 0.7,
                                      None, _dev(d_scores), _dev(d_rows), _dev(d_out_ids), _dev(d_n))
        assert rc == L.DA_ERR_INVALID
