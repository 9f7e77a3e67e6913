"""Multi-process parity check, launched by torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tests/mr_check.py

Each rank holds a row shard of one synthetic corpus, every rank runs the same queries through the
C ABI (local fused scan, NCCL all-gather, merge), and every rank checks the merged answer against
the CPU oracle on the WHOLE corpus."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist

    import oracle
    from bench import shard_range
    from doc_agents_b200 import GpuStore

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idt = torch.tensor(list(GpuStore.nccl_unique_id()), dtype=torch.uint8, device="cuda")
    dist.broadcast(idt, 0)

    S = oracle.Synth(n_clusters=8)
    n, dim, k = 30_000, 3072, 10
    lo, hi = shard_range(rank, world, n, S.chunks_per_doc)
    st = GpuStore(dim=dim, capacity_rows=hi - lo, devices=(local,), world_size=world, rank=rank,
                  nccl_id=bytes(idt.cpu().tolist()))
    st.generate_synthetic(S, lo, hi - lo)
    mat = oracle.gen_rows(S, 0, n, dim)
    rd = oracle.row_docs(S, 0, n)
    docs = oracle.doc_ids(S, 0, n // S.chunks_per_doc)
    chunk = oracle.chunk_ids(S, 0, n)
    qs = np.stack([oracle.gen_query(S, i, i % 8, dim) for i in range(9)])

    picked = list(range(0, n // S.chunks_per_doc, 3))
    allowed = np.zeros(n // S.chunks_per_doc, np.uint8)
    allowed[picked] = 1
    for kw, al in ((dict(all_docs=True), None), (dict(doc_ids=docs[picked]), allowed)):
        ids, sc, rows = st.search(qs, k, **kw)                      # batch of 9 => two passes (8 + 1)
        for i in range(len(qs)):
            o_rows, o_sc, _ = oracle.topk(mat, rd, al, qs[i], k)
            assert np.array_equal(rows[i], o_rows), (rank, i, rows[i], o_rows)
            assert np.allclose(sc[i], o_sc, rtol=1e-5), (rank, i)
            assert np.array_equal(ids[i], chunk[o_rows]), (rank, i)
        one = st.search(qs[3], k, **kw)
        assert np.array_equal(one[2][0], rows[3])
    # a batch large enough for the tensor-core path: candidates per rank, all-gather, merge
    qs40 = np.stack([oracle.gen_query(S, 100 + i, i % 8, dim) for i in range(40)])
    ids, sc, rows = st.search(qs40, k, all_docs=True)
    for i in (0, 13, 39):
        o_rows, o_sc, _ = oracle.topk(mat, rd, None, qs40[i], k)
        assert np.array_equal(rows[i], o_rows), (rank, i, rows[i], o_rows)
        assert np.allclose(sc[i], o_sc, rtol=1e-5), (rank, i)
    # IVFFlat across ranks: k-means is data-parallel (NCCL broadcast of the seeds, all-reduce of the sums),
    # so every rank holds the same centroids; probing every list equals the exact scan
    st.build_ivf(lists=24, iters=5, seed=7)
    _, _, cent = st.debug_ivf(0)
    digest = torch.tensor([float(np.abs(cent).sum()), float(cent[3, 5])], dtype=torch.float64, device="cuda")
    gathered = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(gathered, digest)
    assert all(torch.equal(g, gathered[0]) for g in gathered), "centroids differ between ranks"
    for i in range(4):
        exact = st.search(qs[i], k, all_docs=True)
        ivf_all = st.search(qs[i], k, all_docs=True, ivf=True, probes=24)
        assert np.array_equal(exact[2][0], ivf_all[2][0]) and np.array_equal(exact[0][0], ivf_all[0][0])
        ivf_some = st.search(qs[i], k, all_docs=True, ivf=True, probes=8)
        assert len(np.intersect1d(ivf_some[2][0], exact[2][0])) >= k - 1
    # an unknown document everywhere => nil on every rank, collective still matched
    assert len(st.search(qs[0], k, doc_ids=oracle.doc_ids(S, 10_000, 1))[2][0]) == 0
    ms = st.bench_resident(qs[0], k, 5, all_docs=True)
    assert ms > 0
    st.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print(f"mr_check ok: world={world}, merged top-{k} equals the oracle on the whole corpus")


if __name__ == "__main__":
    main()
