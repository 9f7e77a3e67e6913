"""N>1 path.  CPU (gloo, world_size 2): the host-side sharding and merge logic that bench.py and the
device merge implement.  GPU (>= 2 devices): the real thing under torchrun, via tests/mr_check.py."""
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
from bench import merge_candidates, shard_range
from gpu_util import device_count

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions_on_document_boundaries():
    for n, world, cpd in ((10_000_000, 8, 200), (10_000_000, 3, 200), (30_000, 2, 200), (1000, 4, 200), (100, 8, 200)):
        spans = [shard_range(r, world, n, cpd) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        for (a, b), (c, d) in zip(spans, spans[1:]):
            assert b == c and a <= b
        for a, b in spans[:-1]:
            assert (a % cpd == 0 or a == n) and (b % cpd == 0 or b == n)       # a document never straddles two GPUs


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        S = oracle.Synth(n_clusters=4)
        n, dim, k = 4000, 256, 10
        lo, hi = shard_range(rank, world, n, S.chunks_per_doc)
        query = oracle.gen_query(S, 2, 1, dim)
        local = oracle.gen_rows(S, lo, hi - lo, dim)
        rows, scores, _ = oracle.topk(local, None, None, query, k)            # the rank's local top-k
        gathered = [None] * world
        dist.all_gather_object(gathered, (scores, rows.astype(np.int64) + lo))   # the all-gather step
        m_scores, m_rows = merge_candidates(gathered, k)                      # the merge step
        whole = oracle.gen_rows(S, 0, n, dim)
        o_rows, o_scores, _ = oracle.topk(whole, None, None, query, k)
        ok = np.array_equal(m_rows, o_rows) and np.array_equal(m_scores, o_scores)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_local_topk_allgather_merge_equals_global_topk_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 400
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


def test_merge_candidates_tie_rule():
    s, r = merge_candidates([(np.array([0.9, 0.8], np.float32), [5, 9]), (np.array([0.9, 0.85], np.float32), [2, 7])], 3)
    assert list(r) == [2, 5, 7] and list(s) == [np.float32(0.9), np.float32(0.9), np.float32(0.85)]


@pytest.mark.gpu
@pytest.mark.skipif(device_count() < 2, reason="needs 2 GPUs")
def test_two_ranks_under_torchrun_match_the_oracle(built_lib):      # the library is built once, before the ranks start
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "mr_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "mr_check ok" in out.stdout
