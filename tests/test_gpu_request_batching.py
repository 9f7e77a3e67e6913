"""GPU test of request batching: concurrent da_store_topk calls, as the query agent
issues them from one goroutine per HTTP request (cmd/query/main.go:34), are merged into batched passes
-- each call keeping its own document set -- with results identical to unbatched calls."""
import threading
import uuid

import numpy as np
import pytest

import oracle
from doc_agents_b200 import GpuStore


@pytest.fixture(autouse=True, scope="module")
def _built(built_lib):
    """Compile the CUDA library once if it is stale (the GPU box normally receives it pre-built)."""
    return built_lib


pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("sidecar", [False, True], ids=["fp32", "int8_sidecar"])
def test_concurrent_topk_calls_are_batched_and_results_unchanged(sidecar):
    """`sidecar`: the batched store also keeps the int8 copy, so merged calls (floor 0.7) ride the kind::i8 contraction
    while the plain store answers each call with the fp32 scan: the results must still be identical."""
    S = oracle.Synth(n_clusters=8)
    n, dim = 60_000, 1536
    docs_a = [uuid.UUID(bytes=bytes(b)) for b in oracle.doc_ids(S, 0, n // 200)[::2]]
    docs_b = [uuid.UUID(bytes=bytes(b)) for b in oracle.doc_ids(S, 0, n // 200)[1::2]]
    n_threads = 48
    qs = [oracle.gen_query(S, i, i % 8, dim) for i in range(n_threads)]
    ks = [5 if i % 3 else 10 for i in range(n_threads)]
    with GpuStore(dim=dim, capacity_rows=n) as plain, \
            GpuStore(dim=dim, capacity_rows=n, batch_window_us=20_000, batch_max=32, q8_sidecar=sidecar) as batched:
        plain.generate_synthetic(S, 0, n)
        batched.generate_synthetic(S, 0, n)
        want = [plain.TopK(docs_a if i % 2 == 0 else docs_b, qs[i], ks[i]) for i in range(n_threads)]
        got = [None] * n_threads
        errors = []
        start = threading.Barrier(n_threads)

        def worker(i):
            try:
                start.wait()
                got[i] = batched.TopK(docs_a if i % 2 == 0 else docs_b, qs[i], ks[i])
            except Exception as e:      # noqa: BLE001
                errors.append(e)

        threads = [threading.Thread(target=worker, args=(i,)) for i in range(n_threads)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=120)
        assert not errors, errors
        for i in range(n_threads):
            assert (got[i] is None) == (want[i] is None)
            if want[i] is None:
                continue
            assert [r.Chunk.ID for r in got[i]] == [r.Chunk.ID for r in want[i]], i
            assert [r.Score for r in got[i]] == [r.Score for r in want[i]], i
            assert len(got[i]) <= ks[i]
        s = batched.stats()
        assert s.topk_calls == n_threads
        assert s.topk_passes <= n_threads // 8, (s.topk_calls, s.topk_passes)       # calls with DIFFERENT document sets merged too
        # errors still reach the caller that caused them
        with pytest.raises(Exception):
            batched.TopK(docs_a, qs[0][:100], 5)
        assert batched.TopK([], qs[0], 5) is None
        sp = plain.stats()
        assert sp.topk_calls == sp.topk_passes == n_threads
        # a search that reads little does not wait for the window (20 ms here): it goes at once, on its own slot
        import time
        few = docs_a[:3]
        batched.TopK(few, qs[1], 5)
        t0 = time.perf_counter()
        light = batched.TopK(few, qs[1], 5)
        dt = time.perf_counter() - t0
        assert dt < 0.010, dt
        ref = plain.TopK(few, qs[1], 5)
        assert (light is None) == (ref is None)
        if ref is not None:
            assert [r.Chunk.ID for r in light] == [r.Chunk.ID for r in ref] and [r.Score for r in light] == [r.Score for r in ref]
        s2 = batched.stats()
        assert s2.topk_calls - s.topk_calls == s2.topk_passes - s.topk_passes
