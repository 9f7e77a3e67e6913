"""doc_agents_b200 -- B200-native replacement for ONE path of tomerlieber/doc-agents:
the Query Agent's semantic search (`Store.TopK`, internal/store/store.go:66) and its write
side (`Store.SaveEmbeddings`, :64).  The product is csrc/ (CUDA for sm_100a behind the C ABI
of include/docagents_b200.h); this package is the thin host-side mirror of the reference's
store interface used by tests and bench.  No PyTorch, no CPU fallback on the search path.
"""
from .store import (Chunk, Embedding, GpuStore, SearchResult, StoreError, Summary,  # noqa: F401
                    calculate_avg_similarity)
from . import _lib  # noqa: F401

__all__ = ["Chunk", "Embedding", "GpuStore", "SearchResult", "StoreError", "Summary",
           "calculate_avg_similarity"]
