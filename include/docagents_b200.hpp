// docagents_b200.hpp -- header-only C++ mirror of the reference's `store.Store` interface for the
// one path this library replaces (internal/store/store.go:57-67): SaveChunks, SaveEmbeddings, TopK,
// with the reference's names, argument meaning and error behaviour, over the C ABI of
// docagents_b200.h.  It is what the Go decorator in INTEGRATION.md does, written in the language
// this image can compile (there is no Go toolchain here); tests/cpp/contract_test.cpp drives it.
#pragma once

#include "docagents_b200.h"

#include <array>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace docagents {

using UUID = std::array<uint8_t, 16>;              // google/uuid's [16]byte
using Vector = std::vector<float>;                 // embeddings.Vector (internal/embeddings/embeddings.go:4)

struct Chunk {                                     // store.go:30-36 (text lives in Postgres)
    UUID ID{};
    UUID DocumentID{};
    int Index = 0;
};
struct Embedding {                                 // store.go:44-48
    UUID ChunkID{};
    Vector Vec;
    std::string Model;
};
struct SearchResult {                              // store.go:50-54 (Chunk/Summary hydrated from Postgres by ChunkID)
    UUID ChunkID{};
    float Score = 0.f;
};

// The Go methods return `error`; here a failed call throws StoreError carrying da_last_error().
struct StoreError : std::runtime_error {
    int code;
    StoreError(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

class GpuStore {
public:
    struct Options {
        int dim = 3072;                            // vector(3072): postgres.go:84
        std::vector<int> devices{0};
        uint64_t capacity_rows = 1 << 16;
        double min_similarity = 0.7;               // postgres.go:223, a float64 constant; compared as (double)score >= floor
        uint32_t batch_window_us = 0;
    };

    explicit GpuStore(const Options &o)
    {
        da_config cfg{};
        cfg.struct_size = sizeof cfg;
        cfg.dim = o.dim;
        cfg.n_devices = (int32_t)o.devices.size();
        for (size_t i = 0; i < o.devices.size() && i < DA_MAX_DEVICES; i++) cfg.device_ids[i] = o.devices[i];
        cfg.capacity_rows = o.capacity_rows;
        cfg.min_similarity = (float)o.min_similarity;
        cfg.min_similarity_f64 = o.min_similarity;
        cfg.batch_window_us = o.batch_window_us;
        check(da_store_create(&cfg, &h_));
    }
    ~GpuStore() { da_store_destroy(h_); }
    GpuStore(const GpuStore &) = delete;
    GpuStore &operator=(const GpuStore &) = delete;

    // Store.SaveChunks (store.go:61): ids are assigned by the caller's relational store; the device
    // store learns chunk -> document.
    std::vector<Chunk> SaveChunks(const UUID &docID, std::vector<Chunk> chunks)
    {
        std::vector<uint8_t> ids(chunks.size() * 16);
        for (size_t i = 0; i < chunks.size(); i++) {
            chunks[i].DocumentID = docID;
            if (chunks[i].Index == 0) chunks[i].Index = (int)i;
            memcpy(&ids[16 * i], chunks[i].ID.data(), 16);
        }
        check(da_store_save_chunks(h_, docID.data(), ids.data(), chunks.size()));
        return chunks;
    }

    // Store.SaveEmbeddings (store.go:64; postgres.go:176-201): upsert by chunk id, empty slice is a no-op.
    void SaveEmbeddings(const std::vector<Embedding> &embs)
    {
        if (embs.empty()) return;
        const size_t dim = embs[0].Vec.size();
        std::vector<uint8_t> ids(embs.size() * 16);
        std::vector<float> vecs(embs.size() * dim);
        for (size_t i = 0; i < embs.size(); i++) {
            if (embs[i].Vec.size() != dim) throw StoreError(DA_ERR_DIM, "embeddings of different dimensions in one batch");
            memcpy(&ids[16 * i], embs[i].ChunkID.data(), 16);
            memcpy(&vecs[dim * i], embs[i].Vec.data(), dim * sizeof(float));
        }
        check(da_store_save_embeddings(h_, ids.data(), vecs.data(), embs.size(), (int32_t)dim));
    }

    // Resident embeddings per document (0 = never given to this store): the reader-side check of the Go decorator
    // (integration/gpustore.go: ensureResident) before it searches documents another process embedded.
    std::vector<uint32_t> DocRows(const std::vector<UUID> &docIDs)
    {
        std::vector<uint8_t> docs(docIDs.size() * 16);
        for (size_t i = 0; i < docIDs.size(); i++) memcpy(&docs[16 * i], docIDs[i].data(), 16);
        std::vector<uint32_t> out(docIDs.size());
        check(da_store_doc_rows(h_, docs.data(), docIDs.size(), out.data()));
        return out;
    }

    // Store.TopK (store.go:66; postgres.go:218-285).  An empty result is the reference's nil slice.  An empty
    // query vector is an error (the reference sends '[]' to pgvector, which rejects it); an empty document list
    // matches nothing (document_id = ANY('{}'), postgres.go:239).
    std::vector<SearchResult> TopK(const std::vector<UUID> &docIDs, const Vector &vector, int k)
    {
        if (vector.empty()) throw StoreError(DA_ERR_DIM, "vector must have at least 1 dimension");
        if (docIDs.empty()) return {};
        std::vector<uint8_t> docs(docIDs.size() * 16);
        for (size_t i = 0; i < docIDs.size(); i++) memcpy(&docs[16 * i], docIDs[i].data(), 16);
        const int kk = k > 0 ? k : 1;
        std::vector<uint8_t> out_ids((size_t)kk * 16);
        std::vector<float> out_scores(kk);
        int32_t n = 0;
        check(da_store_topk(h_, docs.empty() ? nullptr : docs.data(), docIDs.size(), vector.data(), (int32_t)vector.size(), k,
                            out_ids.data(), out_scores.data(), &n));
        std::vector<SearchResult> res(n);
        for (int i = 0; i < n; i++) {
            memcpy(res[i].ChunkID.data(), &out_ids[16 * i], 16);
            res[i].Score = out_scores[i];
        }
        return res;
    }

    da_store *handle() const { return h_; }

private:
    static void check(int rc)
    {
        if (rc != DA_OK) throw StoreError(rc, da_last_error());
    }
    da_store *h_ = nullptr;
};

// cmd/query/main.go:161-170: float32 running sum over the results, divided by n.
inline float calculateAvgSimilarity(const std::vector<SearchResult> &results)
{
    if (results.empty()) return 0.0f;
    float total = 0.0f;
    for (const SearchResult &r : results) total += r.Score;
    return total / (float)results.size();
}

}  // namespace docagents
