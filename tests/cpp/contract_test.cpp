// Replays, in C++ through include/docagents_b200.hpp, what the reference's own tests pin about this path
// with its mock store (cmd/query/main_test.go:72-79 TopK receives ids and k, :107-120 default k = 5,
// :190-204 a TopK error is an error, :226-254 empty result; cmd/analysis/main_test.go:76-78 SaveEmbeddings
// once with all chunks, :231 with an empty slice, :204-205 error propagates) -- against the real library
// and the CPU oracle (oracle/liboracle.so, loaded here as the checker).
//   g++ -std=c++17 -I include tests/cpp/contract_test.cpp -L doc_agents_b200 -ldocagents_b200 -L oracle -loracle
#include "docagents_b200.hpp"

#include <cmath>
#include <cstdio>
#include <cstdlib>

extern "C" {
struct orc_synth { uint64_t seed_data, seed_centroid, seed_query, seed_ids; uint32_t chunks_per_doc, n_clusters; int32_t cmul, nmul;
                   int32_t smul; uint32_t super_group; };
void orc_gen_rows(const orc_synth *, uint64_t, uint64_t, int, float *);
void orc_gen_query(const orc_synth *, uint64_t, uint32_t, int, float *);
void orc_chunk_id(const orc_synth *, uint64_t, uint8_t *);
void orc_doc_id(const orc_synth *, uint64_t, uint8_t *);
int orc_topk(const float *, uint64_t, int, const uint32_t *, const uint8_t *, const float *, int, double, int, int,
             uint32_t *, float *, double *);
float orc_avg_similarity(const float *, int);
}

#define EXPECT(cond)                                                                   \
    do {                                                                               \
        if (!(cond)) { fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); return 1; } \
    } while (0)

using namespace docagents;

int main()
{
    const orc_synth S{0xD0C5EED, 0xC0FFEE, 0xBEEF, 0x1D5EED, 200, 1024, 2, 1, 0, 0};
    const int n = 200, dim = 3072;                          // BASELINE configs[0]
    std::vector<float> mat((size_t)n * dim), q(dim);
    orc_gen_rows(&S, 0, n, dim, mat.data());
    orc_gen_query(&S, 0, 0, dim, q.data());
    UUID doc{}, other{};
    orc_doc_id(&S, 0, doc.data());
    orc_doc_id(&S, 999, other.data());

    GpuStore::Options o;
    o.dim = dim; o.capacity_rows = 512;
    GpuStore st(o);

    std::vector<Chunk> chunks(n);
    for (int i = 0; i < n; i++) orc_chunk_id(&S, i, chunks[i].ID.data());
    chunks = st.SaveChunks(doc, chunks);
    EXPECT(chunks.size() == (size_t)n && chunks[7].DocumentID == doc);

    st.SaveEmbeddings({});                                   // empty slice: no-op, no error
    EXPECT(st.TopK({doc}, q, 5).empty());                    // nothing stored: nil slice, no error
    // what the Go decorator asks before it searches: a document this process was never given has no rows
    // (it was embedded by the analysis agent, another process) and is loaded from Postgres first
    EXPECT(st.DocRows({doc, other}) == (std::vector<uint32_t>{0, 0}));

    std::vector<Embedding> embs(n);
    for (int i = 0; i < n; i++) {
        embs[i].ChunkID = chunks[i].ID;
        embs[i].Vec.assign(mat.begin() + (size_t)i * dim, mat.begin() + (size_t)(i + 1) * dim);
        embs[i].Model = "text-embedding-3-large";
    }
    st.SaveEmbeddings(embs);                                 // once, with all chunks of the document

    EXPECT(st.DocRows({other, doc, doc}) == (std::vector<uint32_t>{0, (uint32_t)n, (uint32_t)n}));

    uint32_t o_rows[20]; float o_scores[20]; double o_dist[20];
    const int on = orc_topk(mat.data(), n, dim, nullptr, nullptr, q.data(), 5, 0.7, 0, 1, o_rows, o_scores, o_dist);
    std::vector<SearchResult> res = st.TopK({doc}, q, 5);    // the handler's default top_k
    EXPECT(on == 5 && res.size() == 5);
    for (int i = 0; i < 5; i++) {
        EXPECT(res[i].ChunkID == chunks[o_rows[i]].ID);
        EXPECT(std::fabs(res[i].Score - o_scores[i]) <= 1e-5f * std::fabs(o_scores[i]));
    }
    float sc[5];
    for (int i = 0; i < 5; i++) sc[i] = res[i].Score;
    EXPECT(calculateAvgSimilarity(res) == orc_avg_similarity(sc, 5));
    EXPECT(calculateAvgSimilarity({}) == 0.0f);

    EXPECT(st.TopK({other}, q, 5).empty());                  // unknown document: nil, no error
    EXPECT(st.TopK({}, q, 5).empty());                       // ANY('{}')
    EXPECT(st.TopK({other, doc, doc}, q, 20).size() == 20);  // API cap 20, duplicate ids harmless

    EXPECT(st.TopK({doc}, q, 0).empty());                    // LIMIT 0
    bool threw = false;
    try { st.TopK({doc}, Vector{}, 5); } catch (const StoreError &e) { threw = e.code == DA_ERR_DIM; }
    EXPECT(threw);                                           // the reference sends '[]', which pgvector rejects: an error, not nil
    threw = false;
    try { st.TopK({doc}, q, -1); } catch (const StoreError &e) { threw = e.code == DA_ERR_INVALID; }
    EXPECT(threw);                                           // LIMIT must not be negative
    threw = false;
    try { st.TopK({doc}, Vector(q.begin(), q.begin() + 1536), 5); } catch (const StoreError &e) { threw = e.code == DA_ERR_DIM; }
    EXPECT(threw);                                           // dimension mismatch -> error (-> HTTP 500)
    threw = false;
    Embedding stray;
    orc_chunk_id(&S, 123456, stray.ChunkID.data());
    stray.Vec = q;
    try { st.SaveEmbeddings({stray}); } catch (const StoreError &e) { threw = e.code == DA_ERR_UNKNOWN_CHUNK; }
    EXPECT(threw);                                           // FK embeddings.chunk_id -> chunks.id

    // upsert: overwrite the best row with the worst row's vector; it must leave the top-5
    Embedding up = embs[o_rows[0]];
    up.Vec = embs[(o_rows[0] + 1) % n].Vec;
    st.SaveEmbeddings({up});
    res = st.TopK({doc}, q, 5);
    EXPECT(res.size() == 5 && res[0].ChunkID == chunks[o_rows[1]].ID);
    EXPECT(st.DocRows({doc})[0] == (uint32_t)n);             // an overwrite adds no row
    printf("contract_test ok\n");
    return 0;
}
