/*
 * docagents_b200.h -- C ABI of the B200-native semantic-search store.
 *
 * One shared library (libdocagents_b200.so: CUDA runtime + NCCL, no PyTorch,
 * no CPU fallback) replaces ONE path of tomerlieber/doc-agents: the Query
 * Agent's `Store.TopK` and its write side `Store.SaveEmbeddings`.  Each entry
 * point below names the reference interface it stands in for (paths relative
 * to the reference tree).  The cgo binding a maintainer adds on the Go side is
 * in INTEGRATION.md.
 *
 * Conventions
 *   - every function returns 0 on success or a negative da_status; the message
 *     for the calling thread's most recent failure is da_last_error().  This is
 *     the channel that becomes a Go `error` (TopK errors surface as HTTP 500:
 *     cmd/query/main.go:102-105).
 *   - the caller owns every in/out buffer; nothing is retained after return
 *     (cgo pointer rules); the library owns all device memory.
 *   - thread-safe: any OS thread may call any function on a store at any time
 *     (one goroutine per request: cmd/query/main.go:34); calls on one store
 *     are serialised internally.  The CUDA device is selected per call.
 *   - ids are 16 raw bytes (google/uuid's [16]byte: internal/store/store.go:23-48).
 */
#ifndef DOCAGENTS_B200_H
#define DOCAGENTS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DA_ABI_VERSION 1
#define DA_MAX_DEVICES 8
#define DA_MAX_K 128            /* reference API caps top_k at 20 (cmd/query/main.go:23) */
#define DA_MAX_BATCH 4096
#define DA_NCCL_ID_BYTES 128
#define DA_DEFAULT_MIN_SIMILARITY 0.7f   /* internal/store/postgres.go:223 */

typedef enum {
    DA_OK = 0,
    DA_ERR_INVALID = -1,        /* bad argument (NULL, k<0, ...) */
    DA_ERR_DIM = -2,            /* vector length != store dimension (pgvector: "expected N dimensions") */
    DA_ERR_UNKNOWN_CHUNK = -3,  /* embedding for a chunk never announced by save_chunks
                                   (the FK embeddings.chunk_id -> chunks.id: postgres.go:82-83) */
    DA_ERR_CAPACITY = -4,       /* store full / not enough HBM; there is no CPU fallback */
    DA_ERR_CUDA = -5,
    DA_ERR_NCCL = -6,
    DA_ERR_STATE = -7,          /* e.g. IVF search before the index was built */
    DA_ERR_NONFINITE = -8       /* NaN/Inf element (pgvector rejects them) */
} da_status;

typedef struct da_store da_store;

/* Provider configuration: what `buildStore` (internal/app/deps.go:179-202) would pass for
 * STORE_PROVIDER=gpu, from env knobs in the style of internal/config/config.go:11-42. */
typedef struct {
    uint32_t struct_size;           /* sizeof(da_config) */
    int32_t  dim;                   /* 3072 (vector(3072): postgres.go:84) or 1536 */
    int32_t  n_devices;             /* GPUs driven by THIS process (1..8) */
    int32_t  device_ids[DA_MAX_DEVICES];
    uint64_t capacity_rows;         /* per device; HBM is reserved up front */
    float    min_similarity;        /* floor used by da_store_topk; 0.7 for reference parity */
    /* Multi-process row sharding (one process per GPU, e.g. under torchrun):
     * world_size processes, each with n_devices local shards.  world_size<=1 => single process. */
    int32_t  world_size;
    int32_t  rank;
    uint8_t  nccl_id[DA_NCCL_ID_BYTES];   /* from da_nccl_unique_id on rank 0, same on all ranks */
    uint32_t flags;                 /* DA_STORE_* below */
    /* Request batching for da_store_topk (0 = off).  Concurrent calls -- one goroutine per
     * /api/query, cmd/query/main.go:34 -- that arrive within batch_window_us of each other
     * are answered by ONE batched pass over the corpus (at most batch_max, default 64, calls), each call keeping
     * its own document set. */
    uint32_t batch_window_us;
    uint32_t batch_max;
    /* The floor as the float8 it is in the reference (`const minSimilarity = 0.7`, a float64 parameter compared with a
     * float8 expression: postgres.go:223,240).  Non-zero: rows qualify iff (double)score >= min_similarity_f64 -- exactly
     * Postgres' comparison -- and min_similarity above is ignored.  Zero: the float field is used as it is (note that
     * (float)0.7 = 0.69999999 lets a row with that score in, which the SQL would leave out). */
    double   min_similarity_f64;
} da_config;

/* da_config.flags */
#define DA_STORE_SPLIT16 0x1u   /* split-plane layout: every fp32 element is kept as its upper and lower
                                   16 bits in two planes -- lossless, same HBM footprint.  The scan streams only the upper
                                   plane (half the bytes per pass) to generate candidates, which are re-scored from both
                                   planes with the fp32 scan's arithmetic: results are bit-identical to the default layout.
                                   Batches of 16+ queries contract the upper planes as bf16 on the tensor cores.
                                   dim must be a multiple of 16; the IVFFlat path (TMA over fp32 rows) is not available. */
#define DA_STORE_Q8_SIDECAR 0x2u /* int8 sidecar: next to the fp32 rows the store keeps every row as dim int8
                                   values with a per-row scale and the row's own quantisation-error bound (dim + 8 more bytes
                                   per row).  Searches of one or two queries scan the int8 rows (a quarter of the bytes),
                                   keep every row whose score interval can still reach the k-th best, and re-score those
                                   from the fp32 rows with the fp32 scan's arithmetic; batches with a similarity floor of
                                   at least 0.5 contract the int8 rows on the tensor cores (kind::i8) for their candidates.
                                   Results are bit-identical to the default layout.  dim must be a multiple of 32; excludes DA_STORE_SPLIT16. */

/* Fills `out` with a fresh ncclUniqueId to hand to every rank's da_config. */
int da_nccl_unique_id(uint8_t out[DA_NCCL_ID_BYTES]);

/* store.NewPostgres (internal/store/postgres.go:23-33): open the provider. */
int da_store_create(const da_config *cfg, da_store **out);
void da_store_destroy(da_store *st);

/* Store.SaveChunks (internal/store/store.go:61; postgres.go:142-164) as seen by the device
 * store: it only needs chunk -> document, because store.Embedding carries no document id
 * (store.go:44-48) and TopK filters on chunks.document_id (postgres.go:237-239).
 * chunk_ids: n x 16 bytes.  Re-announcing a chunk under another document moves it. */
int da_store_save_chunks(da_store *st, const uint8_t doc_id[16], const uint8_t *chunk_ids, size_t n);

/* Store.SaveEmbeddings (store.go:64; postgres.go:176-201): upsert keyed by chunk_id
 * (ON CONFLICT (chunk_id) DO UPDATE, :197).  n == 0 is a no-op (:177-179).
 * vecs: n x dim row-major float32, the exact bits of embeddings.Vector
 * (vectorToString is lossless: postgres.go:321-330). */
int da_store_save_embeddings(da_store *st, const uint8_t *chunk_ids, const float *vecs,
                             size_t n, int32_t dim);

/* Cold start: the rows of `embeddings JOIN chunks` streamed in binary (what a COPY of
 * chunk_id, document_id, vector delivers), in any order.  Equivalent to SaveChunks + SaveEmbeddings per
 * document, in one call: doc_ids and chunk_ids are n x 16 bytes, vecs n x dim. */
int da_store_bulk_load(da_store *st, const uint8_t *chunk_ids, const uint8_t *doc_ids, const float *vecs,
                       size_t n, int32_t dim);

/* Store.TopK (store.go:66; postgres.go:218-285).  Rows whose chunk belongs to one of
 * doc_ids (n_docs x 16 bytes; n_docs == 0 matches nothing, :239), cosine similarity
 * 1 - (vector <=> query) >= the store's min_similarity (:223,:240), best first (:241),
 * at most k (:242).  out_chunk_ids: k x 16 bytes, out_scores: k floats (Score float32,
 * :259), *out_n: rows returned (0 <=> the reference's nil slice, :250). */
int da_store_topk(da_store *st, const uint8_t *doc_ids, size_t n_docs, const float *query,
                  int32_t dim, int32_t k, uint8_t *out_chunk_ids, float *out_scores, int32_t *out_n);

/* One device store, two processes.  In the reference the writer and the reader of this path are
 * different processes: the analysis agent calls SaveEmbeddings (cmd/analysis/main.go:106), the query agent calls
 * TopK (cmd/query/main.go:101), both against one Postgres (docker-compose.yml:87-128).  Here both name one store:
 *   da_store_export   the process that owns the store (single GPU, default fp32 layout) publishes it under a POSIX
 *                     shared-memory name ("/docagents-gpu0"): CUDA IPC handles of the resident arrays, a log of the
 *                     documents and row runs it holds, a process-shared lock.  It keeps saving as before.
 *   da_store_attach   another process on the same GPU maps the same HBM read-only.  Every search first replays what
 *                     the exporter has appended since, so a document saved there is found by the next TopK here.
 *                     Saving, index building and synthetic fills are refused (DA_ERR_STATE) on an attached store.
 * A search in the attached process and a save in the exporter exclude each other (the shared lock), so a reader never
 * sees a half-written upsert.  Destroying the exporter's store removes the name; attached stores must be destroyed first. */
int da_store_export(da_store *st, const char *shm_name);
int da_store_attach(const char *shm_name, da_store **out);

/* How many embeddings of each of n_docs documents are resident (0 = the store has never been given the document).
 * No SQL counterpart; it is what lets a reader process notice that a document named by a request
 * (document_ids is required: cmd/query/main.go:20-24) was embedded by ANOTHER process (the analysis agent:
 * docker-compose.yml:87-128) and load it from Postgres before searching.  Host-side lookup, no GPU work. */
int da_store_doc_rows(da_store *st, const uint8_t *doc_ids, size_t n_docs, uint32_t *out_rows);

/* Flags of da_search_params. */
#define DA_SEARCH_ALL_DOCS      0x1u  /* ignore doc_ids: every document is requested (bench roofline runs) */
#define DA_SEARCH_READ_ALL_ROWS 0x2u  /* doc mask as a predicate: read every row even if its document is
                                         not requested (configs[1] "HBM-bound" measurement); default skips them */
#define DA_SEARCH_IVF           0x4u  /* probe the IVFFlat lists instead of the exact scan */
#define DA_SEARCH_NO_TENSOR     0x10u /* keep a batch on the fp32 SIMT scan (default from 4 queries up -- 3 with the int8 sidecar, 16 in the
                                         split-plane layout --: tcgen05 TF32
                                         candidate generation + exact fp32 re-score; results are identical) */
#define DA_SEARCH_NO_COMPACT    0x40u /* a filter naming at most 1/16 of the documents is served by compacting the requested rows
                                         into a list and scanning that; this flag keeps the row loop with the bitmap (comparisons) */
#define DA_SEARCH_NO_Q8         0x20u /* DA_STORE_Q8_SIDECAR stores: scan the fp32 rows for this search (comparisons) */

/* Same query as da_store_topk with every knob exposed, batched: `batch` queries share one doc filter
 * (one /api/query has one document set: cmd/query/main.go:20-24) or carry one each (doc_offsets). */
typedef struct {
    uint32_t struct_size;
    uint32_t flags;
    const uint8_t *doc_ids;  size_t n_docs;
    const float *queries;    int32_t batch;   int32_t dim;    /* batch x dim */
    int32_t k;
    float   min_similarity;  /* use -INFINITY for "no floor" */
    int32_t ivf_probes;      /* DA_SEARCH_IVF only; 0 = store default */
    /* Optional per-query document sets (what merged /api/query requests need: each request names its own
     * documents): query i asks for doc_ids[doc_offsets[i] .. doc_offsets[i+1]), doc_offsets has batch+1
     * entries and n_docs == doc_offsets[batch].  NULL: one set shared by the whole batch. */
    const uint64_t *doc_offsets;
    double min_similarity_f64;   /* non-zero: the floor as a double, compared as (double)score >= it (see da_config); else min_similarity */
} da_search_params;

/* Outputs are batch x k (row-major); out_n is batch counts; out_rows (optional, may be NULL)
 * receives the global insertion row of each hit. */
int da_store_search(da_store *st, const da_search_params *p, uint8_t *out_chunk_ids,
                    float *out_scores, uint32_t *out_rows, int32_t *out_n);

/* The IVFFlat index of the reference DDL (postgres.go:95-99: lists = 100, vector_cosine_ops):
 * spherical k-means on a sample, every row assigned to its nearest centroid, rows regrouped
 * by list.  Searches with DA_SEARCH_IVF rank the centroids and scan `probes` lists: a batch of 8 or more queries
 * ranks them with one tensor-core GEMM (queries x centroids^T, TF32 -- it only steers which lists are probed),
 * smaller batches with the fp32 scan of the centroid matrix; then ONE scan launch serves the whole batch, every
 * query walking its own lists.  Returned scores always come from the fp32 rows. */
typedef struct {
    uint32_t struct_size;
    int32_t  lists;          /* 100 in the reference DDL; 4096 in BASELINE configs[4] */
    int32_t  iters;          /* k-means iterations (<= 20) */
    uint64_t sample_rows;    /* 0 = max(lists*50, 10000), pgvector's rule */
    uint64_t seed;
    int32_t  default_probes; /* pgvector default is 1 (ivfflat.probes) */
    uint32_t flags;          /* DA_IVF_* below */
} da_ivf_params;
#define DA_IVF_SIMT_ASSIGN 0x1u  /* assign with the fp32 SIMT GEMM instead of tcgen05 TF32 */
#define DA_IVF_NO_SPLIT    0x2u  /* skip the balancing rounds after k-means (heavy lists split into empty ones); comparisons */
int da_store_build_ivf(da_store *st, const da_ivf_params *p);

/* Device-side synthetic corpus (bench/tests only; SURVEY.md section 8d, DESIGN.md
 * "Synthetic corpus").  Fills global rows [row_lo, row_lo + n) -- split evenly over this
 * store's shards -- bit-identically to the CPU oracle's generator, and registers their
 * documents.  No counterpart in the reference (its data comes from OpenAI embeddings). */
typedef struct {
    uint32_t struct_size;
    uint64_t seed_data, seed_centroid, seed_ids;
    uint32_t chunks_per_doc, n_clusters;
    int32_t  cmul, nmul;
    uint64_t row_lo, n_rows;
    /* Optional second level (0 = off, the BASELINE corpus): clusters are grouped `super_group` at a time and every
     * vector also carries smul x the direction of its cluster's group, so clusters of one group overlap (cosine
     * smul^2 / (smul^2 + cmul^2 + nmul^2) between them) -- a corpus on which an approximate index can lose recall. */
    int32_t  smul;
    uint32_t super_group;
} da_synth_params;
int da_store_generate_synthetic(da_store *st, const da_synth_params *p);

/* Host-side pieces of the same synthetic corpus, so benches need nothing but this library: query number `qi`
 * aimed at `cluster` (noise stream seeded by seed_query), and the 16-byte ids of chunks (kind 0, by global row)
 * or documents (kind 1, by document index).  p->row_lo / n_rows are ignored.  No GPU involved. */
int da_synth_query(const da_synth_params *p, uint64_t seed_query, uint64_t qi, uint32_t cluster, int32_t dim, float *out);
int da_synth_ids(const da_synth_params *p, int32_t kind, uint64_t first, uint64_t n, uint8_t *out);

typedef struct {
    uint32_t struct_size;
    int32_t  dim, n_devices, world_size, rank;
    uint64_t rows;                  /* resident on this process's shards */
    uint64_t rows_per_device[DA_MAX_DEVICES];
    uint64_t hbm_bytes;             /* reserved by this store */
    uint64_t n_docs;
    /* last search, this process: */
    float    last_scan_ms;          /* CUDA-event time of the scan kernel(s) on shard 0's stream */
    float    last_search_ms;        /* events around H2D + kernels + collective + D2H */
    uint64_t last_rows_read;        /* rows whose vectors were actually read */
    uint64_t last_bytes_algorithmic;/* rows_scanned*(4*dim+8) or rows_read*(4*dim+4)+4*rows: DESIGN.md */
    uint64_t kernel_launches;       /* cumulative count of this library's kernel launches */
    int32_t  ivf_lists;             /* 0 = no index */
    int32_t  scan_grid;             /* CTAs of the last scan launch */
    double   scan_ms_sum;           /* cumulative CUDA-event time of the searches' scan stages on shard 0 (all passes of a search) */
    uint64_t scan_launches;         /* how many searches scan_ms_sum covers (last_bytes_algorithmic is per search too) */
    /* last da_store_build_ivf, CUDA events on shard 0's stream: */
    float    ivf_build_ms;          /* whole build */
    float    ivf_kmeans_ms;         /* all k-means iterations (assignment GEMM + accumulate + finalize + all-reduce) */
    float    ivf_assign_ms;         /* final assignment GEMM over every resident row of shard 0 */
    float    ivf_assign_tflops;     /* 2 * rows * lists * dim / ivf_assign_ms */
    int32_t  peer_gather;           /* 1: shards deliver their records into shard 0 over NVLink (no collective); 2: one process per
                                       GPU and the scan's last CTA exchanges with every other GPU and merges (no collective, no merge kernel) */
    uint64_t tensor_overflows;      /* searches answered by the fp32 SIMT scan because a candidate list (tensor path, int8 sidecar) overflowed */
    uint64_t topk_calls;            /* da_store_topk calls answered */
    uint64_t topk_passes;           /* searches run for them (== topk_calls unless request batching merged some) */
    uint64_t small_searches;        /* searches answered by the one-copy / one-kernel / polled-result path (one query, one shard);
                                       for those last_search_ms is a host clock around copy + scan + pick-up */
    uint64_t last_bench_launches;   /* kernels of this library launched inside the timed region of the last da_store_bench_resident */
} da_stats;
int da_store_stats(da_store *st, da_stats *out);

/* Bench helper, no counterpart in the reference: uploads the queries of `p` once, then runs
 * `iters` complete searches (scan, collective, merge) back to back with inputs and results
 * resident in HBM, timed by CUDA events on shard 0's stream.  *out_total_ms covers all iters. */
int da_store_bench_resident(da_store *st, const da_search_params *p, int32_t iters, float *out_total_ms);

/* Raw device views for the test harness (row-major matrix, 1/|row| vector, doc slots). */
int da_store_debug_read_rows(da_store *st, int32_t device_index, uint64_t row_lo, uint64_t n,
                             float *out_vecs, float *out_inv_norm, uint32_t *out_doc_slot,
                             uint8_t *out_chunk_ids);

/* Test-harness view of the IVFFlat index on one shard: out_offsets lists+1 entries, out_list_rows
 * one entry per resident row (rows grouped by list), out_centroids lists x dim.  Any may be NULL. */
int da_store_debug_read_ivf(da_store *st, int32_t device_index, uint32_t *out_offsets,
                            uint32_t *out_list_rows, float *out_centroids);

/* Test-harness view of the last DA_SEARCH_IVF search on one shard: the lists each of its `batch` queries probed
 * (batch x probes list numbers, 0xFFFFFFFF = none), in no particular order. */
int da_store_debug_read_probes(da_store *st, int32_t device_index, int32_t batch, int32_t probes, uint32_t *out_lists);

/* No counterpart in the reference.  The exact scan of da_store_search for hosts that must not
 * allocate: a serving loop that cannot afford cudaMalloc's device-wide synchronisation, a process
 * whose HBM is already fully committed (a late cudaMalloc would fail), or code running where its
 * framework forbids the CUDA allocator.  The caller owns every buffer (all DEVICE pointers, on
 * `device`); the kernel's scratch (a ticket counter, the k result keys and k candidate keys per CTA) lives
 * in a fixed __device__ array inside the library (about 258 KB per GPU, held until exit), so this
 * call never runs cudaMalloc, cudaFree or any allocating routine.
 *   d_mat n_rows x dim row-major; d_inv_norm 1/|row| (0 for a zero row); d_chunk_ids 16 B per row;
 *   d_doc_bitmap (bit per document slot) with d_doc_slot (slot per row), or both NULL for "every
 *   document"; d_query dim floats.  d_mat, d_query, d_chunk_ids, d_out_chunk_ids 16-byte aligned.
 *   Outputs, best first: d_out_scores[k], d_out_rows[k] (row_base + row; may be NULL),
 *   d_out_chunk_ids[k x 16], *d_out_n = hits.  `stream` is a cudaStream_t (NULL = default stream).
 * Limits: one query, 1 <= k <= DA_NOALLOC_MAX_K, a GPU of at most DA_NOALLOC_MAX_CTAS SMs; larger
 * requests fail with DA_ERR_CAPACITY and launch nothing.
 * The first call on a device resolves the scratch array's device address and loads the kernels:
 * make it at start-up, before any stream capture on that device.  Asynchronous: outputs are valid
 * when `stream` reaches this point.  All calls on a device share the one scratch array (each call
 * clears the parts that must start at zero); calls that can overlap on the same GPU race on it --
 * serialise them, e.g. on one stream. */
#define DA_NOALLOC_MAX_K 128
#define DA_NOALLOC_MAX_CTAS 256
int da_scan_topk_noalloc(int32_t device, const float *d_mat, const float *d_inv_norm,
                         const uint32_t *d_doc_slot, const uint8_t *d_chunk_ids,
                         const uint32_t *d_doc_bitmap, const float *d_query, uint64_t n_rows,
                         uint32_t row_base, int32_t dim, int32_t k, float min_similarity, void *stream,
                         float *d_out_scores, uint32_t *d_out_rows, uint8_t *d_out_chunk_ids,
                         int32_t *d_out_n);

const char *da_last_error(void);
int da_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* DOCAGENTS_B200_H */
