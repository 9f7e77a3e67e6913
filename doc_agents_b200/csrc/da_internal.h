// Internal declarations shared by the .cu translation units of libdocagents_b200.so.
// Not part of the C ABI (include/docagents_b200.h is).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace da {

// ---------------------------------------------------------------------------
// Candidate key: one 64-bit integer orders (score desc, global row asc).
//   hi 32 = order-preserving image of the fp32 similarity
//   lo 32 = ~global_row                 (lower row => larger key)
// 0 is "empty".  Ordering ties the way DESIGN.md "Tie rule" states.
// ---------------------------------------------------------------------------
typedef unsigned long long key_t64;

__host__ __device__ inline uint32_t score_to_ordered(float f)
{
#ifdef __CUDA_ARCH__
    uint32_t b = __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; uint32_t b = c.u;
#endif
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ inline float ordered_to_score(uint32_t o)
{
    uint32_t b = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    union { float f; uint32_t u; } c; c.u = b; return c.f;
#endif
}
__host__ __device__ inline key_t64 make_key(float score, uint32_t global_row)
{
    return ((key_t64)score_to_ordered(score) << 32) | (key_t64)(~global_row);
}
__host__ __device__ inline uint32_t key_row(key_t64 k) { return ~(uint32_t)(k & 0xFFFFFFFFull); }
__host__ __device__ inline float key_score(key_t64 k) { return ordered_to_score((uint32_t)(k >> 32)); }

// ---------------------------------------------------------------------------
// Exact scan
// ---------------------------------------------------------------------------
constexpr int kScanThreads = 512;
constexpr int kScanWarps = kScanThreads / 32;
constexpr int kMaxNQ = 8;          // queries served by one pass of the SIMT scan

// Cross-GPU exchange fused into the scan (one process per GPU): every rank's last CTA stores its records into its
// slot of EVERY rank's exchange buffer (plain stores through CUDA IPC mappings over NVLink), announces them on
// each rank's arrival counter, waits until its own counter shows all ranks, and merges -- no collective kernel,
// no merge kernel, no host round trip.  Buffer layout per rank: [counter parity 0][counter parity 1] (128 B apart)
// then data[parity][rank] of slot_bytes each, from byte 256.  Searches alternate parity; counters only grow.
constexpr int kXchgMaxRanks = 8;
struct XchgArgs {
    unsigned char *peer[kXchgMaxRanks];     // every rank's buffer as mapped in THIS process (peer[rank] = own)
    int n_ranks, rank;                      // n_ranks <= 1: no exchange
    uint32_t slot_bytes;
    uint32_t parity;
    unsigned long long target;              // own counter value that means "all ranks delivered"
    key_t64 *final_keys;                    // merged result of this rank: nq x k keys, nq x k x 16 ids
    uint8_t *final_ids;
    uint32_t *timeout_flag;                 // set to 1 if a peer did not deliver within the bound (then the result is invalid)
};

struct ScanArgs {
    const float *mat;              // n_rows x dim, row-major, 16-byte aligned rows (nullptr in split-plane stores)
    const uint16_t *hi;            // split-plane stores: upper 16 bits of every element, n_rows x dim (dim % 16 == 0)
    float *approx;                 // split-plane stores: nq x n_rows similarities computed from the hi plane
    const float *inv_norm;         // 1/|row| (0 for a zero row)
    const uint32_t *doc_slot;      // document slot of each row
    const uint8_t *chunk_ids;      // 16 bytes per row
    const uint32_t *doc_bitmap;    // bit per document slot; nullptr = every document
    uint32_t bitmap_stride;        // words between the bitmaps of consecutive queries; 0 = one bitmap for all
    const float *queries;          // nq x dim (device)
    const float *q_inv_norm;       // nq; nullptr = the kernel computes 1/|q| itself from the staged queries
    const uint32_t *row_list;      // optional: scan only these rows (compacted document filter); nullptr = [0,n_rows)
    const uint32_t *n_rows_dev;    // with row_list: device-resident count of entries
    // IVFFlat probe mode (one launch for a whole batch, blockIdx.y = query; every query scans its own lists):
    // probe_keys[q][probes] name the lists (key row = list id, 0 = none), list_offsets/list_rows are the index.
    // Per-query pieces (queries, q_inv_norm, doc_bitmap with bitmap_stride, counter, cta_out, out_keys, out_ids)
    // are indexed by blockIdx.y inside the kernel; nq must be 1.
    const key_t64 *probe_keys;
    int probes;
    int probe_batch;               // queries of the launch (grid.y)
    const uint32_t *list_offsets;  // lists + 1
    const uint32_t *list_rows;     // rows grouped by list
    // Range mode (one query): the rows are the concatenation of `probes` row ranges, ranges[i] = (first row, count)
    // -- the rows of the requested documents, which arrive contiguously (one SaveEmbeddings per document).  No
    // bitmap, no document slots, no compaction pass.  Same machinery as probe mode with an identity row list.
    const uint2 *ranges;
    // Small-search extras, all optional, done by the last CTA after the records are written: the rows-read counter is
    // copied to *rows_read_out and cleared, *n_rows_reset is cleared (the compaction count of the next search starts
    // at zero), and *done_flag = done_value after a system-wide fence -- out_keys/out_ids/rows_read_out/done_flag may
    // then be mapped host memory that the caller polls instead of issuing a copy and a synchronisation.
    unsigned long long *rows_read_out;
    uint32_t *n_rows_reset;
    volatile unsigned long long *done_flag;
    unsigned long long done_value;
    XchgArgs xg;
    // Optional: after the final keys of query q are known the last CTA writes gate_out[q] = ordered(max(gate_floor,
    // score of the k-th key)) -- or ordered(gate_floor) while fewer than k rows qualified (int8 sidecar: the gate of
    // the candidate re-score, which round 1 computed in a kernel of its own).
    uint32_t *gate_out;
    float gate_floor;
    uint32_t n_rows;               // rows to visit when row_list is null
    uint32_t row_base;             // global row of local row 0
    int dim;
    int nq;                        // <= kMaxNQ
    int k;
    int P;                         // pow2 >= kScanWarps * k   (smem sort width)
    float min_sim;
    int read_all;                  // 1: doc mask applied after reading the row
    key_t64 *cta_out;              // gridDim.x x nq x k   scratch
    unsigned int *counter;         // scratch: last-block ticket (left at 0)
    key_t64 *out_keys;             // nq x k, best first, 0-padded
    uint8_t *out_ids;              // nq x k x 16; nullptr = do not gather ids
    unsigned long long *rows_read; // optional statistics counter
};

size_t scan_smem_bytes(int dim, int nq, int k, int *P_out);
int scan_pick_nq(int dim, int k, int batch);   // how many queries one pass may carry
// Launches the fused scan on `stream`; grid is chosen inside (persistent, multiple of SM count).
cudaError_t launch_scan(const ScanArgs &a, int sm_count, cudaStream_t stream, int *grid_out);
int scan_grid_for(int sm_count);
// Forces the (lazy) loading, on the current device, of every kernel a one-query, non-indirect launch_scan can pick.
cudaError_t scan_preload_single_query_variants();

// Rows whose document bit is set in `bitmap`, compacted (in no particular order) into rows[0 .. *n_out).
cudaError_t launch_filter_compact(const uint32_t *doc_slot, const uint32_t *bitmap, uint32_t n_rows, uint32_t *rows,
                                  uint32_t *n_out, cudaStream_t stream);

// Merge of G gathered candidate blocks (one block per rank, `block_bytes` apart -- 0 = packed --, layout: batch*k ids
// of 16 B, then batch*k keys, then an optional trailer whose first word is the rank's overflow flag; the OR of the
// trailers goes to *out_flags when it is not null).
cudaError_t launch_merge(const uint8_t *gathered, int G, int batch, int k, key_t64 *out_keys,
                         uint8_t *out_ids, cudaStream_t stream, size_t block_bytes = 0, uint32_t *out_flags = nullptr);

// ---------------------------------------------------------------------------
// Row maintenance
// ---------------------------------------------------------------------------
// inv_norm / doc_slot / chunk id of the rows listed in `rows`.
cudaError_t launch_finalize_rows(const float *mat, int dim, const uint32_t *rows, const uint32_t *slots,
                                 const uint8_t *ids, uint32_t n, float *inv_norm, uint32_t *doc_slot,
                                 uint8_t *chunk_ids, cudaStream_t stream);
cudaError_t launch_query_inv_norm(const float *queries, int dim, int nq, float *out, cudaStream_t stream);
cudaError_t launch_validate_stage(const float *stage, uint64_t n_floats, uint64_t base, unsigned long long *first_bad,
                                  cudaStream_t stream);
cudaError_t launch_commit_rows(const float *stage, int dim, const uint32_t *rows, const uint32_t *slots, const uint8_t *ids,
                               uint32_t n, float *mat, float *inv_norm, uint32_t *doc_slot, uint8_t *chunk_ids,
                               cudaStream_t stream);

struct SynthArgs {
    uint64_t seed_data, seed_centroid, seed_ids;
    uint32_t chunks_per_doc, n_clusters;
    int32_t cmul, nmul;
    int32_t smul; uint32_t super_group;   // optional second level: + smul * direction of the cluster's group (0 = off)
    uint64_t row_lo;               // global row of local row 0
    uint32_t n_rows;
    uint32_t doc_slot_base;        // global doc index mapped to slot 0
    int dim;
};
cudaError_t launch_synth(const SynthArgs &a, float *mat, float *inv_norm, uint32_t *doc_slot,
                         uint8_t *chunk_ids, cudaStream_t stream);

// ---------------------------------------------------------------------------
// IVFFlat (da_ivf.cu)
// ---------------------------------------------------------------------------
#define DA_IVF_MAX_PROBES 128      // = DA_MAX_K: the centroid ranking is a top-k of the fused scan
cudaError_t ivf_launch_assign(const float *mat, int D, const uint32_t *row_idx, uint32_t n, const float *cent, int L,
                              uint32_t *out_assign, int out_by_row, cudaStream_t stream);
cudaError_t ivf_launch_sample(uint32_t n_rows, uint32_t n_sample, uint64_t seed, uint32_t *out, cudaStream_t stream);
cudaError_t ivf_launch_init_centroids(const float *mat, int D, const float *inv_norm, const uint32_t *sample,
                                      uint32_t n_sample, int L, uint64_t seed, float *cent, float *cent_inv, cudaStream_t stream);
// counts[assign[i]] += 1 over the sample (counts zeroed by the caller)
cudaError_t ivf_launch_count(const uint32_t *assign, uint32_t n, float *counts, cudaStream_t stream);
// One balancing round after k-means: the heaviest lists (by `counts`, the same on every rank) each split into
// themselves and one empty list, along a direction taken from one of their members in this shard's sample.
size_t ivf_split_scratch_bytes(int L);
cudaError_t ivf_launch_split_round(const float *counts, int L, void *scratch, const float *mat, int D, const float *inv_norm,
                                   const uint32_t *sample, uint32_t n_sample, const uint32_t *assign, uint64_t seed,
                                   float *cent, cudaStream_t stream, float heavy_x = 1.5f, float light_x = 0.25f);
// (heavy_x, light_x: a donor holds more than heavy_x times the mean, a receiver less than light_x times)
// Small batches: out[q][l] = <queries[q], cent[l]> in fp32 for nq <= 8 queries (one pass over the centroid matrix)
cudaError_t ivf_launch_centroid_scores(const float *cent, int L, int D, const float *queries, int nq, float *out, int sm_count,
                                       cudaStream_t stream);
// Query -> centroid ranking: partial[ks][batch][L] scores (the GEMM of da_tc.cu, or ks = 1 from the kernel above) -> out_keys[batch][probes]
// One query: scores + selection in one launch (`ticket`: a zeroed device word, left zero)
cudaError_t ivf_launch_rank_one_query(const float *cent, int L, int D, const float *query, float *scores, int probes,
                                      key_t64 *out_keys, unsigned int *ticket, int sm_count, cudaStream_t stream);
// Raises a kernel's dynamic shared-memory limit once per (function, device) instead of per launch (da_scan.cu)
cudaError_t ensure_dynamic_smem(const void *func, size_t smem);
constexpr int kIvfSelectMaxLists = 49152;          // the select kernel keeps a query's L scores in shared memory
cudaError_t ivf_launch_select_probes(const float *partial, int ks, int batch, int L, int probes, key_t64 *out_keys,
                                     cudaStream_t stream);
// scores[split][m][n] = <A[m], B[n]> over the split's slice of the dimension, TF32 on the tensor cores (da_tc.cu).
// *ks_out = number of slices written (the caller sizes `out` with tc_scores_max_splits()).
int tc_scores_splits(int M, int N, int D, int sm_count);
cudaError_t tc_launch_scores(const float *A, int M, const float *B, int N, int D, int ks, float *out, cudaStream_t stream);
cudaError_t ivf_launch_accumulate(const float *mat, int D, const float *inv_norm, const uint32_t *sample,
                                  uint32_t n_sample, const uint32_t *assign, float *sums, float *counts,
                                  cudaStream_t stream);
cudaError_t ivf_launch_finalize(const float *sums, const float *counts, int L, int D, float *cent, cudaStream_t stream);
cudaError_t ivf_launch_build_lists(const uint32_t *assign, uint32_t n, int L, uint32_t *counts, uint32_t *offsets,
                                   uint32_t *cursor, uint32_t *list_rows, cudaStream_t stream);
// CTAs per query of a probe-mode scan over `batch` queries (the scratch of the launch is batch x this x k keys).
int scan_probe_ctas_per_query(int sm_count, int batch);

// Tensor-core assignment (da_tc.cu): out_assign[m] = argmax_l <mat[m], cent[l]>, tcgen05 kind::tf32.
size_t tc_assign_scratch_bytes(uint32_t M);
cudaError_t tc_launch_assign(const float *mat, uint32_t M, int D, const float *cent, int L,
                             unsigned long long *packed_scratch, uint32_t *out_assign, cudaStream_t stream);
cudaError_t tc_launch_gather_rows(const float *mat, int D, const uint32_t *rows, uint32_t n, float *out, cudaStream_t stream);

// int8 sidecar operands of the batched path (kind::i8): then `inv_norm` = scale/|row| per row, `q_inv` = scale/|q| per
// query, `eps` = the largest quantisation-error bound of a resident row, q_err = the queries' own bounds.
struct TcInt8Operands { const int8_t *queries; const int8_t *rows; const float *q_err; };

// Batched queries on the tensor cores (da_tc.cu): TF32 candidate generation behind a global per-query gate,
// then exact fp32 re-score and per-query selection (da_scan.cu).  eps bounds |cos_tf32 - cos_fp32|.
constexpr float kTf32Eps = 2.5e-3f;
#define DA_TC_MAX_BATCH 4096
constexpr int kSurvivorCap = 16384;       // re-scored survivors kept per query (a 10M-row cluster mass within the
                                          // bf16 slack of the k-th best is ~5 000 rows)
int tc_batch_cap(int k);
int tc_batch_m_tiles(int batch);
size_t tc_batch_hist_words(int batch);
cudaError_t tc_launch_batch_candidates(const float *queries, int batch, const float *q_inv, const float *mat,
                                       uint32_t n_rows, int D, const float *inv_norm, const uint32_t *doc_slot,
                                       const uint32_t *doc_bitmap, uint32_t bitmap_stride, int k, float min_sim, float eps, int sm_count,
                                       unsigned long long *cand, uint32_t *cand_count, uint32_t *hist,
                                       uint32_t *gate_ordered, uint32_t *overflow, uint32_t *group_sync,
                                       cudaStream_t stream, const uint16_t *q_hi = nullptr, const uint16_t *x_hi = nullptr,
                                       const struct TcInt8Operands *i8 = nullptr);
// upper 16 bits of n floats (truncated bf16), e.g. the queries of a batch for the bf16 contraction
cudaError_t tc_launch_truncate_bf16(const float *src, size_t n, uint16_t *dst, cudaStream_t stream);
cudaError_t launch_rescore_select(const key_t64 *cand, const uint32_t *cand_count, int cap, const uint32_t *gate_ordered,
                                  key_t64 *surv, uint32_t *surv_count, uint32_t *overflow, int batch,
                                  const float *mat, int dim, const float *inv_norm, const float *queries,
                                  const float *q_inv, const uint8_t *chunk_ids, uint32_t row_base, float min_sim,
                                  int k, key_t64 *out_keys, uint8_t *out_ids, cudaStream_t stream,
                                  const uint16_t *hi = nullptr, const uint16_t *lo = nullptr,
                                  bool gate_bounds_exact = false);   // true: the gate also bounds the k-th best EXACT score

// ---------------------------------------------------------------------------
// Split-plane layout (da_split.cu): every fp32 element lives as two 16-bit halves in two planes (lossless,
// same bytes).  The scan reads only the upper plane (truncated-bf16 values, half the HBM traffic) and is a
// candidate generator; candidates are re-scored from both planes with the fp32 scan's arithmetic.
// eps bounds |cos_hi - cos_fp32|: truncation moves an element by < 2^-8 relative, so by Cauchy-Schwarz the
// cosine moves by < 2^-8; the slack adds fp32 summation noise.
// ---------------------------------------------------------------------------
constexpr float kSplitEps = 4.0e-3f;
// the bf16 contraction truncates the queries too: two truncations of < 2^-8 each
constexpr float kSplitBf16Eps = 8.5e-3f;
cudaError_t launch_commit_rows_split(const float *stage, int dim, const uint32_t *rows, const uint32_t *slots,
                                     const uint8_t *ids, uint32_t n, uint16_t *hi, uint16_t *lo, float *inv_norm,
                                     uint32_t *doc_slot, uint8_t *chunk_ids, cudaStream_t stream);
cudaError_t launch_split_rows(const float *src, uint32_t n, int dim, uint16_t *hi, uint16_t *lo, cudaStream_t stream);
cudaError_t launch_join_rows(const uint16_t *hi, const uint16_t *lo, uint32_t n, int dim, float *dst, cudaStream_t stream);
// After the hi-plane scan (top-k by approximate score in approx_keys, all approximate scores in approx):
// collect every row that can still be in the exact top-k, re-score it exactly from both planes, select.
cudaError_t launch_split_finish(const float *approx, uint32_t n_rows, int nq, int k, const key_t64 *approx_keys,
                                float min_sim, float eps, key_t64 *cand, uint32_t cap, uint32_t *cand_count,
                                uint32_t *overflow, const uint16_t *hi, const uint16_t *lo, int dim, const float *inv_norm,
                                const float *queries, const float *q_inv, const uint8_t *chunk_ids, uint32_t row_base,
                                key_t64 *out_keys, uint8_t *out_ids, cudaStream_t stream);

// Host restatement of the id hash of the synthetic corpus (spec: DESIGN.md).
void synth_doc_id(uint64_t seed_ids, uint64_t doc, uint8_t out[16]);
void synth_vector_host(uint64_t seed_noise, uint64_t noise_index, uint64_t seed_centroid, uint64_t cluster,
                       int cmul, int nmul, int dim, float *out, int smul = 0, uint32_t super_group = 0);
void synth_id_host(uint64_t seed_ids, int kind, uint64_t index, uint8_t out[16]);

// ---------------------------------------------------------------------------
// int8 sidecar (da_q8.cu): candidates from one byte per element, exact answers after the fp32 re-score
// ---------------------------------------------------------------------------
// rows [first, first + n), or the n rows listed in `rows` (device), of `mat` -> q8 / factor / err
cudaError_t launch_q8_quantize_rows(const float *mat, int dim, const uint32_t *rows, uint32_t first, uint32_t n,
                                    const float *inv_norm, int8_t *q8, float *factor, float *err, float *err_max,
                                    cudaStream_t stream);    // err_max: running maximum of err over every row written
// quantise a.nq queries, scan the int8 rows, collect every row whose interval reaches the k-th best lower end;
// candidates are left for launch_rescore_select (cand / cand_count / gate_ordered)
// queries -> int8 + scale/|q| + error bound (one warp per query)
cudaError_t launch_q8_quantize_queries(const float *queries, int nq, int dim, const float *q_inv, int8_t *q8, float *factor,
                                       float *err, cudaStream_t stream);
cudaError_t launch_q8_candidates(const ScanArgs &a, const int8_t *q8, const float *factor, const float *err,
                                 int8_t *q8_queries, float *qfactor, float *qerr, key_t64 *cand, uint32_t cap,
                                 uint32_t *cand_count, uint32_t *gate_ordered, uint32_t *overflow, int sm_count,
                                 cudaStream_t stream, int *grid_out, int *launches_out = nullptr);
// true: a search with this floor takes the one-kernel form (the scan quantises the queries itself, appends its
// candidates and leaves the gate); then a.q_inv_norm may be null and launch_rescore_select may be given q_inv = nullptr
bool q8_one_kernel_form(float min_sim);

}  // namespace da
