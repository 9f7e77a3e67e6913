// Optional int8 sidecar of the resident matrix (DA_STORE_Q8_SIDECAR) -- SURVEY.md section 8(f) rank 4,
// "compressed resident format + fp32 re-rank of candidates", done so that answers stay exact.
//
// Next to the fp32 rows the store keeps, per row, dim int8 values q = rint(x / s) with s = max|x| / 127, the
// factor s / |x| and the row's OWN quantisation error e_x = |x - s q| / |x| (computed when the row is written).
// A query is quantised the same way (e_q).  Then, exactly (Cauchy-Schwarz on the two residuals),
//
//     | cos(q, x)  -  s_q s_x (q^ . x^) / (|q| |x|) |   <=   e_x + e_q + e_x e_q  =: E(row, query),
//
// and the integer dot product q^ . x^ is exact in int32 (<= 127^2 * dim).  The scan below reads ONE byte per
// element (dp4a, 32 elements per 256-bit load) and gives every row an interval [approx - E, approx + E] that
// holds its true cosine.  The k-th largest LOWER end T_k bounds the k-th best exact score from below, so only
// rows whose UPPER end reaches max(T_k, floor) can be in the answer: they are collected (4 B per row pass),
// re-scored in fp32 from the fp32 rows with the scan kernel's exact arithmetic and ranked
// (rescore_candidates_kernel / select_candidates_kernel, da_scan.cu).  Results are bit-identical to the fp32
// scan; only the bytes read change: dim + 12 per row instead of 4 dim + 8.
#include "da_internal.h"
#include "da_topk.cuh"

#include <cstdlib>

namespace da {

namespace {

constexpr float kQ8Margin = 2.0e-5f;      // fp32 rounding of the approximate score, of E and of the exact re-score

// One warp quantises one vector of `dim` floats (dim % 4 == 0).  inv = 1/|x| (0: zero vector -> everything 0).
__device__ void q8_quantize_vector(const float *x, int dim, float inv, int8_t *out, float *factor, float *err, int lane,
                                   float *err_max = nullptr)
{
    const float4 *x4 = reinterpret_cast<const float4 *>(x);
    const int D4 = dim >> 2;
    float amax = 0.f;
    for (int c = lane; c < D4; c += 32) {
        const float4 v = x4[c];
        amax = fmaxf(fmaxf(amax, fabsf(v.x)), fmaxf(fabsf(v.y), fmaxf(fabsf(v.z), fabsf(v.w))));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, o));
    const bool live = amax > 0.f && inv > 0.f && amax < 3.0e38f;
    const float scale = live ? amax / 127.f : 0.f;
    const float to_q = live ? 127.f / amax : 0.f;
    float e2 = 0.f;
    for (int c = lane; c < D4; c += 32) {
        const float4 v = x4[c];
        const float f[4] = {v.x, v.y, v.z, v.w};
        uint32_t packed = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            int q = __float2int_rn(f[j] * to_q);
            q = max(-127, min(127, q));
            const float d = fmaf(-(float)q, scale, f[j]);       // residual of this element
            e2 = fmaf(d, d, e2);
            packed |= ((uint32_t)q & 0xFFu) << (8 * j);
        }
        reinterpret_cast<uint32_t *>(out)[c] = packed;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e2 += __shfl_xor_sync(0xFFFFFFFFu, e2, o);
    if (lane == 0) {
        *factor = scale * inv;
        // rounded up: the sum of squares and the square root carry a few ulps each
        const float e = live ? sqrtf(e2) * inv * 1.0001f + 1.0e-7f : 0.f;
        *err = e;
        if (err_max && e > 0.f) atomicMax(reinterpret_cast<unsigned int *>(err_max), __float_as_uint(e));   // e >= 0: bits order as u32
    }
}

// rows [first, first + n) of the resident matrix, or the rows listed in `rows` (upserts)
__global__ void q8_quantize_rows_kernel(const float *mat, int dim, const uint32_t *rows, uint32_t first, uint32_t n,
                                        const float *inv_norm, int8_t *q8, float *factor, float *err, float *err_max)
{
    const int lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = w; i < n; i += nw) {
        const uint32_t r = rows ? rows[i] : first + i;
        q8_quantize_vector(mat + (size_t)r * dim, dim, inv_norm[r], q8 + (size_t)r * dim, factor + r, err + r, lane, err_max);
    }
}

__global__ void q8_quantize_queries_kernel(const float *queries, int nq, int dim, const float *q_inv, int8_t *q8,
                                           float *factor, float *err)
{
    const int lane = threadIdx.x & 31, q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (q >= nq) return;
    q8_quantize_vector(queries + (size_t)q * dim, dim, q_inv[q], q8 + (size_t)q * dim, factor + q, err + q, lane);
}

struct Q8Args {
    ScanArgs s;                 // rows/filter/top-k plumbing; s.approx receives the UPPER ends, s.out_keys the k best LOWER ends
    const int8_t *rows;         // n_rows x dim
    const float *factor;        // per row: scale / |row|   (0: zero row, never matches)
    const float *err;           // per row: e_x
    const int8_t *queries;      // nq x dim
    const float *qfactor;       // per query: scale / |q|
    const float *qerr;          // per query: e_q
    // One-kernel form (searches with a similarity floor): the CTA quantises the fp32 queries itself (raw_queries; the
    // three fields above are then unused), every wanted row whose upper end reaches `floor` is appended to the query's
    // candidate list right away (instead of writing an upper end per row for a collect pass), and the last CTA leaves
    // the gate max(floor, k-th best lower end) for the re-score (ScanArgs::gate_out).
    const float *raw_queries;   // nq x dim fp32, or nullptr for the two-pass form
    key_t64 *cand;              // [nq][cap]: ordered(upper end) << 32 | local row
    uint32_t cap;
    uint32_t *cand_count;       // [nq], zeroed by the caller
    uint32_t *overflow;
    float floor;
};

// Same frame as scan_topk_kernel (da_scan.cu): persistent CTAs, a warp takes R rows at a time, 256-bit streaming
// loads (32 int8 each), per-warp bounded lists, CTA sort, last-CTA merge.
// QS > 0 (one query, dim <= 1024 * QS): the lane keeps the int8 query words of ITS column slots (lane, lane + 32, ...:
// QS slots of 8 words) in registers for the whole scan.  The shared-memory form re-reads two 128-bit words of the query
// per slot and row group, and with 8 dp4a per 32 bytes loaded the kernel was issue-bound on exactly those
// (round 1: 81 % DRAM throughput where the fp32 scan reaches 92 %).
template <int NQ, int R, int U, int QS>
__global__ void __launch_bounds__(kScanThreads, 1) q8_scan_kernel(const Q8Args qa)
{
    static_assert(QS == 0 || NQ == 1, "the register-resident query serves one query");
    const ScanArgs &a = qa.s;
    const int DV = a.dim >> 5;                  // lane slots (32 bytes) per row
    const int DW = a.dim >> 2;                  // 32-bit words per row
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *q_s = reinterpret_cast<uint32_t *>(smem_raw);                                     // NQ x DW
    key_t64 *lists = reinterpret_cast<key_t64 *>(smem_raw + (size_t)NQ * a.dim);                // NQ x P
    __shared__ float qf_s[NQ], qe_s[NQ];
    __shared__ unsigned int is_last_s;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int K = a.k, P = a.P;
    if (qa.raw_queries) {
        // each CTA quantises the queries for itself: one warp per query, q8_quantize_queries_kernel's arithmetic
        if (warp < NQ) {
            if (warp < a.nq) {
                const float *qf = qa.raw_queries + (size_t)warp * a.dim;
                const float inv = warp_inv_norm(qf, a.dim, lane);
                q8_quantize_vector(qf, a.dim, inv, reinterpret_cast<int8_t *>(q_s) + (size_t)warp * a.dim, &qf_s[warp], &qe_s[warp], lane);
            } else {
                for (int i = lane; i < DW; i += 32) q_s[warp * DW + i] = 0u;
                if (lane == 0) { qf_s[warp] = 0.f; qe_s[warp] = 0.f; }
            }
        }
    } else {
        for (int i = tid; i < NQ * DW; i += kScanThreads)
            q_s[i] = (i / DW < a.nq) ? reinterpret_cast<const uint32_t *>(qa.queries)[i] : 0u;
        if (tid < NQ) { qf_s[tid] = tid < a.nq ? qa.qfactor[tid] : 0.f; qe_s[tid] = tid < a.nq ? qa.qerr[tid] : 0.f; }
    }
    for (int i = tid; i < NQ * P; i += kScanThreads) lists[i] = 0ull;
    __syncthreads();
    uint32_t qreg[QS > 0 ? QS : 1][8];
    if constexpr (QS > 0) {
#pragma unroll
        for (int j = 0; j < QS; j++) {
            const int c = j * 32 + lane;
#pragma unroll
            for (int w = 0; w < 8; w++) qreg[j][w] = c < DV ? q_s[c * 8 + w] : 0u;
        }
    }

    int cnt[NQ];
    key_t64 wmin[NQ];
#pragma unroll
    for (int q = 0; q < NQ; q++) { cnt[q] = 0; wmin[q] = 0ull; }

    const uint32_t n_rows = a.n_rows;
    const uint32_t n_groups = (n_rows + R - 1) / R;
    const uint32_t gstride = gridDim.x * kScanWarps;
    unsigned int my_rows_read = 0;

    for (uint32_t g = blockIdx.x * kScanWarps + warp; g < n_groups; g += gstride) {
        const uint32_t row0 = g * R;
        const uint32_t my_row = row0 + lane;
        const bool valid = (lane < R) && (my_row < n_rows);
        uint32_t want_bits = 0u;
        float my_factor = 0.f, my_err = 0.f;
        if (valid) {
            my_factor = qa.factor[my_row];
            my_err = qa.err[my_row];
            if (a.doc_bitmap) {
                const uint32_t slot = a.doc_slot[my_row];
                if (a.bitmap_stride == 0) {
                    want_bits = ((a.doc_bitmap[slot >> 5] >> (slot & 31)) & 1u) ? 0xFFFFFFFFu : 0u;
                } else {
#pragma unroll
                    for (int q = 0; q < NQ; q++)
                        if (q < a.nq)
                            want_bits |= ((a.doc_bitmap[(size_t)q * a.bitmap_stride + (slot >> 5)] >> (slot & 31)) & 1u) << q;
                }
            } else {
                want_bits = 0xFFFFFFFFu;
            }
            if (!(my_factor > 0.f)) want_bits = 0u;             // zero rows never match
        }
        const uint32_t valid_mask = __ballot_sync(0xFFFFFFFFu, valid);
        const uint32_t want_mask = __ballot_sync(0xFFFFFFFFu, want_bits != 0u);
        const uint32_t load_mask = a.read_all ? valid_mask : want_mask;
        if (load_mask == 0) {                                   // skipped rows still need an (impossible) upper end
            if (valid && !qa.cand)
                for (int q = 0; q < a.nq; q++) a.approx[(size_t)q * n_rows + my_row] = -3.0e38f;
            continue;
        }
        my_rows_read += __popc(load_mask);

        int acc[R][NQ];
#pragma unroll
        for (int r = 0; r < R; r++)
#pragma unroll
            for (int q = 0; q < NQ; q++) acc[r][q] = 0;

        const float4 *rp = reinterpret_cast<const float4 *>(qa.rows + (size_t)row0 * a.dim);
        if constexpr (QS > 0) {
#pragma unroll
            for (int j0 = 0; j0 < QS; j0 += U) {
                float4 v[R][U][2];
#pragma unroll
                for (int r = 0; r < R; r++) {
#pragma unroll
                    for (int u = 0; u < U; u++) {
                        const int c = (j0 + u) * 32 + lane;
                        if (j0 + u < QS && ((load_mask >> r) & 1u) && c < DV) {
                            ldg_stream8(rp + (size_t)r * (DW >> 2) + (size_t)c * 2, v[r][u][0], v[r][u][1]);
                        } else {
                            v[r][u][0] = make_float4(0.f, 0.f, 0.f, 0.f);
                            v[r][u][1] = make_float4(0.f, 0.f, 0.f, 0.f);
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < U; u++) {
                    if (j0 + u < QS) {
                        const uint32_t *qw = qreg[j0 + u < QS ? j0 + u : 0];
#pragma unroll
                        for (int r = 0; r < R; r++) {
                            const float4 x0 = v[r][u][0], x1 = v[r][u][1];
                            int s = acc[r][0];
                            s = __dp4a((int)__float_as_uint(x0.x), (int)qw[0], s);
                            s = __dp4a((int)__float_as_uint(x0.y), (int)qw[1], s);
                            s = __dp4a((int)__float_as_uint(x0.z), (int)qw[2], s);
                            s = __dp4a((int)__float_as_uint(x0.w), (int)qw[3], s);
                            s = __dp4a((int)__float_as_uint(x1.x), (int)qw[4], s);
                            s = __dp4a((int)__float_as_uint(x1.y), (int)qw[5], s);
                            s = __dp4a((int)__float_as_uint(x1.z), (int)qw[6], s);
                            s = __dp4a((int)__float_as_uint(x1.w), (int)qw[7], s);
                            acc[r][0] = s;
                        }
                    }
                }
            }
        } else {
        for (int c0 = 0; c0 < DV; c0 += 32 * U) {
            float4 v[R][U][2];
#pragma unroll
            for (int r = 0; r < R; r++) {
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const int c = c0 + u * 32 + lane;
                    if (((load_mask >> r) & 1u) && c < DV) {
                        ldg_stream8(rp + (size_t)r * (DW >> 2) + (size_t)c * 2, v[r][u][0], v[r][u][1]);
                    } else {
                        v[r][u][0] = make_float4(0.f, 0.f, 0.f, 0.f);
                        v[r][u][1] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                int c = c0 + u * 32 + lane;
                if (c >= DV) c = 0;                              // v is zero there
#pragma unroll
                for (int q = 0; q < NQ; q++) {
                    const uint4 qa4 = *reinterpret_cast<const uint4 *>(q_s + q * DW + c * 8);
                    const uint4 qb4 = *reinterpret_cast<const uint4 *>(q_s + q * DW + c * 8 + 4);
#pragma unroll
                    for (int r = 0; r < R; r++) {
                        const float4 x0 = v[r][u][0], x1 = v[r][u][1];
                        int s = acc[r][q];
                        s = __dp4a((int)__float_as_uint(x0.x), (int)qa4.x, s);
                        s = __dp4a((int)__float_as_uint(x0.y), (int)qa4.y, s);
                        s = __dp4a((int)__float_as_uint(x0.z), (int)qa4.z, s);
                        s = __dp4a((int)__float_as_uint(x0.w), (int)qa4.w, s);
                        s = __dp4a((int)__float_as_uint(x1.x), (int)qb4.x, s);
                        s = __dp4a((int)__float_as_uint(x1.y), (int)qb4.y, s);
                        s = __dp4a((int)__float_as_uint(x1.z), (int)qb4.z, s);
                        s = __dp4a((int)__float_as_uint(x1.w), (int)qb4.w, s);
                        acc[r][q] = s;
                    }
                }
            }
        }
        }
#pragma unroll
        for (int r = 0; r < R; r++)
#pragma unroll
            for (int q = 0; q < NQ; q++) acc[r][q] = __reduce_add_sync(0xFFFFFFFFu, acc[r][q]);

#pragma unroll
        for (int r = 0; r < R; r++) {
            if (!((valid_mask >> r) & 1u)) continue;
            const float f = __shfl_sync(0xFFFFFFFFu, my_factor, r);
            const float ex = __shfl_sync(0xFFFFFFFFu, my_err, r);
            const uint32_t wb = __shfl_sync(0xFFFFFFFFu, want_bits, r);
            const uint32_t row = row0 + r;
#pragma unroll
            for (int q = 0; q < NQ; q++) {
                if (q >= a.nq) continue;
                const bool wanted = (wb >> q) & 1u;
                const float approx = (float)acc[r][q] * f * qf_s[q];
                const float E = ex + qe_s[q] + ex * qe_s[q] + kQ8Margin;
                const float upper = fminf(1.f, approx + E);
                if (qa.cand) {
                    if (wanted && upper >= qa.floor && lane == 0) {
                        const uint32_t pos = atomicAdd(qa.cand_count + q, 1u);
                        if (pos < qa.cap) qa.cand[(size_t)q * qa.cap + pos] = ((key_t64)score_to_ordered(upper) << 32) | (key_t64)row;
                        else *qa.overflow = 1u;
                    }
                } else if (lane == 0) {
                    a.approx[(size_t)q * n_rows + row] = wanted ? upper : -3.0e38f;
                }
                if (!wanted) continue;
                const key_t64 key = make_key(fmaxf(-1.f, approx - E), a.row_base + row);
                warp_list_insert(lists + (size_t)q * P + warp * K, K, cnt[q], wmin[q], key, lane);
            }
        }
    }
    if (a.rows_read) {
        if (lane == 0 && my_rows_read) atomicAdd(a.rows_read, (unsigned long long)my_rows_read);
    }
    scan_cta_epilogue(a, scan_per_query(a), lists, &is_last_s, reinterpret_cast<key_t64 *>(smem_raw),
                      (int)(((size_t)NQ * a.dim + (size_t)NQ * P * sizeof(key_t64)) / sizeof(key_t64)));
}

// Every row whose upper end reaches max(T_k, floor) joins the query's candidate list, in the layout the re-score
// kernel of the batched path reads: ordered(upper) << 32 | local row, gate as an ordered score per query.
__global__ void q8_collect_kernel(const float *upper, uint32_t n_rows, int k, const key_t64 *lower_keys, float min_sim,
                                  key_t64 *cand, uint32_t cap, uint32_t *cand_count, uint32_t *gate_ordered, uint32_t *overflow)
{
    const int q = blockIdx.y;
    const key_t64 kth = lower_keys[(size_t)q * k + (k - 1)];            // 0 when fewer than k rows qualified
    const float gate = kth ? fmaxf(min_sim, key_score(kth)) : min_sim;
    if (blockIdx.x == 0 && threadIdx.x == 0) gate_ordered[q] = score_to_ordered(gate);
    const float *up = upper + (size_t)q * n_rows;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
        const float v = up[r];
        if (v > -1.0e38f && v >= gate) {        // -3e38 marks rows outside the document set / zero rows
            const uint32_t pos = atomicAdd(cand_count + q, 1u);
            if (pos < cap) cand[(size_t)q * cap + pos] = ((key_t64)score_to_ordered(v) << 32) | (key_t64)r;
            else *overflow = 1u;
        }
    }
}

template <int NQ, int R, int U, int QS = 0>
cudaError_t q8_launch_one(const Q8Args &qa, int grid, size_t smem, cudaStream_t stream)
{
    auto kern = q8_scan_kernel<NQ, R, U, QS>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kern<<<grid, kScanThreads, smem, stream>>>(qa);
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_q8_quantize_rows(const float *mat, int dim, const uint32_t *rows, uint32_t first, uint32_t n,
                                    const float *inv_norm, int8_t *q8, float *factor, float *err, float *err_max,
                                    cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    int blocks = (int)((n + 7) / 8);
    if (blocks > 148 * 8) blocks = 148 * 8;
    q8_quantize_rows_kernel<<<blocks, 256, 0, stream>>>(mat, dim, rows, first, n, inv_norm, q8, factor, err, err_max);
    return cudaGetLastError();
}

cudaError_t launch_q8_quantize_queries(const float *queries, int nq, int dim, const float *q_inv, int8_t *q8, float *factor,
                                       float *err, cudaStream_t stream)
{
    if (nq <= 0) return cudaSuccess;
    q8_quantize_queries_kernel<<<(nq * 32 + 255) / 256, 256, 0, stream>>>(queries, nq, dim, q_inv, q8, factor, err);
    return cudaGetLastError();
}

// Queries -> int8 (+ factor, error), the int8 scan, the candidate pass.  Leaves the candidates in `cand` /
// `cand_count` / `gate_ordered` for launch_rescore_select (da_scan.cu).  `a` as for launch_scan, with a.approx set
// (nq x n_rows floats) and a.out_keys pointing at nq x k scratch keys.
// A floor of 0.5 or more bounds the candidate set from the first row on (only rows near the query can reach it), so the
// scan can append candidates as it goes; without one the bound must come from the global k-th best lower end, which
// exists only after the scan: those searches keep the two-pass form (upper ends per row, then a collect pass).
bool q8_one_kernel_form(float min_sim)
{
    static const bool off = [] { const char *v = getenv("DA_Q8_TWO_PASS"); return v && v[0] == '1'; }();
    return !off && min_sim >= 0.5f;
}

cudaError_t launch_q8_candidates(const ScanArgs &a_in, const int8_t *q8, const float *factor, const float *err,
                                 int8_t *q8_queries, float *qfactor, float *qerr, key_t64 *cand, uint32_t cap,
                                 uint32_t *cand_count, uint32_t *gate_ordered, uint32_t *overflow, int sm_count,
                                 cudaStream_t stream, int *grid_out, int *launches_out)
{
    Q8Args qa{};
    qa.s = a_in;
    ScanArgs &a = qa.s;
    const bool one_kernel = q8_one_kernel_form(a.min_sim);
    if ((a.dim % 32) || a.nq < 1 || a.nq > kMaxNQ || (!one_kernel && !a.approx) || a.row_list) return cudaErrorInvalidValue;
    const int nqt = a.nq <= 1 ? 1 : a.nq <= 2 ? 2 : a.nq <= 4 ? 4 : 8;
    int P = next_pow2(kScanWarps * a.k);
    if (P < 2 * a.k) P = next_pow2(2 * a.k);
    a.P = P;
    const size_t smem = (size_t)nqt * a.dim + (size_t)nqt * P * sizeof(key_t64);
    if (smem > 220 * 1024) return cudaErrorInvalidValue;
    int launches = 0;
    if (one_kernel) {
        cudaError_t e0 = cudaMemsetAsync(cand_count, 0, (size_t)a.nq * 4, stream);
        if (e0 != cudaSuccess) return e0;
        qa.raw_queries = a.queries;
        qa.cand = cand; qa.cap = cap; qa.cand_count = cand_count; qa.overflow = overflow; qa.floor = a.min_sim;
        a.gate_out = gate_ordered; a.gate_floor = a.min_sim;
    } else {
        q8_quantize_queries_kernel<<<(a.nq * 32 + 255) / 256, 256, 0, stream>>>(a.queries, a.nq, a.dim, a.q_inv_norm, q8_queries,
                                                                                qfactor, qerr);
        launches++;
    }
    qa.rows = q8; qa.factor = factor; qa.err = err;
    qa.queries = q8_queries; qa.qfactor = qfactor; qa.qerr = qerr;
    a.out_ids = nullptr;
    int grid = sm_count;
    const long long need = (((long long)a.n_rows + 1) / 2 + kScanWarps - 1) / kScanWarps;
    if (need < grid) grid = need < 1 ? 1 : (int)need;
    if (grid_out) *grid_out = grid;
    cudaError_t e;
    switch (nqt) {
        // measured at 10M x 3072 (ms per search): (R,U) = (1,3) 6.28, (2,3) 4.96, (3,3) 4.55, (4,2) 4.55, (4,3) 4.52 with spills
        case 1: {
            // one query: its int8 words stay in registers when the row has at most 4 slots per lane (dim <= 4096)
            static const int ru = [] { const char *v = getenv("DA_Q8_RU"); return v ? atoi(v) : 42; }();    // measurement knob: 42, 23, 43, 41, 81
            static const bool smem_q = [] { const char *v = getenv("DA_Q8_SMEM_QUERY"); return v && v[0] == '1'; }();
            const int qs = (a.dim / 32 + 31) / 32;
            if (smem_q || qs > 4) { e = q8_launch_one<1, 4, 2>(qa, grid, smem, stream); break; }
#define DA_Q8_QS(R_, U_)                                                                  \
            (qs == 1 ? q8_launch_one<1, R_, U_, 1>(qa, grid, smem, stream)                   \
             : qs == 2 ? q8_launch_one<1, R_, U_, 2>(qa, grid, smem, stream)                 \
             : qs == 3 ? q8_launch_one<1, R_, U_, 3>(qa, grid, smem, stream)                 \
                       : q8_launch_one<1, R_, U_, 4>(qa, grid, smem, stream))
            if (ru == 23) e = DA_Q8_QS(2, 3);
            else if (ru == 43) e = DA_Q8_QS(4, 3);
            else if (ru == 41) e = DA_Q8_QS(4, 1);
            else if (ru == 81) e = DA_Q8_QS(8, 1);
            else e = DA_Q8_QS(4, 2);
#undef DA_Q8_QS
            break;
        }
        // two queries: (2,3) 6.04 ms, (4,2) 4.89, (4,1) 5.05, (3,3) 4.91; four: (2,3) 8.58, (4,2) 6.93 (spills), (4,1) 6.49, (2,2) 8.15
        case 2: e = q8_launch_one<2, 4, 2>(qa, grid, smem, stream); break;
        case 4: e = q8_launch_one<4, 4, 1>(qa, grid, smem, stream); break;
        default: e = q8_launch_one<8, 4, 1>(qa, grid, smem, stream); break;
    }
    if (e != cudaSuccess) return e;
    launches++;
    if (launches_out) *launches_out = launches + (one_kernel ? 0 : 1);
    if (one_kernel) return cudaSuccess;
    e = cudaMemsetAsync(cand_count, 0, (size_t)a.nq * 4, stream);
    if (e != cudaSuccess) return e;
    uint32_t bx = (a.n_rows + 1023) / 1024;
    if (bx > 148 * 4) bx = 148 * 4;
    if (bx < 1) bx = 1;
    q8_collect_kernel<<<dim3(bx, a.nq), 256, 0, stream>>>(a.approx, a.n_rows, a.k, a.out_keys, a.min_sim, cand, cap, cand_count,
                                                          gate_ordered, overflow);
    return cudaGetLastError();
}

}  // namespace da
