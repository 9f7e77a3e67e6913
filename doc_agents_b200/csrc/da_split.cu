// Split-plane resident format ("compressed resident formats + fp32 re-rank"), lossless:
// every fp32 element is stored as its upper and its lower 16 bits in two planes of n_rows x dim uint16.
// Same HBM footprint as the fp32 matrix, but the scan (da_scan.cu, VEC == 16) streams only the upper plane --
// truncated-bf16 values, half the bytes per pass -- and is a candidate generator.  This file holds the rest:
// writing the planes, and turning approximate scores into the exact answer:
//   1. gate per query  = max(floor - eps, T_k - 2 eps),  T_k = k-th best approximate score of the scan
//      (eps >= |cos_hi - cos_fp32|, so no row below the gate can be in the exact top-k);
//   2. collect the rows at or above the gate (one pass over 4 bytes per row);
//   3. re-score them from BOTH planes, i.e. the original fp32 bits, with the fp32 scan's arithmetic
//      (same lane ownership, FMA order, shuffle tree) -- results are bit-identical to the fp32 layout;
//   4. select the top-k and gather the chunk ids.
#include "da_internal.h"
#include "da_topk.cuh"

namespace da {

namespace {

__device__ __forceinline__ double warp_sum_double(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}

// fp32 rows -> planes.  One warp per row; 4 elements per lane per step (two packed 16-bit pairs per plane).
__device__ __forceinline__ double split_row(const float *src, int dim, uint16_t *hi, uint16_t *lo, int lane)
{
    double ss = 0.0;
    const float4 *s4 = reinterpret_cast<const float4 *>(src);
    uint2 *h2 = reinterpret_cast<uint2 *>(hi), *l2 = reinterpret_cast<uint2 *>(lo);
    for (int c = lane; c < (dim >> 2); c += 32) {
        const float4 x = s4[c];
        ss += (double)x.x * x.x + (double)x.y * x.y + (double)x.z * x.z + (double)x.w * x.w;
        const uint32_t b0 = __float_as_uint(x.x), b1 = __float_as_uint(x.y), b2 = __float_as_uint(x.z), b3 = __float_as_uint(x.w);
        h2[c] = make_uint2((b0 >> 16) | (b1 & 0xFFFF0000u), (b2 >> 16) | (b3 & 0xFFFF0000u));
        l2[c] = make_uint2((b0 & 0xFFFFu) | (b1 << 16), (b2 & 0xFFFFu) | (b3 << 16));
    }
    return warp_sum_double(ss);
}

__global__ void commit_rows_split_kernel(const float *stage, int dim, const uint32_t *rows, const uint32_t *slots,
                                         const uint8_t *ids, uint32_t n, uint16_t *hi, uint16_t *lo, float *inv_norm,
                                         uint32_t *doc_slot, uint8_t *chunk_ids)
{
    const int lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = w; i < n; i += nw) {
        const uint32_t row = rows[i];
        const double ss = split_row(stage + (size_t)i * dim, dim, hi + (size_t)row * dim, lo + (size_t)row * dim, lane);
        if (lane == 0) {
            inv_norm[row] = ss > 0.0 ? (float)(1.0 / sqrt(ss)) : 0.f;
            doc_slot[row] = slots[i];
            reinterpret_cast<uint4 *>(chunk_ids)[row] = reinterpret_cast<const uint4 *>(ids)[i];
        }
    }
}

__global__ void split_rows_kernel(const float *src, uint32_t n, int dim, uint16_t *hi, uint16_t *lo)
{
    const int lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = w; i < n; i += nw) split_row(src + (size_t)i * dim, dim, hi + (size_t)i * dim, lo + (size_t)i * dim, lane);
}

__global__ void join_rows_kernel(const uint16_t *hi, const uint16_t *lo, uint32_t n, int dim, float *dst)
{
    const size_t total = (size_t)n * dim;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = __uint_as_float(((uint32_t)hi[i] << 16) | (uint32_t)lo[i]);
}

// gate per query from the scan's approximate top-k, then every row at or above it joins the candidate list
__global__ void split_collect_kernel(const float *approx, uint32_t n_rows, int k, const key_t64 *approx_keys, float floor_eps,
                                     float two_eps, key_t64 *cand, uint32_t cap, uint32_t *cand_count, uint32_t *overflow)
{
    const int q = blockIdx.y;
    const key_t64 kth = approx_keys[(size_t)q * k + (k - 1)];          // 0 when fewer than k rows qualified
    const float gate = kth ? fmaxf(floor_eps, key_score(kth) - two_eps) : floor_eps;
    const float *ap = approx + (size_t)q * n_rows;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
        const float v = ap[r];
        if (v > -1.0e38f && v >= gate) {        // -3e38 marks rows outside the document set / zero rows
            const uint32_t pos = atomicAdd(cand_count + q, 1u);
            if (pos < cap) cand[(size_t)q * cap + pos] = (key_t64)r;
            else *overflow = 1u;
        }
    }
}

// exact fp32 score of every candidate from both planes; the entry becomes its final key (0 if below the floor)
__global__ void __launch_bounds__(256) split_rescore_kernel(key_t64 *cand, const uint32_t *cand_count, uint32_t cap,
                                                            const uint16_t *hi, const uint16_t *lo, int dim,
                                                            const float *inv_norm, const float *queries, const float *q_inv,
                                                            uint32_t row_base, float min_sim)
{
    const int q = blockIdx.y;
    const uint32_t cnt = min(cand_count[q], cap);
    const int lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    const float4 *qv = reinterpret_cast<const float4 *>(queries + (size_t)q * dim);
    const float qi = q_inv[q];
    key_t64 *lst = cand + (size_t)q * cap;
    const int D8 = dim >> 3;
    for (uint32_t i = w; i < cnt; i += nw) {
        const uint32_t row = (uint32_t)lst[i];
        const uint4 *h = reinterpret_cast<const uint4 *>(hi + (size_t)row * dim);
        const uint4 *l = reinterpret_cast<const uint4 *>(lo + (size_t)row * dim);
        float acc = 0.f;
        // the fp32 scan's order: a lane owns 8 consecutive elements, lane + 32 j, x-y-z-w FMAs
        for (int c = lane; c < D8; c += 32) {
            const uint4 hw = __ldg(h + c), lw = __ldg(l + c);
            const uint32_t hh[4] = {hw.x, hw.y, hw.z, hw.w}, ll[4] = {lw.x, lw.y, lw.z, lw.w};
            float x[8];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                x[2 * j] = __uint_as_float((hh[j] << 16) | (ll[j] & 0xFFFFu));
                x[2 * j + 1] = __uint_as_float((hh[j] & 0xFFFF0000u) | (ll[j] >> 16));
            }
            const float4 w0 = qv[2 * c], w1 = qv[2 * c + 1];
            acc = fmaf(x[0], w0.x, acc); acc = fmaf(x[1], w0.y, acc); acc = fmaf(x[2], w0.z, acc); acc = fmaf(x[3], w0.w, acc);
            acc = fmaf(x[4], w1.x, acc); acc = fmaf(x[5], w1.y, acc); acc = fmaf(x[6], w1.z, acc); acc = fmaf(x[7], w1.w, acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
        float sim = acc * inv_norm[row] * qi;
        sim = fminf(1.f, fmaxf(-1.f, sim));
        if (lane == 0) lst[i] = (sim >= min_sim) ? make_key(sim, row_base + row) : 0ull;
    }
}

// one CTA per query: top-k of its re-scored candidates, chunk-id gather
__global__ void __launch_bounds__(256) split_select_kernel(const key_t64 *cand, const uint32_t *cand_count, uint32_t cap, int k,
                                                           int P, const uint8_t *chunk_ids, uint32_t row_base,
                                                           key_t64 *out_keys, uint8_t *out_ids)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    key_t64 *buf = reinterpret_cast<key_t64 *>(smem_raw);
    const int q = blockIdx.x, tid = threadIdx.x;
    const uint32_t n = min(cand_count[q], cap);
    const key_t64 *lst = cand + (size_t)q * cap;
    for (int i = tid; i < k; i += blockDim.x) buf[i] = 0ull;
    const uint32_t room = (uint32_t)(P - k);
    uint32_t done = 0;
    do {
        __syncthreads();
        for (uint32_t i = tid; i < room; i += blockDim.x) buf[k + i] = (done + i < n) ? lst[done + i] : 0ull;
        cta_bitonic_desc(buf, P);
        done += room;
    } while (done < n);
    for (int i = tid; i < k; i += blockDim.x) {
        const key_t64 key = buf[i];
        out_keys[(size_t)q * k + i] = key;
        uint4 id = make_uint4(0, 0, 0, 0);
        if (key) id = reinterpret_cast<const uint4 *>(chunk_ids)[key_row(key) - row_base];
        reinterpret_cast<uint4 *>(out_ids)[(size_t)q * k + i] = id;
    }
}

}  // namespace

cudaError_t launch_commit_rows_split(const float *stage, int dim, const uint32_t *rows, const uint32_t *slots,
                                     const uint8_t *ids, uint32_t n, uint16_t *hi, uint16_t *lo, float *inv_norm,
                                     uint32_t *doc_slot, uint8_t *chunk_ids, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    int blocks = (int)((n + 7) / 8);
    if (blocks > 148 * 8) blocks = 148 * 8;
    commit_rows_split_kernel<<<blocks, 256, 0, stream>>>(stage, dim, rows, slots, ids, n, hi, lo, inv_norm, doc_slot, chunk_ids);
    return cudaGetLastError();
}

cudaError_t launch_split_rows(const float *src, uint32_t n, int dim, uint16_t *hi, uint16_t *lo, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    split_rows_kernel<<<148 * 8, 256, 0, stream>>>(src, n, dim, hi, lo);
    return cudaGetLastError();
}

cudaError_t launch_join_rows(const uint16_t *hi, const uint16_t *lo, uint32_t n, int dim, float *dst, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    join_rows_kernel<<<148 * 4, 256, 0, stream>>>(hi, lo, n, dim, dst);
    return cudaGetLastError();
}

cudaError_t launch_split_finish(const float *approx, uint32_t n_rows, int nq, int k, const key_t64 *approx_keys,
                                float min_sim, float eps, key_t64 *cand, uint32_t cap, uint32_t *cand_count,
                                uint32_t *overflow, const uint16_t *hi, const uint16_t *lo, int dim, const float *inv_norm,
                                const float *queries, const float *q_inv, const uint8_t *chunk_ids, uint32_t row_base,
                                key_t64 *out_keys, uint8_t *out_ids, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(cand_count, 0, (size_t)nq * 4, stream);
    if (e != cudaSuccess) return e;
    uint32_t bx = (n_rows + 1023) / 1024;
    if (bx > 148 * 4) bx = 148 * 4;
    if (bx < 1) bx = 1;
    split_collect_kernel<<<dim3(bx, nq), 256, 0, stream>>>(approx, n_rows, k, approx_keys, min_sim - eps, 2.f * eps, cand, cap,
                                                           cand_count, overflow);
    split_rescore_kernel<<<dim3(148, nq), 256, 0, stream>>>(cand, cand_count, cap, hi, lo, dim, inv_norm, queries, q_inv,
                                                            row_base, min_sim);
    int P = next_pow2(2 * k);
    if (P < 2048) P = 2048;
    split_select_kernel<<<nq, 256, (size_t)P * sizeof(key_t64), stream>>>(cand, cand_count, cap, k, P, chunk_ids, row_base,
                                                                        out_keys, out_ids);
    return cudaGetLastError();
}

}  // namespace da
