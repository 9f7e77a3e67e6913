// Row maintenance kernels: per-row 1/|v|, document slot and chunk id after an upsert
// (write side of the path: Store.SaveEmbeddings, internal/store/postgres.go:176-201),
// query norms, and the device-side synthetic corpus used by tests and bench.
#include "da_internal.h"
#include "da_topk.cuh"

namespace da {

namespace {

__global__ void finalize_rows_kernel(const float *mat, int dim, const uint32_t *rows, const uint32_t *slots,
                                     const uint8_t *ids, uint32_t n, float *inv_norm, uint32_t *doc_slot,
                                     uint8_t *chunk_ids)
{
    const int lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = w; i < n; i += nw) {
        const uint32_t row = rows[i];
        float inv = warp_inv_norm(mat + (size_t)row * dim, dim, lane);
        if (lane == 0) {
            inv_norm[row] = inv;
            doc_slot[row] = slots[i];
            reinterpret_cast<uint4 *>(chunk_ids)[row] = reinterpret_cast<const uint4 *>(ids)[i];
        }
    }
}

// Finite check of staged vectors on the device (pgvector rejects NaN/Inf at insert): smallest flat index of a
// non-finite element, or ~0 if all are finite.
__global__ void validate_stage_kernel(const float *stage, unsigned long long n_floats, unsigned long long base,
                                      unsigned long long *first_bad)
{
    const unsigned long long n4 = n_floats >> 2;
    unsigned long long bad = ~0ull;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n4;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const uint4 v = reinterpret_cast<const uint4 *>(stage)[i];
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; j++)
            if ((w[j] & 0x7F800000u) == 0x7F800000u && base + i * 4 + j < bad) bad = base + i * 4 + j;
    }
    if (bad != ~0ull) atomicMin(first_bad, bad);
}

// Staged rows -> their final rows, with 1/|row|, document slot and chunk id: one warp per row, one pass.
__global__ void commit_rows_kernel(const float *stage, int dim, const uint32_t *rows, const uint32_t *slots,
                                   const uint8_t *ids, uint32_t n, float *mat, float *inv_norm, uint32_t *doc_slot,
                                   uint8_t *chunk_ids)
{
    const int lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = w; i < n; i += nw) {
        const uint32_t row = rows[i];
        const float4 *src = reinterpret_cast<const float4 *>(stage + (size_t)i * dim);
        float4 *dst = reinterpret_cast<float4 *>(mat + (size_t)row * dim);
        double ss = 0.0;
        for (int c = lane; c < (dim >> 2); c += 32) {
            const float4 x = src[c];
            dst[c] = x;
            ss += (double)x.x * x.x + (double)x.y * x.y + (double)x.z * x.z + (double)x.w * x.w;
        }
        ss = warp_sum_d(ss);
        if (lane == 0) {
            inv_norm[row] = ss > 0.0 ? (float)(1.0 / sqrt(ss)) : 0.f;
            doc_slot[row] = slots[i];
            reinterpret_cast<uint4 *>(chunk_ids)[row] = reinterpret_cast<const uint4 *>(ids)[i];
        }
    }
}

__global__ void query_inv_norm_kernel(const float *queries, int dim, int nq, float *out)
{
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= nq) return;
    float inv = warp_inv_norm(queries + (size_t)w * dim, dim, lane);
    if (lane == 0) out[w] = inv;
}

// ---------------------------------------------------------------------------
// Synthetic corpus (spec in DESIGN.md; integer-exact, so the result is bit-identical
// to the CPU oracle's independent statement of the same spec).
// ---------------------------------------------------------------------------
constexpr unsigned long long kGold = 0x9E3779B97F4A7C15ull;
constexpr unsigned long long kPair = 0xD1B54A32D192ED03ull;
constexpr unsigned long long kSuperTag = 0x5C0DE5C0DE5C0DE5ull;     // seed_centroid ^ this = the stream of group directions

__host__ __device__ inline unsigned long long mix64(unsigned long long z)
{
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}
__host__ __device__ inline unsigned long long stream_key(unsigned long long seed, unsigned long long index)
{
    return mix64(seed ^ ((index + 1) * kGold));
}
__host__ __device__ inline int ih4(uint32_t w)
{
    return (int)((w & 255u) + ((w >> 8) & 255u) + ((w >> 16) & 255u) + (w >> 24)) - 510;
}
__host__ __device__ inline void id16(unsigned long long seed, unsigned long long tag, unsigned long long index,
                                     unsigned long long &a, unsigned long long &b)
{
    a = mix64(seed ^ mix64(tag + 2 * index));
    b = mix64(seed ^ mix64(tag + 2 * index + 1));
    a = (a & ~(0xFFull << 48)) | (((((a >> 48) & 0x0Full)) | 0x40ull) << 48);   // byte 6: version 4
    b = (b & ~0xFFull) | ((b & 0x3Full) | 0x80ull);                             // byte 8: RFC 4122 variant
}

__global__ void synth_kernel(const SynthArgs a, float *mat, float *inv_norm, uint32_t *doc_slot, uint8_t *chunk_ids)
{
    const int lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
    const int pairs = a.dim >> 1;
    for (uint32_t r = w; r < a.n_rows; r += nw) {
        const unsigned long long grow = a.row_lo + r;
        const uint32_t doc = (uint32_t)(grow / a.chunks_per_doc);
        const unsigned long long kn = stream_key(a.seed_data, grow);
        const unsigned long long kc = stream_key(a.seed_centroid, doc % a.n_clusters);
        const bool two_level = a.smul != 0 && a.super_group != 0;
        const unsigned long long ks = two_level ? stream_key(a.seed_centroid ^ kSuperTag, (doc % a.n_clusters) / a.super_group) : 0ull;
        float2 *out = reinterpret_cast<float2 *>(mat + (size_t)r * a.dim);
        long long ss = 0;
        for (int p = lane; p < pairs; p += 32) {
            unsigned long long hn = mix64(kn + (unsigned long long)(p + 1) * kPair);
            unsigned long long hc = mix64(kc + (unsigned long long)(p + 1) * kPair);
            int u0 = a.cmul * ih4((uint32_t)hc) + a.nmul * ih4((uint32_t)hn);
            int u1 = a.cmul * ih4((uint32_t)(hc >> 32)) + a.nmul * ih4((uint32_t)(hn >> 32));
            if (two_level) {
                const unsigned long long hs = mix64(ks + (unsigned long long)(p + 1) * kPair);
                u0 += a.smul * ih4((uint32_t)hs);
                u1 += a.smul * ih4((uint32_t)(hs >> 32));
            }
            ss += (long long)u0 * u0 + (long long)u1 * u1;
            out[p] = make_float2((float)u0, (float)u1);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xFFFFFFFFu, ss, o);
        if (ss != 0) {
            // float32 magnitude from the exact float64 sum of squares, then an IEEE divide per
            // element: the normalisation of internal/embeddings/openai.go:146-158.
            const float mag = (float)sqrt((double)ss);
            for (int p = lane; p < pairs; p += 32) {
                float2 v = out[p];
                out[p] = make_float2(__fdiv_rn(v.x, mag), __fdiv_rn(v.y, mag));
            }
        }
        __syncwarp();
        float inv = warp_inv_norm(mat + (size_t)r * a.dim, a.dim, lane);
        if (lane == 0) {
            inv_norm[r] = inv;
            doc_slot[r] = doc - a.doc_slot_base;
            unsigned long long ia, ib;
            id16(a.seed_ids, 0xC4A9000000000000ull, grow, ia, ib);
            reinterpret_cast<ulonglong2 *>(chunk_ids)[r] = make_ulonglong2(ia, ib);
        }
    }
}

}  // namespace

cudaError_t launch_finalize_rows(const float *mat, int dim, const uint32_t *rows, const uint32_t *slots,
                                 const uint8_t *ids, uint32_t n, float *inv_norm, uint32_t *doc_slot,
                                 uint8_t *chunk_ids, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    int blocks = (int)((n + 7) / 8);
    if (blocks > 148 * 8) blocks = 148 * 8;
    finalize_rows_kernel<<<blocks, 256, 0, stream>>>(mat, dim, rows, slots, ids, n, inv_norm, doc_slot, chunk_ids);
    return cudaGetLastError();
}

cudaError_t launch_validate_stage(const float *stage, uint64_t n_floats, uint64_t base, unsigned long long *first_bad,
                                  cudaStream_t stream)
{
    if (n_floats == 0) return cudaSuccess;
    validate_stage_kernel<<<148 * 4, 256, 0, stream>>>(stage, n_floats, base, first_bad);
    return cudaGetLastError();
}

cudaError_t launch_commit_rows(const float *stage, int dim, const uint32_t *rows, const uint32_t *slots, const uint8_t *ids,
                               uint32_t n, float *mat, float *inv_norm, uint32_t *doc_slot, uint8_t *chunk_ids,
                               cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    int blocks = (int)((n + 7) / 8);
    if (blocks > 148 * 8) blocks = 148 * 8;
    commit_rows_kernel<<<blocks, 256, 0, stream>>>(stage, dim, rows, slots, ids, n, mat, inv_norm, doc_slot, chunk_ids);
    return cudaGetLastError();
}

cudaError_t launch_query_inv_norm(const float *queries, int dim, int nq, float *out, cudaStream_t stream)
{
    if (nq <= 0) {              // nothing to compute: only make sure the (lazily loaded) kernel is resident
        cudaFuncAttributes fa;
        return cudaFuncGetAttributes(&fa, query_inv_norm_kernel);
    }
    int blocks = (nq + 7) / 8;
    query_inv_norm_kernel<<<blocks, 256, 0, stream>>>(queries, dim, nq, out);
    return cudaGetLastError();
}

cudaError_t launch_synth(const SynthArgs &a, float *mat, float *inv_norm, uint32_t *doc_slot,
                         uint8_t *chunk_ids, cudaStream_t stream)
{
    if (a.n_rows == 0) return cudaSuccess;
    long long blocks = ((long long)a.n_rows + 7) / 8;
    if (blocks > 148 * 16) blocks = 148 * 16;
    synth_kernel<<<(int)blocks, 256, 0, stream>>>(a, mat, inv_norm, doc_slot, chunk_ids);
    return cudaGetLastError();
}

// Host statement of one synthetic vector (queries for benches/tests): same integer-exact spec as synth_kernel.
void synth_vector_host(uint64_t seed_noise, uint64_t noise_index, uint64_t seed_centroid, uint64_t cluster,
                       int cmul, int nmul, int dim, float *out, int smul, uint32_t super_group)
{
    const unsigned long long kn = stream_key(seed_noise, noise_index), kc = stream_key(seed_centroid, cluster);
    const bool two_level = smul != 0 && super_group != 0;
    const unsigned long long ks = two_level ? stream_key(seed_centroid ^ kSuperTag, cluster / super_group) : 0ull;
    long long ss = 0;
    for (int c = 0; c < dim; c++) {
        const unsigned long long hn = mix64(kn + (unsigned long long)((c >> 1) + 1) * kPair);
        const unsigned long long hc = mix64(kc + (unsigned long long)((c >> 1) + 1) * kPair);
        const uint32_t wn = (c & 1) ? (uint32_t)(hn >> 32) : (uint32_t)hn;
        const uint32_t wc = (c & 1) ? (uint32_t)(hc >> 32) : (uint32_t)hc;
        int u = cmul * ih4(wc) + nmul * ih4(wn);
        if (two_level) {
            const unsigned long long hs = mix64(ks + (unsigned long long)((c >> 1) + 1) * kPair);
            u += smul * ih4((c & 1) ? (uint32_t)(hs >> 32) : (uint32_t)hs);
        }
        ss += (long long)u * u;
        out[c] = (float)u;
    }
    if (ss == 0) return;
    const float mag = (float)sqrt((double)ss);
    for (int c = 0; c < dim; c++) out[c] /= mag;
}

void synth_id_host(uint64_t seed_ids, int kind, uint64_t index, uint8_t out[16])
{
    unsigned long long a, b;
    id16(seed_ids, kind == 0 ? 0xC4A9000000000000ull : 0xD0C0000000000000ull, index, a, b);
    for (int i = 0; i < 8; i++) { out[i] = (uint8_t)(a >> (8 * i)); out[8 + i] = (uint8_t)(b >> (8 * i)); }
}

void synth_doc_id(uint64_t seed_ids, uint64_t doc, uint8_t out[16])
{
    unsigned long long a, b;
    id16(seed_ids, 0xD0C0000000000000ull, doc, a, b);
    for (int i = 0; i < 8; i++) { out[i] = (uint8_t)(a >> (8 * i)); out[8 + i] = (uint8_t)(b >> (8 * i)); }
}

}  // namespace da
