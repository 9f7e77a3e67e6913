// IVFFlat kernels: the index the reference asks Postgres for with
//   CREATE INDEX ... USING ivfflat (vector vector_cosine_ops) WITH (lists = 100)
// (internal/store/postgres.go:95-99).  pgvector builds it by k-means on a sample and then
// assigns every row to its nearest centroid; a query ranks the centroids and scans the
// `probes` nearest lists.  Here:
//   - k-means is spherical (centroids kept at unit length, nearest = largest dot product);
//   - "assign rows to centroids" is a GEMM (rows x centroids^T) fused with a per-row argmax:
//     the score matrix is never written;
//   - a list is an array of row numbers -- rows stay where they are in HBM (a 12 KB row is
//     read contiguously anyway), so building or refreshing the index moves no vectors;
//   - a query = score the centroids (fp32 here for small batches -- one launch with the selection for a single
//     query --, the tcgen05 GEMM of da_tc.cu for batches), keep the `probes` best lists by an exact radix select, and
//     ONE probe-mode launch of the fused scan (da_scan.cu) in which every query walks its own lists.
#include "da_internal.h"

#include <algorithm>

namespace da {

namespace {

constexpr int BM = 128, BN = 128, BK = 16, TM = 8, TN = 8;
constexpr int kAssignThreads = 256;

__host__ __device__ inline unsigned long long ivf_mix64(unsigned long long z)
{
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

// SIMT fp32 GEMM + row argmax.  out[i] = argmax_j  <row_i, centroid_j>.
// rows are mat[row_idx[i]] when row_idx != nullptr, else mat[i].  The result for item i is written
// to out_assign[out_by_row ? source row : i].
__global__ void __launch_bounds__(kAssignThreads) assign_simt_kernel(const float *__restrict__ mat, int D,
                                                                     const uint32_t *__restrict__ row_idx, uint32_t n,
                                                                     const float *__restrict__ cent, int L,
                                                                     uint32_t *__restrict__ out_assign, int out_by_row)
{
    __shared__ __align__(16) float As[BK][BM + 4];
    __shared__ __align__(16) float Bs[BK][BN + 4];
    __shared__ float red_val[BM][17];
    __shared__ int red_idx[BM][17];

    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const uint32_t m0 = blockIdx.x * BM;
    const int lr = tid >> 2, kq = tid & 3;     // loader: rows lr and lr + 64, float4 number kq of the BK slice

    const float *a_src[2];
    bool a_ok[2];
#pragma unroll
    for (int i = 0; i < 2; i++) {
        uint32_t g = m0 + lr + 64 * i;
        a_ok[i] = g < n;
        uint32_t src = a_ok[i] ? (row_idx ? row_idx[g] : g) : 0u;
        a_src[i] = mat + (size_t)src * D;
    }

    float best_val[TM];
    int best_idx[TM];
#pragma unroll
    for (int i = 0; i < TM; i++) { best_val[i] = -3.0e38f; best_idx[i] = 0; }

    for (int n0 = 0; n0 < L; n0 += BN) {
        float acc[TM][TN];
#pragma unroll
        for (int i = 0; i < TM; i++)
#pragma unroll
            for (int j = 0; j < TN; j++) acc[i][j] = 0.f;

        const float *b_src[2];
        bool b_ok[2];
#pragma unroll
        for (int i = 0; i < 2; i++) {
            int c = n0 + lr + 64 * i;
            b_ok[i] = c < L;
            b_src[i] = cent + (size_t)(b_ok[i] ? c : 0) * D;
        }

        for (int k0 = 0; k0 < D; k0 += BK) {
            const int k = k0 + kq * 4;
            const bool kin = k < D;
#pragma unroll
            for (int i = 0; i < 2; i++) {
                float4 va = (a_ok[i] && kin) ? *reinterpret_cast<const float4 *>(a_src[i] + k) : make_float4(0.f, 0.f, 0.f, 0.f);
                float4 vb = (b_ok[i] && kin) ? *reinterpret_cast<const float4 *>(b_src[i] + k) : make_float4(0.f, 0.f, 0.f, 0.f);
                const int r = lr + 64 * i;
                As[kq * 4 + 0][r] = va.x; As[kq * 4 + 1][r] = va.y; As[kq * 4 + 2][r] = va.z; As[kq * 4 + 3][r] = va.w;
                Bs[kq * 4 + 0][r] = vb.x; Bs[kq * 4 + 1][r] = vb.y; Bs[kq * 4 + 2][r] = vb.z; Bs[kq * 4 + 3][r] = vb.w;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < BK; kk++) {
                float a[TM], b[TN];
                const float4 a0 = *reinterpret_cast<const float4 *>(&As[kk][ty * TM]);
                const float4 a1 = *reinterpret_cast<const float4 *>(&As[kk][ty * TM + 4]);
                const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[kk][tx * TN]);
                const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[kk][tx * TN + 4]);
                a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
                b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
                for (int i = 0; i < TM; i++)
#pragma unroll
                    for (int j = 0; j < TN; j++) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            }
            __syncthreads();
        }
#pragma unroll
        for (int i = 0; i < TM; i++)
#pragma unroll
            for (int j = 0; j < TN; j++) {
                const int c = n0 + tx * TN + j;
                if (c < L && acc[i][j] > best_val[i]) { best_val[i] = acc[i][j]; best_idx[i] = c; }
            }
    }

    // reduce over the 16 threads (tx) that share a row; ties go to the lower centroid index
#pragma unroll
    for (int i = 0; i < TM; i++) { red_val[ty * TM + i][tx] = best_val[i]; red_idx[ty * TM + i][tx] = best_idx[i]; }
    __syncthreads();
    if (tid < BM) {
        const uint32_t g = m0 + tid;
        if (g < n) {
            float bv = red_val[tid][0];
            int bi = red_idx[tid][0];
            for (int t = 1; t < 16; t++) {
                float v = red_val[tid][t];
                int ix = red_idx[tid][t];
                if (v > bv || (v == bv && ix < bi)) { bv = v; bi = ix; }
            }
            const uint32_t dst = out_by_row ? (row_idx ? row_idx[g] : g) : g;
            out_assign[dst] = (uint32_t)bi;
        }
    }
}

// sample row numbers: evenly spaced with a hashed offset inside each stride (distinct rows)
__global__ void sample_rows_kernel(uint32_t n_rows, uint32_t n_sample, unsigned long long seed, uint32_t *out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_sample) return;
    const unsigned long long step = n_rows / n_sample;          // >= 1
    const unsigned long long off = step > 1 ? ivf_mix64(seed ^ (i + 1) * 0x9E3779B97F4A7C15ull) % step : 0;
    out[i] = (uint32_t)(i * step + off);
}

// centroid l starts as the unit-length copy of a sample row: stratum l of the sample (n_sample / L consecutive
// entries), hashed position inside it -- distinct rows, and no fixed stride that could beat against a period of the
// corpus (the evenly spaced picks of round 1 did, against 200-row documents dealt round-robin to clusters)
__global__ void init_centroids_kernel(const float *mat, int D, const float *inv_norm, const uint32_t *sample,
                                      uint32_t n_sample, int L, unsigned long long seed, float *cent)
{
    const int l = blockIdx.x;
    const unsigned long long lo = (unsigned long long)l * n_sample / L, hi = (unsigned long long)(l + 1) * n_sample / L;
    const unsigned long long span = hi > lo ? hi - lo : 1;
    const uint32_t row = sample[lo + ivf_mix64(seed ^ (0xC3A5C85C97CB3127ull * (unsigned long long)(l + 1))) % span];
    const float inv = inv_norm[row];
    for (int c = threadIdx.x; c < D; c += blockDim.x) cent[(size_t)l * D + c] = mat[(size_t)row * D + c] * inv;
}

// sums[assign[i]] += unit(row_i); counts[assign[i]] += 1      (one warp per sample row)
__global__ void kmeans_accumulate_kernel(const float *mat, int D, const float *inv_norm, const uint32_t *sample,
                                         uint32_t n_sample, const uint32_t *assign, float *sums, float *counts)
{
    const int lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = w; i < n_sample; i += nw) {
        const uint32_t row = sample[i];
        const uint32_t a = assign[i];
        const float inv = inv_norm[row];
        const float *src = mat + (size_t)row * D;
        float *dst = sums + (size_t)a * D;
        for (int c = lane; c < D; c += 32) atomicAdd(dst + c, src[c] * inv);
        if (lane == 0) atomicAdd(counts + a, 1.0f);
    }
}

// centroid = sums / |sums| (spherical k-means); an empty or degenerate cluster keeps its centroid
__global__ void kmeans_finalize_kernel(const float *sums, const float *counts, int D, float *cent)
{
    const int l = blockIdx.x;
    __shared__ float red[32];
    __shared__ float inv_s;
    float ss = 0.f;
    for (int c = threadIdx.x; c < D; c += blockDim.x) { float v = sums[(size_t)l * D + c]; ss += v * v; }
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xFFFFFFFFu, ss, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); i++) t += red[i];
        inv_s = (counts[l] > 0.f && t > 0.f) ? rsqrtf(t) : 0.f;
    }
    __syncthreads();
    if (inv_s == 0.f) return;
    for (int c = threadIdx.x; c < D; c += blockDim.x) cent[(size_t)l * D + c] = sums[(size_t)l * D + c] * inv_s;
}

__global__ void fill_ones_kernel(float *p, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = 1.0f;
}

__global__ void list_hist_kernel(const uint32_t *assign, uint32_t n, uint32_t *counts)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) atomicAdd(counts + assign[i], 1u);
}

// exclusive prefix sum of counts[0..L) into offsets[0..L]; cursor = copy of offsets.  One CTA.
__global__ void list_scan_kernel(const uint32_t *counts, int L, uint32_t *offsets, uint32_t *cursor)
{
    __shared__ uint32_t part[1024];
    const int t = threadIdx.x, T = blockDim.x;
    const int per = (L + T - 1) / T;
    uint32_t s = 0;
    for (int i = t * per; i < min(L, (t + 1) * per); i++) s += counts[i];
    part[t] = s;
    __syncthreads();
    if (t == 0) {
        uint32_t run = 0;
        for (int i = 0; i < T; i++) { uint32_t v = part[i]; part[i] = run; run += v; }
        offsets[L] = run;
    }
    __syncthreads();
    uint32_t run = part[t];
    for (int i = t * per; i < min(L, (t + 1) * per); i++) { offsets[i] = run; cursor[i] = run; run += counts[i]; }
}

__global__ void list_scatter_kernel(const uint32_t *assign, uint32_t n, uint32_t *cursor, uint32_t *list_rows)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        list_rows[atomicAdd(cursor + assign[i], 1u)] = i;
}

// ---------------------------------------------------------------------------
// Splitting heavy lists into empty ones (after the k-means iterations).
//
// Spherical k-means cannot keep two centroids inside one isotropic cluster in high dimension: the mean of more
// members is a better representative of EVERY member (its noise part is smaller), so the larger sibling wins
// rows from the smaller one until that one is empty -- measured in round 1: 2 716 of 4 096 lists empty on the
// BASELINE corpus, whole 9 800-row clusters in single lists.  An index wants balanced lists, so after the
// iterations the heaviest lists donate to the empty (or nearly empty) ones: donor centroid c and the off-centroid part
// d of one of its members become  normalize(c + d / sqrt(D))  and  normalize(c - d / sqrt(D)).
// The step is the point.  |d| / sqrt(D) is the spread of the donor's rows along ONE direction, i.e. the distance at
// which the two halves' own means would sit: the pair is the donor's 2-means split, symmetric (neither sibling has the
// larger-sibling advantage), and NEUTRAL towards every other list -- measured alternatives: a fixed step of 0.25
// renormalised costs 3 % of every score and the rows leave for starved stubs that score 1 % less than the donor did;
// the same step left un-normalised adds 0.25 |<x, d>| to everybody's score, and the pair robs its neighbours until
// a few lineages hold the corpus.  With the natural step both effects are second order (1e-5).  Between the two
// siblings the assignment is as fine as TF32 resolves (for an isotropic 3072-d cluster: a coin flip) -- which halves
// the list, the purpose.  The pair is NOT re-averaged (that would start the collapse again).

__global__ void count_assign_kernel(const uint32_t *assign, uint32_t n, float *counts)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) atomicAdd(counts + assign[i], 1.0f);
}

// One CTA.  keys (P = pow2 >= L entries of scratch): (count << 32 | ~list), sorted descending; then the j-th
// heaviest list is paired with the j-th list from the light end while the donor holds more than 1.5 x the mean
// and the receiver less than a quarter of it (an empty list, or the stub a starved sibling has shrunk to: a few
// rows around its seed, which simply join their next-nearest list).  pairs[2j] = donor, pairs[2j+1] = receiver;
// *n_pairs = number of pairs.
__global__ void __launch_bounds__(1024) plan_splits_kernel(const float *counts, int L, int P, unsigned long long *keys,
                                                          uint32_t *pairs, uint32_t *n_pairs, float heavy_x, float light_x)
{
    __shared__ float total_s;
    __shared__ uint32_t n_s;
    const int tid = threadIdx.x;
    if (tid == 0) { total_s = 0.f; n_s = 0u; }
    __syncthreads();
    float part = 0.f;
    for (int i = tid; i < P; i += blockDim.x) {
        unsigned long long k = 0ull;
        if (i < L) { const float c = counts[i]; part += c; k = ((unsigned long long)(uint32_t)c << 32) | (unsigned long long)(~(uint32_t)i); }
        keys[i] = k;
    }
    atomicAdd(&total_s, part);
    for (int size = 2; size <= P; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int t = tid; t < (P >> 1); t += blockDim.x) {
                const int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                const bool desc = ((lo & size) == 0);
                const unsigned long long a = keys[lo], b = keys[hi];
                if ((a < b) == desc) { keys[lo] = b; keys[hi] = a; }
            }
        }
    __syncthreads();
    const float heavy = heavy_x * total_s / (float)L, light = light_x * total_s / (float)L;
    for (int j = tid; j < L / 2; j += blockDim.x) {
        const unsigned long long d = keys[j], r = keys[L - 1 - j];
        if ((float)(uint32_t)(d >> 32) > heavy && (float)(uint32_t)(r >> 32) < light) {
            // pairs are a prefix of j (counts descend from the head, ascend from the tail), so slot j is dense
            pairs[2 * j] = ~(uint32_t)(d & 0xFFFFFFFFull);
            pairs[2 * j + 1] = ~(uint32_t)(r & 0xFFFFFFFFull);
            atomicAdd(&n_s, 1u);
        }
    }
    __syncthreads();
    if (tid == 0) *n_pairs = n_s;
}

// Block j splits pair j.  The direction comes from the first member of the donor list found in the sample from a
// hashed start (a heavy list holds many sample rows, so the walk is short).
__global__ void __launch_bounds__(256) split_lists_kernel(const uint32_t *pairs, const uint32_t *n_pairs, const float *mat, int D,
                                                          const float *inv_norm, const uint32_t *sample, uint32_t n_sample,
                                                          const uint32_t *assign, unsigned long long seed, float *cent)
{
    if (blockIdx.x >= *n_pairs) return;
    const uint32_t donor = pairs[2 * blockIdx.x], recv = pairs[2 * blockIdx.x + 1];
    __shared__ uint32_t found_s;
    __shared__ float red[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) found_s = 0xFFFFFFFFu;
    __syncthreads();
    const uint32_t start = (uint32_t)(ivf_mix64(seed ^ (0x9E3779B97F4A7C15ull * (donor + 1ull))) % n_sample);
    for (uint32_t base = 0; base < n_sample; base += blockDim.x) {
        const uint32_t off = base + (uint32_t)tid;
        if (off < n_sample && assign[(start + off) % n_sample] == donor) atomicMin(&found_s, off);
        __syncthreads();
        const uint32_t f = found_s;
        __syncthreads();                                     // everybody has read it before the next round may write
        if (f != 0xFFFFFFFFu) break;
    }
    if (found_s == 0xFFFFFFFFu) return;                      // no member in this sample: leave the pair as it is
    const uint32_t row = sample[(start + found_s) % n_sample];
    const float inv = inv_norm[row];
    const float *x = mat + (size_t)row * D;
    float *c = cent + (size_t)donor * D, *e = cent + (size_t)recv * D;
    auto block_sum = [&](float v) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
        __syncthreads();
        if (lane == 0) red[warp] = v;
        __syncthreads();
        float t = 0.f;
        for (int i = 0; i < 8; i++) t += red[i];
        return t;
    };
    float dot = 0.f, cc = 0.f;
    for (int i = tid; i < D; i += blockDim.x) { const float ci = c[i]; dot += x[i] * inv * ci; cc += ci * ci; }
    dot = block_sum(dot);
    cc = block_sum(cc);
    if (!(cc > 0.f)) return;
    const float proj = dot / cc;                             // the member's component along the centroid
    float dd = 0.f;                                          // |x/|x| - proj c|^2
    for (int i = tid; i < D; i += blockDim.x) { const float d = x[i] * inv - proj * c[i]; dd += d * d; }
    dd = block_sum(dd);
    if (!(dd > 1e-12f)) return;                              // the member IS the centroid: nothing to split along
    const float step = rsqrtf((float)D);                     // |d| / sqrt(D): the rows' spread along one direction
    const float scale = rsqrtf(cc + dd * step * step);       // both siblings unit length (d is orthogonal to c)
    for (int i = tid; i < D; i += blockDim.x) {
        const float ci = c[i], d = (x[i] * inv - proj * ci) * step;
        e[i] = (ci + d) * scale;
        c[i] = (ci - d) * scale;
    }
}

// ---------------------------------------------------------------------------
// Query -> centroid scores for small batches (fewer than 8 queries; larger ones take the tensor-core GEMM of da_tc.cu):
// out[q][l] = <queries[q], cent[l]>, fp32, one warp per centroid, up to NQ queries per pass staged in shared memory.
// Nothing but the scores: the selection of the `probes` best lists is select_probes_kernel's, as for the GEMM.  (Round 2
// first ranked the centroids with the fused top-k scan: its per-CTA sorts and last-CTA merge of 128 x 32 keys cost
// 45-70 us for a matrix that streams in 7 -- measured, profiles/README.md.)  The centroid matrix is read with plain
// loads so that it stays in L2 between queries (the row scans stream with evict-first).
// ---------------------------------------------------------------------------
template <int NQ>
__global__ void __launch_bounds__(512) centroid_scores_kernel(const float *__restrict__ cent, int L, int D,
                                                              const float *__restrict__ queries, int nq, float *__restrict__ out)
{
    extern __shared__ __align__(16) unsigned char cs_smem[];
    float4 *q_s = reinterpret_cast<float4 *>(cs_smem);                 // NQ x D/4
    const int D4 = D >> 2, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < NQ * D4; i += blockDim.x)
        q_s[i] = (i / D4 < nq) ? reinterpret_cast<const float4 *>(queries)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    for (int l = blockIdx.x * (blockDim.x >> 5) + warp; l < L; l += gridDim.x * (blockDim.x >> 5)) {
        const float4 *row = reinterpret_cast<const float4 *>(cent) + (size_t)l * D4;
        float acc[NQ];
#pragma unroll
        for (int q = 0; q < NQ; q++) acc[q] = 0.f;
        for (int c0 = lane; c0 < D4; c0 += 32 * 8) {
            float4 x[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int c = c0 + 32 * u;
                x[u] = c < D4 ? __ldg(row + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int c = c0 + 32 * u < D4 ? c0 + 32 * u : 0;          // x is zero there
#pragma unroll
                for (int q = 0; q < NQ; q++) {
                    const float4 w = q_s[q * D4 + c];
                    acc[q] = fmaf(x[u].x, w.x, acc[q]); acc[q] = fmaf(x[u].y, w.y, acc[q]);
                    acc[q] = fmaf(x[u].z, w.z, acc[q]); acc[q] = fmaf(x[u].w, w.w, acc[q]);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < NQ; q++) {
            float v = acc[q];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
            if (lane == 0 && q < nq) out[(size_t)q * L + l] = v;
        }
    }
}

// ---------------------------------------------------------------------------
// Batched query -> centroid ranking: the scores come from the tensor-core GEMM (da_tc.cu: queries x centroids^T,
// ks partial sums over slices of the dimension), this kernel adds the partials in a fixed order and keeps each
// query's `probes` best lists.  One CTA per query; the L scores sit in shared memory as order-preserving integers;
// a 4 x 8-bit radix select finds the probes-th largest value exactly; lists above it, then the lowest-numbered
// lists equal to it, are emitted in list order (deterministic).
// ---------------------------------------------------------------------------
__device__ void select_probes_cta(const float *partial, int ks, int batch, int q, int L, int probes, key_t64 *out_keys,
                                  uint32_t *sc /* L words of shared memory */)
{
    __shared__ uint32_t hist[256];
    __shared__ uint32_t sel_prefix, sel_mask, sel_remaining, sel_done;
    __shared__ uint32_t scan_gt[16], scan_eq[16];
    const int tid = threadIdx.x;
    for (int i = tid; i < L; i += blockDim.x) {
        float s = 0.f;
        for (int j = 0; j < ks; j++) s += __ldcg(&partial[((size_t)j * batch + q) * L + i]);
        sc[i] = score_to_ordered(s);
    }
    const int want = probes < L ? probes : L;
    const int lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { sel_prefix = 0u; sel_mask = 0u; sel_remaining = (uint32_t)want; sel_done = 0u; }
    for (int shift = 24; shift >= 0; shift -= 8) {
        if (tid < 256) hist[tid] = 0u;
        __syncthreads();
        const uint32_t prefix = sel_prefix, mask = sel_mask;
        // scores of one query crowd into a few bins (the top byte is the same for almost all of them): lanes that
        // hit the same bin are counted together, one shared-memory atomic per distinct bin and warp
        for (int i0 = 0; i0 < L; i0 += blockDim.x) {
            const int i = i0 + tid;
            const uint32_t v = i < L ? sc[i] : 0u;
            const bool in = i < L && (v & mask) == prefix;
            const uint32_t bin = in ? ((v >> shift) & 255u) : 256u;
            const uint32_t peers = __match_any_sync(0xFFFFFFFFu, bin);
            if (in && lane == __ffs(peers) - 1) atomicAdd(&hist[bin], (uint32_t)__popc(peers));
        }
        __syncthreads();
        if (warp == 0) {
            // which bin, counted from the top, holds the remaining-th value: lane l owns bins 255-8l .. 248-8l
            uint32_t mine[8], tot = 0;
#pragma unroll
            for (int j = 0; j < 8; j++) { mine[j] = hist[255 - 8 * lane - j]; tot += mine[j]; }
            uint32_t incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                if (lane >= o) incl += t;
            }
            const uint32_t rem = sel_remaining, before = incl - tot;      // values in the bins above this lane's
            const bool here = before < rem && rem <= incl;
            const uint32_t found = __ballot_sync(0xFFFFFFFFu, here);
            if (found == 0u) {                       // fewer matching values than wanted (cannot happen: want <= L)
                if (lane == 0) { sel_prefix = prefix; sel_mask = mask | (255u << shift); }
            } else if (here) {
                uint32_t r = rem - before, b = 0;
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (b == 0u && r <= mine[j]) b = 0x100u | (uint32_t)(255 - 8 * lane - j);   // flag + bin
                    else if (b == 0u) r -= mine[j];
                }
                sel_remaining = r;
                sel_prefix = prefix | ((b & 255u) << shift);
                sel_mask = mask | (255u << shift);
                // The whole bin is wanted (nearly always so after two passes: the bin then holds a single value): every
                // value >= the bin's lowest possible one is selected, the lower digits need not be resolved.
                if (r == hist[b & 255u]) sel_done = 1u;
            }
        }
        __syncthreads();
        if (sel_done) break;
    }
    const uint32_t T = sel_prefix, take_eq = sel_remaining;     // the want-th largest value; how many of its equals to keep
    const int per = (L + (int)blockDim.x - 1) / (int)blockDim.x;
    const int i0 = tid * per, i1 = min(L, i0 + per);
    uint32_t gt = 0, eq = 0;
    for (int i = i0; i < i1; i++) { const uint32_t v = sc[i]; gt += v > T; eq += v == T; }
    // exclusive prefix over the threads (in thread order = list order): warp scans, then the warps' totals
    uint32_t ig = gt, ie = eq;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t tg = __shfl_up_sync(0xFFFFFFFFu, ig, o), te = __shfl_up_sync(0xFFFFFFFFu, ie, o);
        if (lane >= o) { ig += tg; ie += te; }
    }
    if (lane == 31) { scan_gt[warp] = ig; scan_eq[warp] = ie; }
    __syncthreads();
    uint32_t base_g = 0, base_e = 0, n_gt = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) {
        if (w < warp) { base_g += scan_gt[w]; base_e += scan_eq[w]; }
        n_gt += scan_gt[w];
    }
    uint32_t pg = base_g + ig - gt, pe = base_e + ie - eq;
    key_t64 *out = out_keys + (size_t)q * probes;
    for (int i = i0; i < i1; i++) {
        const uint32_t v = sc[i];
        const key_t64 key = ((key_t64)v << 32) | (key_t64)(~(uint32_t)i);
        if (v > T) out[pg++] = key;
        else if (v == T) { if (pe < take_eq) out[n_gt + pe] = key; pe++; }
    }
    for (int i = want + tid; i < probes; i += blockDim.x) out[i] = 0ull;
}

__global__ void __launch_bounds__(512) select_probes_kernel(const float *partial, int ks, int batch, int L, int probes,
                                                            key_t64 *out_keys)
{
    extern __shared__ uint32_t sel_sc[];                     // L ordered scores
    select_probes_cta(partial, ks, batch, (int)blockIdx.x, L, probes, out_keys, sel_sc);
}

// One query: scores and selection in ONE launch.  The CTAs score the centroids as centroid_scores_kernel does; the
// last one to finish (ticket) ranks the L scores, which it finds in L2, in the shared memory the query no longer needs.
// (As two kernels the selection paid a launch and a cold start: 13 us of a 44 us probes = 1 search.)
__global__ void __launch_bounds__(512) rank_one_query_kernel(const float *__restrict__ cent, int L, int D, const float *__restrict__ query,
                                                             float *scores, int probes, key_t64 *out_keys, unsigned int *ticket)
{
    extern __shared__ __align__(16) unsigned char rk_smem[];
    __shared__ unsigned int rk_last;
    float4 *q_s = reinterpret_cast<float4 *>(rk_smem);
    const int D4 = D >> 2, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < D4; i += blockDim.x) q_s[i] = reinterpret_cast<const float4 *>(query)[i];
    __syncthreads();
    for (int l = blockIdx.x * (blockDim.x >> 5) + warp; l < L; l += gridDim.x * (blockDim.x >> 5)) {
        const float4 *row = reinterpret_cast<const float4 *>(cent) + (size_t)l * D4;
        float acc = 0.f;
        for (int c0 = lane; c0 < D4; c0 += 32 * 8) {
            float4 x[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int c = c0 + 32 * u;
                x[u] = c < D4 ? __ldg(row + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int c = c0 + 32 * u < D4 ? c0 + 32 * u : 0;          // x is zero there
                const float4 w = q_s[c];
                acc = fmaf(x[u].x, w.x, acc); acc = fmaf(x[u].y, w.y, acc);
                acc = fmaf(x[u].z, w.z, acc); acc = fmaf(x[u].w, w.w, acc);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
        if (lane == 0) scores[l] = acc;
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) rk_last = atomicAdd(ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
    __syncthreads();
    if (!rk_last) return;
    __threadfence();
    select_probes_cta(scores, 1, 1, 0, L, probes, out_keys, reinterpret_cast<uint32_t *>(rk_smem));
    if (tid == 0) *ticket = 0u;
}

}  // namespace

cudaError_t ivf_launch_assign(const float *mat, int D, const uint32_t *row_idx, uint32_t n, const float *cent, int L,
                              uint32_t *out_assign, int out_by_row, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    assign_simt_kernel<<<(n + BM - 1) / BM, kAssignThreads, 0, stream>>>(mat, D, row_idx, n, cent, L, out_assign, out_by_row);
    return cudaGetLastError();
}

cudaError_t ivf_launch_sample(uint32_t n_rows, uint32_t n_sample, uint64_t seed, uint32_t *out, cudaStream_t stream)
{
    if (n_sample == 0) return cudaSuccess;
    sample_rows_kernel<<<(n_sample + 255) / 256, 256, 0, stream>>>(n_rows, n_sample, seed, out);
    return cudaGetLastError();
}

cudaError_t ivf_launch_init_centroids(const float *mat, int D, const float *inv_norm, const uint32_t *sample,
                                      uint32_t n_sample, int L, uint64_t seed, float *cent, float *cent_inv, cudaStream_t stream)
{
    init_centroids_kernel<<<L, 256, 0, stream>>>(mat, D, inv_norm, sample, n_sample, L, seed, cent);
    fill_ones_kernel<<<(L + 255) / 256, 256, 0, stream>>>(cent_inv, L);
    return cudaGetLastError();
}

cudaError_t ivf_launch_accumulate(const float *mat, int D, const float *inv_norm, const uint32_t *sample,
                                  uint32_t n_sample, const uint32_t *assign, float *sums, float *counts,
                                  cudaStream_t stream)
{
    if (n_sample == 0) return cudaSuccess;
    int blocks = (int)((n_sample + 7) / 8);
    if (blocks > 148 * 8) blocks = 148 * 8;
    kmeans_accumulate_kernel<<<blocks, 256, 0, stream>>>(mat, D, inv_norm, sample, n_sample, assign, sums, counts);
    return cudaGetLastError();
}

cudaError_t ivf_launch_finalize(const float *sums, const float *counts, int L, int D, float *cent, cudaStream_t stream)
{
    kmeans_finalize_kernel<<<L, 256, 0, stream>>>(sums, counts, D, cent);
    return cudaGetLastError();
}

cudaError_t ivf_launch_build_lists(const uint32_t *assign, uint32_t n, int L, uint32_t *counts, uint32_t *offsets,
                                   uint32_t *cursor, uint32_t *list_rows, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(counts, 0, (size_t)L * 4, stream);
    if (e != cudaSuccess) return e;
    if (n) list_hist_kernel<<<148 * 4, 256, 0, stream>>>(assign, n, counts);
    list_scan_kernel<<<1, 1024, 0, stream>>>(counts, L, offsets, cursor);
    if (n) list_scatter_kernel<<<148 * 4, 256, 0, stream>>>(assign, n, cursor, list_rows);
    return cudaGetLastError();
}

cudaError_t ivf_launch_count(const uint32_t *assign, uint32_t n, float *counts, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    count_assign_kernel<<<148 * 2, 256, 0, stream>>>(assign, n, counts);
    return cudaGetLastError();
}

size_t ivf_split_scratch_bytes(int L)
{
    int P = 1;
    while (P < L) P <<= 1;
    return (size_t)P * 8 + (size_t)L * 4 + 16;             // sort keys, (donor, receiver) pairs, pair count
}

cudaError_t ivf_launch_split_round(const float *counts, int L, void *scratch, const float *mat, int D, const float *inv_norm,
                                   const uint32_t *sample, uint32_t n_sample, const uint32_t *assign, uint64_t seed,
                                   float *cent, cudaStream_t stream, float heavy_x, float light_x)
{
    if (L < 2) return cudaSuccess;
    int P = 1;
    while (P < L) P <<= 1;
    unsigned long long *keys = static_cast<unsigned long long *>(scratch);
    uint32_t *pairs = reinterpret_cast<uint32_t *>(keys + P);
    uint32_t *n_pairs = pairs + (size_t)(L / 2) * 2;
    plan_splits_kernel<<<1, 1024, 0, stream>>>(counts, L, P, keys, pairs, n_pairs, heavy_x, light_x);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    split_lists_kernel<<<L / 2, 256, 0, stream>>>(pairs, n_pairs, mat, D, inv_norm, sample, n_sample, assign, seed, cent);
    return cudaGetLastError();
}

template <int NQ>
static cudaError_t launch_centroid_scores(const float *cent, int L, int D, const float *queries, int nq, float *out, int sm_count,
                                          cudaStream_t stream)
{
    const size_t smem = (size_t)NQ * D * 4;
    cudaError_t e = cudaFuncSetAttribute(centroid_scores_kernel<NQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int grid = (L + 15) / 16;
    if (grid > sm_count) grid = sm_count;
    centroid_scores_kernel<NQ><<<grid, 512, smem, stream>>>(cent, L, D, queries, nq, out);
    return cudaGetLastError();
}

// out[q][l] for q < nq <= 8 (one pass)
cudaError_t ivf_launch_centroid_scores(const float *cent, int L, int D, const float *queries, int nq, float *out, int sm_count,
                                       cudaStream_t stream)
{
    if (nq < 1 || nq > 8 || (D & 3) || (size_t)8 * D * 4 > 200 * 1024) return cudaErrorInvalidValue;
    if (nq <= 1) return launch_centroid_scores<1>(cent, L, D, queries, nq, out, sm_count, stream);
    if (nq <= 2) return launch_centroid_scores<2>(cent, L, D, queries, nq, out, sm_count, stream);
    if (nq <= 4) return launch_centroid_scores<4>(cent, L, D, queries, nq, out, sm_count, stream);
    return launch_centroid_scores<8>(cent, L, D, queries, nq, out, sm_count, stream);
}

size_t ivf_select_smem_bytes(int L) { return (size_t)L * 4; }

// One query: centroid scores + selection of the `probes` best lists in one launch (`ticket`: a zeroed word, left zero).
cudaError_t ivf_launch_rank_one_query(const float *cent, int L, int D, const float *query, float *scores, int probes,
                                      key_t64 *out_keys, unsigned int *ticket, int sm_count, cudaStream_t stream)
{
    if ((D & 3) || L < 1) return cudaErrorInvalidValue;
    const size_t smem = std::max((size_t)D * 4, ivf_select_smem_bytes(L));
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void *>(rank_one_query_kernel), smem);
    if (e != cudaSuccess) return e;
    int grid = (L + 15) / 16;
    if (grid > sm_count) grid = sm_count;
    rank_one_query_kernel<<<grid, 512, smem, stream>>>(cent, L, D, query, scores, probes, out_keys, ticket);
    return cudaGetLastError();
}

cudaError_t ivf_launch_select_probes(const float *partial, int ks, int batch, int L, int probes, key_t64 *out_keys,
                                     cudaStream_t stream)
{
    const size_t smem = ivf_select_smem_bytes(L);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(select_probes_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    select_probes_kernel<<<batch, 512, smem, stream>>>(partial, ks, batch, L, probes, out_keys);
    return cudaGetLastError();
}

}  // namespace da
