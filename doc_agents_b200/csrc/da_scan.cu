// Fused exact scan for sm_100a: cosine score + document mask + similarity floor + top-k
// in one pass over the resident embedding matrix.
//
// Replaces the per-row `e.vector <=> $1` loop, the `document_id = ANY($2)` filter, the
// `>= 0.7` floor, ORDER BY and LIMIT of the reference's one SQL statement
// (internal/store/postgres.go:225-243).  HBM-bound: 4*dim + 8 bytes per row, 2*dim flops.
//
// Shape of the kernel
//   - persistent grid (multiple of the SM count), 16 warps per CTA;
//   - a warp owns R consecutive rows at a time and streams them with 256-bit
//     ld.global.nc.L1::no_allocate.L2::evict_first loads (128-bit when dim % 8 != 0), R*U of them
//     in flight per lane;
//   - queries live in shared memory (up to 8 per pass, so a small batch re-uses every
//     row fetched from HBM);
//   - each warp keeps its own top-k (unsorted, with a cached minimum) in shared memory;
//     inserts are rare once the list is warm, so bookkeeping stays off the load path;
//   - CTA epilogue: bitonic sort of the warps' lists, k keys per query to global scratch;
//   - the last CTA to finish (ticket counter) merges all CTAs' keys, gathers the chunk
//     ids and writes the final records: one launch per pass, no host round trip.
#include "da_internal.h"
#include "da_topk.cuh"

#include <mutex>

namespace da {

namespace {

// INDIRECT: visit the rows listed in a.row_list (count in *a.n_rows_dev) instead of [0, n_rows):
// the probed IVFFlat lists.  A 12 KB row is still read contiguously, so gathering by row costs
// no bandwidth.
// VEC = elements per lane per load: 8 (256-bit loads of fp32, rows 32-byte aligned, dim % 8 == 0), 4, or 16:
// the split-plane layout, where a 256-bit load brings 16 upper halves (truncated bf16) of a row of a.hi and every
// row's approximate similarity is also stored to a.approx for the candidate pass (da_split.cu).
template <int NQ, int R, int U, int VEC, bool FULL_TILES, bool INDIRECT>
__global__ void __launch_bounds__(kScanThreads, 1) scan_topk_kernel(const ScanArgs a)
{
    DA_TS(0);
    ScanPerQuery pq = scan_per_query(a);
    // IVFFlat probe mode / range mode: this CTA serves query blockIdx.y; its rows are the concatenation of that
    // query's lists (or of the requested documents' row ranges)
    __shared__ uint32_t p_start[DA_IVF_MAX_PROBES + 1];       // first flat index of each piece (+ total)
    __shared__ uint32_t p_src[DA_IVF_MAX_PROBES];             // where the piece starts (in list_rows, or as a row number)
    const bool probe_mode = INDIRECT && (a.probe_keys != nullptr || a.ranges != nullptr);
    if (probe_mode) {
        const size_t qy = blockIdx.y;
        pq.queries += qy * a.dim;
        if (pq.q_inv_norm) pq.q_inv_norm += qy;
        if (pq.doc_bitmap) pq.doc_bitmap += qy * a.bitmap_stride;
        pq.bitmap_stride = 0;
        pq.cta_out += qy * gridDim.x * a.k;
        pq.counter += qy;
        pq.out_keys += qy * a.k;
        if (pq.out_ids) pq.out_ids += qy * a.k * 16;
        if (pq.probe_keys) pq.probe_keys += qy * a.probes;
        if (threadIdx.x < (unsigned)a.probes) {
            uint32_t off = 0, len = 0;
            if (pq.probe_keys) {
                const key_t64 key = pq.probe_keys[threadIdx.x];
                if (key) { const uint32_t l = key_row(key); off = a.list_offsets[l]; len = a.list_offsets[l + 1] - off; }
            } else {
                const uint2 r = a.ranges[threadIdx.x];
                off = r.x; len = r.y;
            }
            p_src[threadIdx.x] = off;
            p_start[threadIdx.x + 1] = len;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t run = 0;
            p_start[0] = 0;
            for (int p = 0; p < a.probes; p++) { run += p_start[p + 1]; p_start[p + 1] = run; }
        }
        // (the barrier after the query staging below publishes the table)
    }
    constexpr int V4 = VEC / 4;                 // float4 words of the query per lane slot
    constexpr int VR = VEC == 16 ? 2 : V4;      // float4 registers holding one loaded slot (raw 16-bit pairs when VEC == 16)
    const int DV = a.dim / VEC;                 // lane slots per row
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int D4 = a.dim >> 2;
    float4 *q_s = reinterpret_cast<float4 *>(smem_raw);                       // NQ x D4
    key_t64 *lists = reinterpret_cast<key_t64 *>(smem_raw + (size_t)NQ * D4 * sizeof(float4));  // NQ x P
    __shared__ float q_inv_s[NQ];
    __shared__ unsigned int is_last_s;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int K = a.k, P = a.P;

    // stage queries (rows beyond a.nq are zero so they never qualify) and clear the lists
    for (int i = tid; i < NQ * D4; i += kScanThreads) {
        int q = i / D4;
        q_s[i] = (q < a.nq) ? reinterpret_cast<const float4 *>(pq.queries)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int i = tid; i < NQ * P; i += kScanThreads) lists[i] = 0ull;
    if (pq.q_inv_norm && tid < NQ) q_inv_s[tid] = (tid < a.nq) ? pq.q_inv_norm[tid] : 0.f;
    __syncthreads();
    if (!pq.q_inv_norm) {        // 1/|q| from the staged copy, with query_inv_norm_kernel's arithmetic (one warp per query)
        if (warp < NQ) {
            const float inv = warp < a.nq ? warp_inv_norm(reinterpret_cast<const float *>(q_s) + (size_t)warp * a.dim, a.dim, lane) : 0.f;
            if (lane == 0) q_inv_s[warp] = inv;
        }
        __syncthreads();
    }

    int cnt[NQ];
    key_t64 wmin[NQ];
    uint32_t live_q = 0u;               // queries that can match at all: a zero query (1/|q| = 0) matches nothing, whatever the floor
#pragma unroll
    for (int q = 0; q < NQ; q++) { cnt[q] = 0; wmin[q] = 0ull; live_q |= (q_inv_s[q] > 0.f ? 1u : 0u) << q; }
    DA_TS(1);

    const uint32_t n_rows = probe_mode ? p_start[a.probes] : INDIRECT ? *a.n_rows_dev : a.n_rows;
    const uint32_t n_groups = (n_rows + R - 1) / R;
    const uint32_t gstride = gridDim.x * kScanWarps;
    unsigned int my_rows_read = 0;

    // What a lane learns about "its" row: 1/|row| and, per query, whether the row's document is requested.
    auto row_facts = [&](uint32_t my_row, bool valid, uint32_t &want_bits, float &my_inv) {
        want_bits = 0u;                          // bit q: this lane's row is requested by query q
        my_inv = 0.f;
        if (valid) {
            my_inv = a.inv_norm[my_row];
            if (pq.doc_bitmap) {
                const uint32_t slot = a.doc_slot[my_row];
                if (pq.bitmap_stride == 0) {
                    want_bits = ((pq.doc_bitmap[slot >> 5] >> (slot & 31)) & 1u) ? 0xFFFFFFFFu : 0u;
                } else {
#pragma unroll
                    for (int q = 0; q < NQ; q++)
                        if (q < a.nq)
                            want_bits |= ((pq.doc_bitmap[(size_t)q * pq.bitmap_stride + (slot >> 5)] >> (slot & 31)) & 1u) << q;
                }
            } else {
                want_bits = 0xFFFFFFFFu;
            }
            want_bits &= live_q;
        }
    };

    // One group of R rows starting at row0; lanes 0..R-1 carry the facts of rows row0 .. row0+R-1.
    auto process_group = [&](const uint32_t row0, const uint32_t my_row, const bool valid, const uint32_t want_bits,
                             const float my_inv) {
        const uint32_t valid_mask = __ballot_sync(0xFFFFFFFFu, valid);
        const uint32_t want_mask = __ballot_sync(0xFFFFFFFFu, want_bits != 0u);
        const uint32_t load_mask = a.read_all ? valid_mask : want_mask;
        if (load_mask == 0) {
            if constexpr (VEC == 16) {              // skipped rows still need an (impossible) approximate score
                if (valid)
                    for (int q = 0; q < a.nq; q++) a.approx[(size_t)q * n_rows + my_row] = -3.0e38f;
            }
            return;
        }
        my_rows_read += __popc(load_mask);

        float acc[R][NQ];
#pragma unroll
        for (int r = 0; r < R; r++)
#pragma unroll
            for (int q = 0; q < NQ; q++) acc[r][q] = 0.f;

        const float4 *rp = reinterpret_cast<const float4 *>(a.mat) + (size_t)row0 * D4;
        const float4 *rpi[R];
        if (INDIRECT) {
#pragma unroll
            for (int r = 0; r < R; r++)
                rpi[r] = reinterpret_cast<const float4 *>(a.mat) + (size_t)__shfl_sync(0xFFFFFFFFu, my_row, r) * D4;
        }

        for (int c0 = 0; c0 < DV; c0 += 32 * U) {
            float4 v[R][U][VR];
#pragma unroll
            for (int r = 0; r < R; r++) {
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const int c = c0 + u * 32 + lane;
                    const bool ok = ((load_mask >> r) & 1u) && (FULL_TILES || c < DV);
                    const float4 *src = (INDIRECT ? rpi[r] : rp + (size_t)r * D4) + (size_t)c * V4;
                    if (ok) {
                        if constexpr (VEC == 16) {
                            // 16 upper halves, kept raw (element 2j in the low 16 bits of word j, 2j+1 in the high 16 bits)
                            ldg_stream8(reinterpret_cast<const float4 *>(a.hi + ((size_t)(row0 + r) * a.dim + (size_t)c * 16)),
                                        v[r][u][0], v[r][u][1]);
                        } else if constexpr (VEC == 8) ldg_stream8(src, v[r][u][0], v[r][u][1]);
                        else v[r][u][0] = ldg_stream(src);
                    } else {
#pragma unroll
                        for (int i = 0; i < VR; i++) v[r][u][i] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                int c = c0 + u * 32 + lane;
                if (!FULL_TILES && c >= DV) c = 0;   // v is zero there; any in-range q slot will do
#pragma unroll
                for (int q = 0; q < NQ; q++) {
#pragma unroll
                    for (int i = 0; i < V4; i++) {
                        const float4 qv = q_s[q * D4 + c * V4 + i];
#pragma unroll
                        for (int r = 0; r < R; r++) {
                            float4 x;
                            if constexpr (VEC == 16) {          // widen the two raw words that hold elements 4i .. 4i+3
                                const float4 raw = v[r][u][i >> 1];
                                const uint32_t w0 = __float_as_uint((i & 1) ? raw.z : raw.x), w1 = __float_as_uint((i & 1) ? raw.w : raw.y);
                                x = make_float4(__uint_as_float(w0 << 16), __uint_as_float(w0 & 0xFFFF0000u),
                                                __uint_as_float(w1 << 16), __uint_as_float(w1 & 0xFFFF0000u));
                            } else {
                                x = v[r][u][i];
                            }
                            acc[r][q] = fmaf(x.x, qv.x, acc[r][q]);
                            acc[r][q] = fmaf(x.y, qv.y, acc[r][q]);
                            acc[r][q] = fmaf(x.z, qv.z, acc[r][q]);
                            acc[r][q] = fmaf(x.w, qv.w, acc[r][q]);
                        }
                    }
                }
            }
        }

#pragma unroll
        for (int r = 0; r < R; r++)
#pragma unroll
            for (int q = 0; q < NQ; q++) {
                float s = acc[r][q];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
                acc[r][q] = s;
            }

#pragma unroll
        for (int r = 0; r < R; r++) {
            const float inv = __shfl_sync(0xFFFFFFFFu, my_inv, r);
            const uint32_t row = INDIRECT ? __shfl_sync(0xFFFFFFFFu, my_row, r) : row0 + r;
            const uint32_t wb = __shfl_sync(0xFFFFFFFFu, want_bits, r);
            if constexpr (VEC == 16) {                                  // every visited row gets its approximate score
                if (lane < NQ && lane < a.nq && ((valid_mask >> r) & 1u)) {
                    float sv = -3.0e38f;
                    if (((wb >> lane) & 1u) && inv > 0.f) {
                        float t = 0.f;
#pragma unroll
                        for (int q = 0; q < NQ; q++) if (q == lane) t = acc[r][q];
                        sv = fminf(1.f, fmaxf(-1.f, t * inv * q_inv_s[lane]));
                    }
                    a.approx[(size_t)lane * n_rows + row] = sv;
                }
            }
            if (!wb || !(inv > 0.f)) continue;                          // zero-norm rows never match
#pragma unroll
            for (int q = 0; q < NQ; q++) {
                if (!((wb >> q) & 1u)) continue;                        // not in this query's document set
                float sim = acc[r][q] * inv * q_inv_s[q];
                sim = fminf(1.f, fmaxf(-1.f, sim));                     // pgvector clamps to [-1, 1]
                if (!(sim >= a.min_sim)) continue;                      // also drops NaN
                key_t64 key = make_key(sim, a.row_base + row);
                warp_list_insert(lists + (size_t)q * P + warp * K, K, cnt[q], wmin[q], key, lane);
            }
        }
    };

    if (!INDIRECT && pq.doc_bitmap && !a.read_all && n_rows >= 32768u) {
        // Document filter with rows skipped: the filter itself is the work when few rows qualify (the reference's
        // own case: one to sixteen documents), so a warp looks at 32 rows per step -- one per lane -- and hands
        // only the groups that contain a requested row to the row loop.  (Small stores keep the plain loop: there the
        // requested rows are most of the rows, and 32 of them in one warp's hands would serialise the search.)
        const uint32_t n_super = (n_rows + 31) / 32;
        constexpr uint32_t kGroupBits = R >= 32 ? 0xFFFFFFFFu : ((1u << R) - 1u);
        for (uint32_t sg = blockIdx.x * kScanWarps + warp; sg < n_super; sg += gstride) {
            const uint32_t row = sg * 32 + lane;
            const bool valid = row < n_rows;
            uint32_t want_bits;
            float inv;
            row_facts(row, valid, want_bits, inv);
            uint32_t wm = __ballot_sync(0xFFFFFFFFu, want_bits != 0u);
            if constexpr (VEC == 16) {              // rows of groups that are skipped still need an (impossible) score
                const uint32_t base = (uint32_t)(lane / R) * R;
                if (valid && !((wm >> base) & kGroupBits))
                    for (int q = 0; q < a.nq; q++) a.approx[(size_t)q * n_rows + row] = -3.0e38f;
            }
            while (wm) {
                const uint32_t base = (uint32_t)((__ffs(wm) - 1) / R) * R;
                wm &= ~(kGroupBits << base);
                const int src = (int)((base + lane) & 31u);
                const bool g_valid = __shfl_sync(0xFFFFFFFFu, valid ? 1 : 0, src) && lane < R;
                const uint32_t g_want = __shfl_sync(0xFFFFFFFFu, want_bits, src);
                const float g_inv = __shfl_sync(0xFFFFFFFFu, inv, src);
                process_group(sg * 32 + base, sg * 32 + base + lane, g_valid, lane < R ? g_want : 0u, lane < R ? g_inv : 0.f);
            }
        }
    } else {
        // The row number of lane r's row of a group: itself, or looked up (compacted filter list, IVFFlat lists, ranges).
        auto resolve = [&](uint32_t flat, bool valid) -> uint32_t {
            if (!INDIRECT) return flat;
            if (!valid) return 0u;
            if (!probe_mode) return a.row_list[flat];
            int lo = 0, hi = a.probes - 1;             // which piece holds the flat index: last p with p_start[p] <= flat
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (p_start[mid] <= flat) lo = mid; else hi = mid - 1;
            }
            const uint32_t r = p_src[lo] + (flat - p_start[lo]);
            return a.list_rows ? a.list_rows[r] : r;
        };
        // The look-up of the NEXT group is issued before the current group's rows are streamed, so its latency (a
        // dependent load in front of every group: binary search, list entry, then the row) hides behind them.
        uint32_t g = blockIdx.x * kScanWarps + warp;
        uint32_t nxt_row = 0u;
        if (g < n_groups) nxt_row = resolve(g * R + lane, lane < R && g * R + lane < n_rows);
        for (; g < n_groups; g += gstride) {
            const uint32_t row0 = g * R;
            // lane r (< R) looks at row0 + r: document mask and 1/|row|
            const bool valid = (lane < R) && (row0 + lane < n_rows);
            const uint32_t my_row = nxt_row;
            if (INDIRECT) {
                const uint32_t gn = g + gstride;
                if (gn < n_groups) nxt_row = resolve(gn * R + lane, lane < R && gn * R + lane < n_rows);
            } else {
                nxt_row = row0 + gstride * R + lane;
            }
            uint32_t want_bits;
            float my_inv;
            row_facts(my_row, valid, want_bits, my_inv);
            process_group(row0, my_row, valid, want_bits, my_inv);
        }
    }

    if (a.rows_read) {
        unsigned int tot = my_rows_read;   // identical in every lane of the warp
        if (lane == 0 && tot) atomicAdd(a.rows_read, (unsigned long long)tot);
    }

    scan_cta_epilogue(a, pq, lists, &is_last_s, reinterpret_cast<key_t64 *>(smem_raw),
                      (int)(((size_t)NQ * D4 * sizeof(float4) + (size_t)NQ * P * sizeof(key_t64)) / sizeof(key_t64)));
}


template <int NQ, int R, int U, int VEC>
cudaError_t launch_one(const ScanArgs &a, int grid, size_t smem, cudaStream_t stream)
{
    const bool full = ((a.dim / VEC) % (32 * U)) == 0;
    auto kern = full ? scan_topk_kernel<NQ, R, U, VEC, true, false> : scan_topk_kernel<NQ, R, U, VEC, false, false>;
    if (a.row_list || a.probe_keys || a.ranges) {
        if constexpr (VEC != 16) kern = full ? scan_topk_kernel<NQ, R, U, VEC, true, true> : scan_topk_kernel<NQ, R, U, VEC, false, true>;
        else return cudaErrorInvalidValue;
    }
    cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void *>(kern), smem);
    if (e != cudaSuccess) return e;
    kern<<<dim3((unsigned)grid, a.probe_keys ? (unsigned)a.probe_batch : 1u), kScanThreads, smem, stream>>>(a);
    return cudaGetLastError();
}

constexpr size_t kSmemBudget = 220 * 1024;

}  // namespace

// cudaFuncSetAttribute costs a driver call per launch; a function's limit only ever needs raising, per device.
// The table is a fixed array: this is also on the path of da_scan_topk_noalloc, which must not allocate anything.
cudaError_t ensure_dynamic_smem(const void *func, size_t smem)
{
    struct Entry { const void *f; int dev; size_t smem; };
    constexpr int kEntries = 256;            // (scan variants) x (devices); beyond that the attribute is simply set every time
    static std::mutex mu;
    static Entry seen[kEntries];
    static int n_seen = 0;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(mu);
    for (int i = 0; i < n_seen; i++)
        if (seen[i].f == func && seen[i].dev == dev) {
            if (seen[i].smem >= smem) return cudaSuccess;
            e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e == cudaSuccess) seen[i].smem = smem;
            return e;
        }
    e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess && n_seen < kEntries) seen[n_seen++] = Entry{func, dev, smem};
    return e;
}

size_t scan_smem_bytes(int dim, int nq_tmpl, int k, int *P_out)
{
    int P = next_pow2(kScanWarps * k);
    if (P < 2 * k) P = next_pow2(2 * k);
    if (P_out) *P_out = P;
    return (size_t)nq_tmpl * dim * sizeof(float) + (size_t)nq_tmpl * P * sizeof(key_t64);
}

static int nq_template_for(int nq) { return nq <= 1 ? 1 : nq <= 2 ? 2 : nq <= 4 ? 4 : 8; }

int scan_pick_nq(int dim, int k, int batch)
{
    int nq = batch < kMaxNQ ? batch : kMaxNQ;
    while (nq > 1 && scan_smem_bytes(dim, nq_template_for(nq), k, nullptr) > kSmemBudget) nq >>= 1;
    return nq;
}

int scan_grid_for(int sm_count) { return sm_count; }

// A probe-mode launch runs batch x this many CTAs (one CTA per SM at a time): a single query gets the whole chip,
// a large batch about four CTAs' worth of work per SM so that the tail of the last wave stays short.
int scan_probe_ctas_per_query(int sm_count, int batch)
{
    if (batch <= 1) return sm_count;
    const int c = (4 * sm_count + batch - 1) / batch;
    return c < 1 ? 1 : (c > sm_count ? sm_count : c);
}

cudaError_t scan_preload_single_query_variants()
{
    // the NQ = 1 row of launch_scan's table: 256-bit and 128-bit loads, whole and ragged tiles
    const void *fns[] = {(const void *)scan_topk_kernel<1, 2, 3, 8, true, false>, (const void *)scan_topk_kernel<1, 2, 3, 8, false, false>,
                         (const void *)scan_topk_kernel<1, 4, 4, 4, true, false>, (const void *)scan_topk_kernel<1, 4, 4, 4, false, false>};
    for (const void *f : fns) {
        cudaFuncAttributes fa;
        cudaError_t e = cudaFuncGetAttributes(&fa, f);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_scan(const ScanArgs &a_in, int sm_count, cudaStream_t stream, int *grid_out)
{
    ScanArgs a = a_in;
    const int nqt = nq_template_for(a.nq);
    size_t smem = scan_smem_bytes(a.dim, nqt, a.k, &a.P);
    if (smem > kSmemBudget || (a.dim & 3) || a.nq < 1 || a.nq > kMaxNQ) return cudaErrorInvalidValue;
    int grid = scan_grid_for(sm_count);
    if (a.probe_keys) {
        if (a.nq != 1 || a.hi || a.probes < 1 || a.probes > DA_IVF_MAX_PROBES || a.probe_batch < 1 || !a.list_offsets || !a.list_rows)
            return cudaErrorInvalidValue;
        grid = scan_probe_ctas_per_query(sm_count, a.probe_batch);
    } else if (a.ranges) {
        if (a.nq != 1 || a.hi || a.probes < 1 || a.probes > DA_IVF_MAX_PROBES) return cudaErrorInvalidValue;
        // n_rows carries the total of the ranges (known on the host): no more CTAs than its row groups need
        const long long groups = ((long long)a.n_rows + 1) / 2, need = (groups + kScanWarps - 1) / kScanWarps;
        if (need < grid) grid = need < 1 ? 1 : (int)need;
    } else if (!a.row_list) {   // small stores (the reference's own scale) and centroid scans: no more CTAs than row groups need
        const int r = 2;                                            // smallest R of any variant below
        const long long groups = ((long long)a.n_rows + r - 1) / r;
        const long long need = (groups + kScanWarps - 1) / kScanWarps;
        if (need < grid) grid = need < 1 ? 1 : (int)need;
    }
    if (grid_out) *grid_out = grid;
    if (a.hi) {   // split-plane store: the scan reads the upper plane only
        if (a.row_list || (a.dim % 16) || !a.approx) return cudaErrorInvalidValue;
        switch (nqt) {
            // measured at 10M x 3072: (R,U) = (4,2) 8.54 ms, (4,3) 8.68, (4,1) 8.94, (2,2) 10.4, (2,6) 11.2
            case 1: return launch_one<1, 4, 2, 16>(a, grid, smem, stream);
            case 2: return launch_one<2, 2, 2, 16>(a, grid, smem, stream);
            case 4: return launch_one<4, 2, 1, 16>(a, grid, smem, stream);
            default: return launch_one<8, 2, 1, 16>(a, grid, smem, stream);
        }
    }
    // 256-bit loads need 32-byte aligned rows; other dimensions keep the 128-bit path
    const bool v8 = (a.dim % 8) == 0 && (reinterpret_cast<uintptr_t>(a.mat) & 31) == 0;
    switch (nqt) {
        // (R, U) chosen by measurement on a B200 (10M x 3072): two rows per warp and three 256-bit loads
        // per lane in flight gave 7.33 TB/s; deeper unrolls (R*U = 12..16) were 2-4 % slower.
        case 1: return v8 ? launch_one<1, 2, 3, 8>(a, grid, smem, stream) : launch_one<1, 4, 4, 4>(a, grid, smem, stream);
        case 2: return v8 ? launch_one<2, 2, 3, 8>(a, grid, smem, stream) : launch_one<2, 4, 4, 4>(a, grid, smem, stream);
        case 4: return v8 ? launch_one<4, 2, 3, 8>(a, grid, smem, stream) : launch_one<4, 4, 2, 4>(a, grid, smem, stream);
        default: return v8 ? launch_one<8, 2, 1, 8>(a, grid, smem, stream) : launch_one<8, 2, 2, 4>(a, grid, smem, stream);
    }
}

// ---------------------------------------------------------------------------
// Merge of the all-gathered candidates: one CTA per query.
// gathered: G blocks of [batch*k ids of 16 B][batch*k keys of 8 B]
// ---------------------------------------------------------------------------
namespace {
constexpr int kMergeThreads = 256;

__global__ void __launch_bounds__(kMergeThreads) merge_kernel(const uint8_t *gathered, int G, int batch, int k,
                                                              int P, key_t64 *out_keys, uint8_t *out_ids,
                                                              size_t block_bytes, uint32_t *out_flags)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    key_t64 *keys = reinterpret_cast<key_t64 *>(smem_raw);        // P
    const int q = blockIdx.x;
    const int total = G * k;
    // trailer word of every rank's block (candidate-list overflow on that rank): every rank sees the OR, so the
    // decision to answer the batch another way is the same everywhere
    if (out_flags && q == 0 && threadIdx.x == 0) {
        uint32_t f = 0;
        for (int g = 0; g < G; g++)
            f |= *reinterpret_cast<const uint32_t *>(gathered + (size_t)g * block_bytes + (size_t)batch * k * 24);
        *out_flags = f;
    }
    for (int i = threadIdx.x; i < P; i += kMergeThreads) {
        key_t64 v = 0ull;
        if (i < total) {
            int g = i / k, j = i % k;
            v = reinterpret_cast<const key_t64 *>(gathered + (size_t)g * block_bytes + (size_t)batch * k * 16)[(size_t)q * k + j];
        }
        keys[i] = v;
    }
    cta_bitonic_desc(keys, P);
    // keys are unique across ranks (global rows), so the id of a winner is found by
    // looking its key up again in the owner's block.
    for (int i = threadIdx.x; i < k; i += kMergeThreads) {
        key_t64 key = keys[i];
        uint4 id = make_uint4(0, 0, 0, 0);
        if (key) {
            for (int g = 0; g < G; g++) {
                const key_t64 *kb = reinterpret_cast<const key_t64 *>(gathered + (size_t)g * block_bytes + (size_t)batch * k * 16) + (size_t)q * k;
                const uint4 *ib = reinterpret_cast<const uint4 *>(gathered + (size_t)g * block_bytes) + (size_t)q * k;
                for (int j = 0; j < k; j++)
                    if (kb[j] == key) id = ib[j];
            }
        }
        out_keys[(size_t)q * k + i] = key;
        reinterpret_cast<uint4 *>(out_ids)[(size_t)q * k + i] = id;
    }
}
}  // namespace

// ---------------------------------------------------------------------------
// Second and third stage of the batched tensor-core path (da_tc.cu produces the candidates).
// ---------------------------------------------------------------------------
namespace {


// Candidates below the query's final gate are dropped without touching their rows; the others get their
// exact fp32 score with the scan kernel's own arithmetic (same lane ownership, x-y-z-w FMA order, xor-shuffle
// tree, then * 1/|row| * 1/|q|), so a batched search returns bit-identical scores to a single one.
__global__ void __launch_bounds__(256) rescore_candidates_kernel(const key_t64 *cand, const uint32_t *cand_count, int cap,
                                                                 const uint32_t *gate_ordered, key_t64 *surv,
                                                                 uint32_t *surv_count, uint32_t *overflow, const float *mat,
                                                                 int dim, const float *inv_norm, const float *queries,
                                                                 const float *q_inv, uint32_t row_base, float min_sim,
                                                                 const uint16_t *hi, const uint16_t *lo, uint32_t blocks_per_query,
                                                                 int gate_bounds_exact)
{
    const uint32_t q = blockIdx.x / blocks_per_query, part = blockIdx.x % blocks_per_query;
    const uint32_t cnt = min(cand_count[q], (uint32_t)cap);
    const uint32_t per = (cnt + blocks_per_query - 1) / blocks_per_query;
    const uint32_t e_lo = part * per, e_hi = min(cnt, e_lo + per);      // this block's share of the list
    if (e_lo >= e_hi) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D4 = dim >> 2;
    const bool vec8 = (dim % 8) == 0 && (reinterpret_cast<uintptr_t>(mat) & 31) == 0;
    const float4 *qv = reinterpret_cast<const float4 *>(queries + (size_t)q * dim);
    float qi;
    if (q_inv) {
        qi = q_inv[q];
    } else {            // no query-norm kernel ran: one warp computes it with that kernel's arithmetic
        __shared__ float qi_s;
        if (warp == 0) {
            const float v = warp_inv_norm(queries + (size_t)q * dim, dim, lane);
            if (lane == 0) qi_s = v;
        }
        __syncthreads();
        qi = qi_s;
    }
    if (!(qi > 0.f)) return;            // a zero query matches nothing (deviation 1): no survivors
    const uint32_t gate = gate_ordered[q];
    // int8 sidecar: the gate is also a lower bound of the k-th best EXACT score, so exact scores below it are out
    const float exact_floor = gate_bounds_exact ? fmaxf(min_sim, ordered_to_score(gate)) : min_sim;
    const key_t64 *lst = cand + (size_t)q * cap;
    for (uint32_t i0 = e_lo + (uint32_t)warp * 32u; i0 < e_hi; i0 += 8u * 32u) {
        const uint32_t i = i0 + (uint32_t)lane;
        const key_t64 e = i < e_hi ? lst[i] : 0ull;
        uint32_t m = __ballot_sync(0xFFFFFFFFu, i < e_hi && (uint32_t)(e >> 32) >= gate);
        while (m) {
            const int src = __ffs(m) - 1;
            m &= m - 1;
            const uint32_t row = (uint32_t)(__shfl_sync(0xFFFFFFFFu, e, src) & 0xFFFFFFFFull);
            const float4 *rp = reinterpret_cast<const float4 *>(mat) + (size_t)row * D4;
            float acc = 0.f;
            if (hi) {            // split-plane store: rebuild the fp32 bits from both planes, same order as below
                const uint4 *h = reinterpret_cast<const uint4 *>(hi + (size_t)row * dim);
                const uint4 *l = reinterpret_cast<const uint4 *>(lo + (size_t)row * dim);
                for (int c = lane; c < (D4 >> 1); c += 32) {
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
            } else if (vec8) {   // the scan's 256-bit path: a lane owns 8 consecutive floats
                // four slots of loads in flight per lane (a warp works on one row at a time, so the row's latency is
                // paid per candidate); the FMAs keep the scan's order
                for (int c0 = lane; c0 < (D4 >> 1); c0 += 32 * 4) {
                    float4 v[4][2];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const int c = c0 + 32 * u;
                        if (c < (D4 >> 1)) { v[u][0] = ldg_stream(rp + 2 * c); v[u][1] = ldg_stream(rp + 2 * c + 1); }
                    }
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const int c = c0 + 32 * u;
                        if (c < (D4 >> 1)) {
#pragma unroll
                            for (int h = 0; h < 2; h++) {
                                const float4 x = v[u][h], w = qv[2 * c + h];
                                acc = fmaf(x.x, w.x, acc); acc = fmaf(x.y, w.y, acc); acc = fmaf(x.z, w.z, acc); acc = fmaf(x.w, w.w, acc);
                            }
                        }
                    }
                }
            } else {
                for (int c = lane; c < D4; c += 32) {
                    const float4 v = ldg_stream(rp + c), w = qv[c];
                    acc = fmaf(v.x, w.x, acc); acc = fmaf(v.y, w.y, acc); acc = fmaf(v.z, w.z, acc); acc = fmaf(v.w, w.w, acc);
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
            float sim = acc * inv_norm[row] * qi;
            sim = fminf(1.f, fmaxf(-1.f, sim));
            if (lane == 0 && sim >= exact_floor) {
                const uint32_t pos = atomicAdd(surv_count + q, 1u);
                if (pos < (uint32_t)kSurvivorCap) surv[(size_t)q * kSurvivorCap + pos] = make_key(sim, row_base + row);
                else *overflow = 1u;
            }
        }
    }
}

// One CTA per query: top-k of its re-scored survivors, then the chunk-id gather.
__global__ void __launch_bounds__(1024) select_candidates_kernel(const key_t64 *surv, const uint32_t *surv_count, int k, int P,
                                                                const uint8_t *chunk_ids, uint32_t row_base,
                                                                key_t64 *out_keys, uint8_t *out_ids)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    key_t64 *buf = reinterpret_cast<key_t64 *>(smem_raw);       // P keys: [0,k) kept, [k,P) incoming
    const uint32_t q = blockIdx.x;
    const int tid = threadIdx.x;
    const int n = (int)min(surv_count[q], (uint32_t)kSurvivorCap);
    const key_t64 *lst = surv + (size_t)q * kSurvivorCap;
    constexpr int kPerThread = kSurvivorCap / 1024;              // 16 survivors per thread at most in the small-k form
    if (k <= 20 && blockDim.x == 1024) {
        // Small k (the reference's API caps top_k at 20; 10 in the BASELINE configs) and a CTA to itself: every thread
        // keeps its survivors in registers and the CTA takes the maximum k times -- k rounds of a block-wide max
        // instead of up to eight 2048-wide sorts (measured on the int8 path: 78 us -> ~10).  Keys are unique.
        key_t64 mine[kPerThread];
#pragma unroll
        for (int j = 0; j < kPerThread; j++) { const int i = tid + j * 1024; mine[j] = i < n ? lst[i] : 0ull; }
        key_t64 *warp_best = buf;                                // 32 keys
        key_t64 *picked = buf + 32;                              // k keys
        const int lane = tid & 31, warp = tid >> 5;
        for (int r = 0; r < k; r++) {
            key_t64 best = 0ull;
#pragma unroll
            for (int j = 0; j < kPerThread; j++) best = mine[j] > best ? mine[j] : best;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { const key_t64 t = __shfl_xor_sync(0xFFFFFFFFu, best, o); best = t > best ? t : best; }
            if (lane == 0) warp_best[warp] = best;
            __syncthreads();
            key_t64 top = warp_best[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { const key_t64 t = __shfl_xor_sync(0xFFFFFFFFu, top, o); top = t > top ? t : top; }
            if (tid == 0) picked[r] = top;
            if (top) {
#pragma unroll
                for (int j = 0; j < kPerThread; j++) if (mine[j] == top) mine[j] = 0ull;
            }
            __syncthreads();
        }
        for (int i = tid; i < k; i += blockDim.x) buf[64 + i] = picked[i];
        __syncthreads();
        for (int i = tid; i < k; i += blockDim.x) buf[i] = buf[64 + i];
        __syncthreads();
    } else {
        for (int i = tid; i < k; i += blockDim.x) buf[i] = 0ull;
        const int room = P - k;
        int done = 0;
        do {
            __syncthreads();
            for (int i = tid; i < room; i += blockDim.x) buf[k + i] = (done + i < n) ? lst[done + i] : 0ull;
            cta_bitonic_desc(buf, P);
            done += room;
        } while (done < n);
    }
    for (int i = tid; i < k; i += blockDim.x) {
        const key_t64 key = buf[i];
        out_keys[(size_t)q * k + i] = key;
        uint4 id = make_uint4(0, 0, 0, 0);
        if (key) id = reinterpret_cast<const uint4 *>(chunk_ids)[key_row(key) - row_base];
        reinterpret_cast<uint4 *>(out_ids)[(size_t)q * k + i] = id;
    }
}

}  // namespace

cudaError_t launch_rescore_select(const key_t64 *cand, const uint32_t *cand_count, int cap, const uint32_t *gate_ordered,
                                  key_t64 *surv, uint32_t *surv_count, uint32_t *overflow, int batch,
                                  const float *mat, int dim, const float *inv_norm, const float *queries,
                                  const float *q_inv, const uint8_t *chunk_ids, uint32_t row_base, float min_sim,
                                  int k, key_t64 *out_keys, uint8_t *out_ids, cudaStream_t stream,
                                  const uint16_t *hi, const uint16_t *lo, bool gate_bounds_exact)
{
    cudaError_t e = cudaMemsetAsync(surv_count, 0, (size_t)batch * 4, stream);
    if (e != cudaSuccess) return e;
    // a few queries with long lists (the int8 sidecar) or thousands with short ones (the tensor path): about two
    // blocks per SM either way, never fewer than four per query
    uint32_t bpq = (uint32_t)((2 * 148 + batch - 1) / batch);
    if (bpq < 4) bpq = 4;
    rescore_candidates_kernel<<<(unsigned)batch * bpq, 256, 0, stream>>>(
        cand, cand_count, cap, gate_ordered, surv, surv_count, overflow, mat, dim, inv_norm, queries, q_inv, row_base, min_sim,
        hi, lo, bpq, gate_bounds_exact ? 1 : 0);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const bool few = batch <= kMaxNQ;           // one CTA per query: give it more threads and wider rounds when CTAs are few
    int P = next_pow2(2 * k);
    if (P < (few ? 2048 : 1024)) P = few ? 2048 : 1024;
    select_candidates_kernel<<<batch, few ? 1024 : 256, (size_t)P * sizeof(key_t64), stream>>>(surv, surv_count, k, P, chunk_ids,
                                                                                            row_base, out_keys, out_ids);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Narrow document filters (the reference's own case: document_id = ANY of one to sixteen ids, postgres.go:237-239).
// The requested rows sit in a few runs of ~200, which the grid-strided row loop would leave to a handful of warps;
// so the filter runs first, alone -- 4 bytes per row -- and compacts the requested rows into a list that the
// INDIRECT scan then spreads over every warp of the chip.
// ---------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) filter_compact_kernel(const uint32_t *doc_slot, const uint32_t *bitmap, uint32_t n_rows,
                                                             uint32_t *rows, uint32_t *n_out)
{
    const int lane = threadIdx.x & 31;
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t rounds = (n_rows + stride - 1) / stride;
    for (uint32_t it = 0, r = blockIdx.x * blockDim.x + threadIdx.x; it < rounds; it++, r += stride) {
        bool want = false;
        if (r < n_rows) {
            const uint32_t slot = doc_slot[r];
            want = (bitmap[slot >> 5] >> (slot & 31)) & 1u;
        }
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, want);
        if (m) {
            uint32_t pos = 0;
            if (lane == 0) pos = atomicAdd(n_out, (uint32_t)__popc(m));
            pos = __shfl_sync(0xFFFFFFFFu, pos, 0);
            if (want) rows[pos + __popc(m & ((1u << lane) - 1u))] = r;
        }
    }
}
}  // namespace

cudaError_t launch_filter_compact(const uint32_t *doc_slot, const uint32_t *bitmap, uint32_t n_rows, uint32_t *rows,
                                  uint32_t *n_out, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(n_out, 0, 4, stream);
    if (e != cudaSuccess || n_rows == 0) return e;
    uint32_t blocks = (n_rows + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    filter_compact_kernel<<<blocks, 256, 0, stream>>>(doc_slot, bitmap, n_rows, rows, n_out);
    return cudaGetLastError();
}

cudaError_t launch_merge(const uint8_t *gathered, int G, int batch, int k, key_t64 *out_keys,
                         uint8_t *out_ids, cudaStream_t stream, size_t block_bytes, uint32_t *out_flags)
{
    int P = next_pow2(G * k);
    if (P < 2) P = 2;
    if (block_bytes == 0) block_bytes = (size_t)batch * k * 24;
    merge_kernel<<<batch, kMergeThreads, (size_t)P * sizeof(key_t64), stream>>>(gathered, G, batch, k, P, out_keys, out_ids,
                                                                                block_bytes, out_flags);
    return cudaGetLastError();
}

}  // namespace da
