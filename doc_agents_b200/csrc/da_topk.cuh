// Device helpers shared by the scan kernels (da_scan.cu, da_split.cu): streaming loads, the per-warp bounded
// list with cached minimum, the CTA-wide bitonic sort.  Internal to the library.
#pragma once

#include "da_internal.h"
#ifdef DA_SCAN_TIMING
#include <cstdio>
#endif

namespace da {
namespace {

__device__ __forceinline__ float4 ldg_stream(const float4 *p)
{
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// 256-bit streaming load (sm_100: LDG.256): one warp instruction covers 1 KB of a row.
__device__ __forceinline__ void ldg_stream8(const float4 *p, float4 &a, float4 &b)
{
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}

__device__ __forceinline__ double warp_sum_d(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}

// 1/|v| of one vector, computed by one warp: double accumulation, fixed order (global or shared memory).
__device__ __forceinline__ float warp_inv_norm(const float *v, int dim, int lane)
{
    const float4 *p = reinterpret_cast<const float4 *>(v);
    double ss = 0.0;
    for (int c = lane; c < (dim >> 2); c += 32) {
        float4 x = p[c];
        ss += (double)x.x * x.x + (double)x.y * x.y + (double)x.z * x.z + (double)x.w * x.w;
    }
    ss = warp_sum_d(ss);
    return ss > 0.0 ? (float)(1.0 / sqrt(ss)) : 0.f;
}

__device__ __forceinline__ key_t64 warp_min_key(key_t64 v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        key_t64 t = __shfl_xor_sync(0xFFFFFFFFu, v, o);
        v = t < v ? t : v;
    }
    return v;
}

// Unsorted bounded list with cached minimum.  All lanes call with identical arguments.
// The rare part (append, or replace the minimum and find the new one) is one out-of-line function: the scan
// kernels expand the call once per (row, query) of their unrolled tails, and inlining the slow path there made
// the multi-query variants instruction-cache bound (ncu: `no_instruction` the top stall at 8 queries per pass).
struct ListState { int cnt; key_t64 wmin; };

__device__ __noinline__ ListState warp_list_insert_slow(key_t64 *list, int K, int cnt, key_t64 wmin, key_t64 key, int lane)
{
    if (cnt < K) {
        if (lane == 0) list[cnt] = key;
        cnt++;
        __syncwarp();
        if (cnt == K) {
            key_t64 m = ~0ull;
            for (int i = lane; i < K; i += 32) { key_t64 t = list[i]; m = t < m ? t : m; }
            wmin = warp_min_key(m);
        }
        return ListState{cnt, wmin};
    }
    // keys are unique (they embed the row), so exactly one slot holds wmin
    for (int i = lane; i < K; i += 32)
        if (list[i] == wmin) list[i] = key;
    __syncwarp();
    key_t64 m = ~0ull;
    for (int i = lane; i < K; i += 32) { key_t64 t = list[i]; m = t < m ? t : m; }
    return ListState{cnt, warp_min_key(m)};
}

__device__ __forceinline__ void warp_list_insert(key_t64 *list, int K, int &cnt, key_t64 &wmin,
                                                 key_t64 key, int lane)
{
    if (cnt >= K && key <= wmin) return;
    const ListState st = warp_list_insert_slow(list, K, cnt, wmin, key, lane);
    cnt = st.cnt;
    wmin = st.wmin;
}

#ifdef DA_SCAN_TIMING
__shared__ long long g_ts[16];
#define DA_TS(i) do { if (threadIdx.x == 0) g_ts[i] = clock64(); } while (0)
#else
#define DA_TS(i) do { } while (0)
#endif

// In-place descending bitonic sort of P (power of two) keys in shared memory by the whole CTA.
__device__ void cta_bitonic_desc(key_t64 *buf, int P)
{
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (P >> 1); t += blockDim.x) {
                int lo = 2 * t - (t & (stride - 1));
                int hi = lo + stride;
                bool desc = ((lo & size) == 0);
                key_t64 a = buf[lo], b = buf[hi];
                if ((a < b) == desc) { buf[lo] = b; buf[hi] = a; }
            }
        }
    }
    __syncthreads();
}

// The K (<= 32) largest of the keys a warp holds in registers (PER per lane, 0 = empty slot; keys are unique), in
// descending order: lane r returns the r-th largest (0 when fewer exist).  K rounds of two warp-wide integer maxima
// (upper word, then lower word among the lanes that hold the winning upper word); the one lane that owned the winner
// retires it.  No shared memory, no barrier: a round is ~100 cycles where a stage of the CTA-wide bitonic sort costs a
// barrier, and a sort of 1 024 keys has 55 of them.
template <int PER>
__device__ __forceinline__ key_t64 warp_topk_rounds(key_t64 (&v)[PER], int K, int lane)
{
    key_t64 best = 0ull, mine = 0ull;
#pragma unroll
    for (int i = 0; i < PER; i++) best = v[i] > best ? v[i] : best;
    for (int r = 0; r < K; r++) {
        const uint32_t hi = (uint32_t)(best >> 32);
        const uint32_t mx = __reduce_max_sync(0xFFFFFFFFu, hi);
        if (mx == 0u) break;                                   // nothing left (uniform)
        const uint32_t ml = __reduce_max_sync(0xFFFFFFFFu, hi == mx ? (uint32_t)best : 0u);
        const key_t64 win = ((key_t64)mx << 32) | (key_t64)ml;
        if (lane == r) mine = win;
        if (best == win) {
            best = 0ull;
#pragma unroll
            for (int i = 0; i < PER; i++) {
                if (v[i] == win) v[i] = 0ull;
                best = v[i] > best ? v[i] : best;
            }
        }
    }
    return mine;
}

constexpr int kSmallK = 20;             // top-k sizes whose merges run as warp tournaments: the reference's API allows 1..20 (cmd/query/main.go:23)
constexpr int kMidK = 16;               // up to here the narrow register arrays below do (measured: the wide ones cost k = 5 / 10 searches 0.8 us)
// keys per lane:                          K <= 16 | K <= 20
constexpr int kTourneyCtaPer = 8,  kTourneyCtaPerWide = 16;     // CTA level: kScanWarps x K <= P <= 256 | 512
constexpr int kTourneyPerLane = 5, kTourneyPerLaneWide = 6;     // level 1 of the last CTA's merge: 16 warps x 32 lanes x 5 | 6 keys >= 148 CTAs x K
constexpr int kTourneyL2Per = 8,   kTourneyL2PerWide = 10;      // level 2: kScanWarps x K <= 256 | 320 keys in one warp
constexpr int kTourneyXgPer = 4,   kTourneyXgPerWide = 5;       // cross-GPU merge: kXchgMaxRanks x K <= 128 | 160 keys in one warp

// The K best of src[0 .. count) (shared memory), one warp; lane r returns the r-th.
template <int PER>
__device__ __forceinline__ key_t64 tourney_smem(const key_t64 *src, int count, int K, int lane)
{
    key_t64 v[PER];
#pragma unroll
    for (int i = 0; i < PER; i++) v[i] = lane + 32 * i < count ? src[lane + 32 * i] : 0ull;
    return warp_topk_rounds<PER>(v, K, lane);
}
// ... of the CTAs' keys [lo, hi) of query q in global scratch (flat index j = cta * K + i).
template <int PER>
__device__ __forceinline__ key_t64 tourney_cta_keys(const key_t64 *cta_out, uint32_t lo, uint32_t hi, int K, int nq, int q, int lane)
{
    key_t64 v[PER];
#pragma unroll
    for (int i = 0; i < PER; i++) {
        const uint32_t j = lo + (uint32_t)lane + 32u * i;
        v[i] = j < hi ? __ldcg(&cta_out[((size_t)(j / K) * nq + q) * K + j % K]) : 0ull;
    }
    return warp_topk_rounds<PER>(v, K, lane);
}
// ... of the ranks' keys of query q in this rank's exchange buffer.
template <int PER>
__device__ __forceinline__ key_t64 tourney_rank_keys(const unsigned char *base, size_t slot_bytes, size_t keys_off, int n_ranks, int K, int q, int lane)
{
    key_t64 v[PER];
#pragma unroll
    for (int i = 0; i < PER; i++) {
        const int j = lane + 32 * i;
        v[i] = j < n_ranks * K ? __ldcv(reinterpret_cast<const key_t64 *>(base + (size_t)(j / K) * slot_bytes + keys_off) + (size_t)q * K + j % K) : 0ull;
    }
    return warp_topk_rounds<PER>(v, K, lane);
}

// The pieces of ScanArgs that probe mode re-bases per query (blockIdx.y).  Kept as a handful of scalars in registers:
// a mutable copy of the whole argument struct would live in local memory and put a load in front of every use in
// the row loop (measured: 5 % of the scan's bandwidth).
struct ScanPerQuery {
    const float *queries;
    const float *q_inv_norm;
    const uint32_t *doc_bitmap;
    uint32_t bitmap_stride;
    key_t64 *cta_out;
    unsigned int *counter;
    key_t64 *out_keys;
    uint8_t *out_ids;
    const key_t64 *probe_keys;
};
__device__ __forceinline__ ScanPerQuery scan_per_query(const ScanArgs &a)
{
    return ScanPerQuery{a.queries, a.q_inv_norm, a.doc_bitmap, a.bitmap_stride, a.cta_out, a.counter, a.out_keys, a.out_ids, a.probe_keys};
}

// Tail of the scan kernels: the k best keys of this CTA per query go to global scratch; the last CTA to arrive
// merges gridDim.x * k keys per query, writes a.out_keys and (if a.out_ids) gathers the chunk ids.
// `lists` = a.nq x a.P keys in shared memory, filled by the warps (unused slots 0); blockDim.x = kScanThreads.
// `mbuf` / `mbuf_keys`: the kernel's whole dynamic shared memory (queries + lists; both are dead once the CTA's keys
// are in global scratch), which lets the merge sort up to 8x more keys per round than one list holds -- the merge
// is the floor of small searches (a 3 200-row filtered search spent 30 of its 47 us in 7 rounds of 256 keys).
__device__ void scan_cta_epilogue(const ScanArgs &a, const ScanPerQuery &pq, key_t64 *lists, unsigned int *is_last_s, key_t64 *mbuf, int mbuf_keys)
{
    const int tid = threadIdx.x, K = a.k, P = a.P;
    DA_TS(3);
    // ---- CTA epilogue: k best of this CTA per query -> global scratch
    // Small k (the reference's 5 and 10): warp q ranks query q's kScanWarps x K keys in registers.  Otherwise: CTA-wide sorts.
    const int lane = tid & 31, warp = tid >> 5;
    const bool narrow = K <= kMidK;
    const bool tourney = K <= kSmallK && P <= 32 * (narrow ? kTourneyCtaPer : kTourneyCtaPerWide) &&
                         (uint32_t)gridDim.x * (uint32_t)K <= (uint32_t)(kScanWarps * 32 * (narrow ? kTourneyPerLane : kTourneyPerLaneWide));
    if (tourney) {
        __syncthreads();                                 // every warp's list is complete
        if (warp < a.nq) {
            const key_t64 mine = narrow ? tourney_smem<kTourneyCtaPer>(lists + (size_t)warp * P, P, K, lane)
                                        : tourney_smem<kTourneyCtaPerWide>(lists + (size_t)warp * P, P, K, lane);
            if (lane < K) pq.cta_out[((size_t)blockIdx.x * a.nq + warp) * K + lane] = mine;
        }
    } else {
        for (int q = 0; q < a.nq; q++) {
            cta_bitonic_desc(lists + (size_t)q * P, P);
            for (int i = tid; i < K; i += kScanThreads)
                pq.cta_out[((size_t)blockIdx.x * a.nq + q) * K + i] = lists[(size_t)q * P + i];
        }
    }
    // Release / acquire through ONE thread (the pattern of a cooperative grid barrier): the CTA barrier orders every
    // thread's stores before thread 0's fence, fences are cumulative, so the ticket publishes the whole CTA's keys; on
    // the other side thread 0's fence after the ticket and the barrier order the last CTA's loads after it.  (A fence in
    // all 512 threads, as before, bought nothing and cost ~1 us per search.)
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        unsigned int t = atomicAdd(pq.counter, 1u);
        *is_last_s = (t == gridDim.x - 1);
        __threadfence();
    }
    __syncthreads();
    if (!*is_last_s) return;
    DA_TS(4);

    // ---- last CTA: merge gridDim.x * K keys per query, then write the records
    const uint32_t total = gridDim.x * (uint32_t)K;
    int MP = P;                                          // sort width of the merge: a power of two that fits mbuf
    while (MP * 2 <= mbuf_keys && (uint32_t)MP < total + (uint32_t)K) MP *= 2;
    key_t64 *buf = MP > P ? mbuf : lists;
    __syncthreads();
    if (tourney) {
        // A handful of CTAs (one or two documents of the reference: 200 chunks each): all their keys fit one warp's registers,
        // so warp q ranks query q's straight from the scratch -- one level instead of two.
        const bool one_level = total <= 32u * (uint32_t)(narrow ? kTourneyL2Per : kTourneyL2PerWide);
        if (!one_level) {
            // level 1: the 16 warps share each query's gridDim.x * K keys (<= 5 per lane), winners into the dead list area
            const uint32_t share = (total + kScanWarps - 1) / kScanWarps;
            const uint32_t lo = (uint32_t)warp * share, hi = lo + share < total ? lo + share : total;
            for (int q = 0; q < a.nq; q++) {
                const key_t64 mine = narrow ? tourney_cta_keys<kTourneyPerLane>(pq.cta_out, lo, hi, K, a.nq, q, lane)
                                            : tourney_cta_keys<kTourneyPerLaneWide>(pq.cta_out, lo, hi, K, a.nq, q, lane);
                if (lane < K) lists[(size_t)q * P + warp * K + lane] = mine;
            }
            __syncthreads();
        }
        // level 2: warp q ranks query q's 16 x K winners (or, one level, every CTA's keys) and writes the records
        if (warp < a.nq) {
            const int q = warp;
            const key_t64 key = one_level ? (narrow ? tourney_cta_keys<kTourneyL2Per>(pq.cta_out, 0u, total, K, a.nq, q, lane)
                                                    : tourney_cta_keys<kTourneyL2PerWide>(pq.cta_out, 0u, total, K, a.nq, q, lane))
                                          : (narrow ? tourney_smem<kTourneyL2Per>(lists + (size_t)q * P, kScanWarps * K, K, lane)
                                                    : tourney_smem<kTourneyL2PerWide>(lists + (size_t)q * P, kScanWarps * K, K, lane));
            if (lane < K) {
                pq.out_keys[(size_t)q * K + lane] = key;
                if (a.gate_out && lane == K - 1)
                    a.gate_out[q] = score_to_ordered(key ? fmaxf(a.gate_floor, key_score(key)) : a.gate_floor);
                if (pq.out_ids) {
                    uint4 id = make_uint4(0, 0, 0, 0);
                    if (key) id = reinterpret_cast<const uint4 *>(a.chunk_ids)[key_row(key) - a.row_base];
                    reinterpret_cast<uint4 *>(pq.out_ids)[(size_t)q * K + lane] = id;
                }
            }
        }
        __syncthreads();
    } else
    for (int q = 0; q < a.nq; q++) {
        for (int i = tid; i < K; i += kScanThreads) buf[i] = 0ull;
        uint32_t done = 0;
        while (done < total) {
            __syncthreads();
            const uint32_t room = (uint32_t)(MP - K);
            for (uint32_t i = tid; i < room; i += kScanThreads) {
                uint32_t j = done + i;
                key_t64 v = 0ull;
                if (j < total) {
                    uint32_t cta = j / K, idx = j % K;
                    v = __ldcg(&pq.cta_out[((size_t)cta * a.nq + q) * K + idx]);
                }
                buf[K + i] = v;
            }
            cta_bitonic_desc(buf, MP);
            done += room;
        }
        for (int i = tid; i < K; i += kScanThreads) {
            key_t64 key = buf[i];
            pq.out_keys[(size_t)q * K + i] = key;
            if (a.gate_out && i == K - 1)
                a.gate_out[q] = score_to_ordered(key ? fmaxf(a.gate_floor, key_score(key)) : a.gate_floor);
            if (pq.out_ids) {
                uint4 id = make_uint4(0, 0, 0, 0);
                if (key) id = reinterpret_cast<const uint4 *>(a.chunk_ids)[key_row(key) - a.row_base];
                reinterpret_cast<uint4 *>(pq.out_ids)[(size_t)q * K + i] = id;
            }
        }
        __syncthreads();
    }
    DA_TS(5);
    // ---- exchange with the other GPUs (see XchgArgs): pq.out_keys / pq.out_ids ARE this rank's slot in its own buffer
    if (a.xg.n_ranks > 1) {
        const XchgArgs &x = a.xg;
        const size_t data0 = 256 + (size_t)x.parity * x.n_ranks * x.slot_bytes;
        const uint32_t rec16 = (uint32_t)(((size_t)a.nq * K * 24 + 15) / 16);       // 16-byte words of one record block
        __syncthreads();
        const uint4 *mine = reinterpret_cast<const uint4 *>(x.peer[x.rank] + data0 + (size_t)x.rank * x.slot_bytes);
        for (int r = 0; r < x.n_ranks; r++) {
            if (r == x.rank) continue;
            uint4 *dst = reinterpret_cast<uint4 *>(x.peer[r] + data0 + (size_t)x.rank * x.slot_bytes);
            for (uint32_t i = tid; i < rec16; i += kScanThreads) dst[i] = mine[i];
        }
        __threadfence_system();
        __syncthreads();
        if (tid == 0) {
            for (int r = 0; r < x.n_ranks; r++)
                atomicAdd_system(reinterpret_cast<unsigned long long *>(x.peer[r] + (size_t)x.parity * 128), 1ull);
            volatile unsigned long long *own = reinterpret_cast<volatile unsigned long long *>(x.peer[x.rank] + (size_t)x.parity * 128);
            unsigned long long t0;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            while (*own < x.target) {
                __nanosleep(64);
                unsigned long long t1;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 4000000000ull) { *x.timeout_flag = 1u; break; }      // 4 s: a peer never launched; do not hang the GPU
            }
            __threadfence_system();
        }
        __syncthreads();
        // merge n_ranks x K keys per query (n_ranks * K <= P: the CTA's own list area is wide enough)
        const unsigned char *base = x.peer[x.rank] + data0;
        const size_t keys_off = (size_t)a.nq * K * 16;
        if (tourney) {
            if (warp < a.nq) {                               // n_ranks x K <= 160 keys per query: warp q ranks query q's
                const int q = warp;
                const key_t64 key = narrow ? tourney_rank_keys<kTourneyXgPer>(base, x.slot_bytes, keys_off, x.n_ranks, K, q, lane)
                                           : tourney_rank_keys<kTourneyXgPerWide>(base, x.slot_bytes, keys_off, x.n_ranks, K, q, lane);
                if (lane < K) {
                    uint4 id = make_uint4(0, 0, 0, 0);
                    if (key) {
                        for (int r = 0; r < x.n_ranks; r++) {
                            const key_t64 *kb = reinterpret_cast<const key_t64 *>(base + (size_t)r * x.slot_bytes + keys_off) + (size_t)q * K;
                            const uint4 *ib = reinterpret_cast<const uint4 *>(base + (size_t)r * x.slot_bytes) + (size_t)q * K;
                            for (int j = 0; j < K; j++)
                                if (__ldcv(kb + j) == key) id = __ldcv(ib + j);
                        }
                    }
                    x.final_keys[(size_t)q * K + lane] = key;
                    reinterpret_cast<uint4 *>(x.final_ids)[(size_t)q * K + lane] = id;
                }
            }
            __syncthreads();
        } else
        for (int q = 0; q < a.nq; q++) {
            key_t64 *mb = lists + (size_t)q * P;
            for (int i = tid; i < P; i += kScanThreads) {
                key_t64 v = 0ull;
                if (i < x.n_ranks * K) {
                    const int r = i / K, j = i % K;
                    v = __ldcv(reinterpret_cast<const key_t64 *>(base + (size_t)r * x.slot_bytes + keys_off) + (size_t)q * K + j);
                }
                mb[i] = v;
            }
            cta_bitonic_desc(mb, P);
            for (int i = tid; i < K; i += kScanThreads) {
                const key_t64 key = mb[i];
                uint4 id = make_uint4(0, 0, 0, 0);
                if (key) {            // keys are unique across ranks (global rows): find the owner's id
                    for (int r = 0; r < x.n_ranks; r++) {
                        const key_t64 *kb = reinterpret_cast<const key_t64 *>(base + (size_t)r * x.slot_bytes + keys_off) + (size_t)q * K;
                        const uint4 *ib = reinterpret_cast<const uint4 *>(base + (size_t)r * x.slot_bytes) + (size_t)q * K;
                        for (int j = 0; j < K; j++)
                            if (__ldcv(kb + j) == key) id = __ldcv(ib + j);
                    }
                }
                x.final_keys[(size_t)q * K + i] = key;
                reinterpret_cast<uint4 *>(x.final_ids)[(size_t)q * K + i] = id;
            }
            __syncthreads();
        }
    }
    // small-search extras (see ScanArgs): counters handed over and cleared, then the completion flag
    __syncthreads();                 // every thread's record stores precede thread 0's system fence below (cumulative)
    DA_TS(6);
#ifdef DA_SCAN_TIMING
    if (tid == 0 && blockIdx.y == 0)
        printf("scan cta %d/%d: prologue %lld  loop %lld  cta-sort+ticket %lld  merge+ids %lld  fence %lld  (cycles)\n", (int)blockIdx.x, (int)gridDim.x,
               g_ts[1] - g_ts[0], g_ts[3] - g_ts[1], g_ts[4] - g_ts[3], g_ts[5] - g_ts[4], g_ts[6] - g_ts[5]);
#endif
    if (tid == 0) {
        if (a.rows_read_out) {
            *a.rows_read_out = a.rows_read ? atomicExch(a.rows_read, 0ull) : 0ull;
        }
        if (a.n_rows_reset) *a.n_rows_reset = 0u;
        *pq.counter = 0u;
        if (a.done_flag) {
            __threadfence_system();
            *a.done_flag = a.done_value;
        }
    }
}

inline int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

}  // namespace
}  // namespace da
