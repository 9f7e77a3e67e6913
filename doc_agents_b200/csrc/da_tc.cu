// Tensor-core GEMM with a fused per-row reduction for sm_100a (tcgen05 + TMEM + TMA).
//
//   scores = A (M x K, fp32, K contiguous) . B^T (N x K, fp32, K contiguous)     as kind::tf32
//   out[m] = argmax_n scores[m][n]                       (the score matrix is never written)
//
// Used for the IVFFlat build (internal/store/postgres.go:95-99 asks Postgres for that index):
// A = embedding rows (or the k-means sample), B = centroids.  TF32 reads the fp32 words as they
// sit in HBM (TMA -> 128B-swizzled shared memory -> tcgen05.mma), so no conversion pass and no
// second copy of the corpus exist.  Assignment only steers recall; scores returned to the caller
// always come from the fp32 scan.
//
// CTA = 192 threads:  warp 0 TMA producer | warp 1 TMEM owner + MMA issuer | warps 2-5 epilogue.
// Tile 128 (M) x 256 (N) x 32 (K, = one 128-byte swizzle row), 4-stage smem ring (48 KB/stage),
// two 256-column accumulators in TMEM so the epilogue of tile t overlaps the MMAs of tile t+1.
// Epilogue thread r owns accumulator lane r = one row of A, streams the 256 columns through
// registers with tcgen05.ld and keeps a running (max, argmax): no cross-thread reduction.
// Grid = M tiles x N partitions; sibling CTAs (same M tile) run together so the A tile is read
// from HBM once and from L2 afterwards; partitions meet through one 64-bit atomicMax per row.
#include "da_internal.h"

#include <cuda.h>

#include <cstdlib>

namespace da {

namespace {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 32, TC_STAGES = 4, TC_UMMA_K = 8;
constexpr int TC_THREADS = 192;
constexpr uint32_t TC_A_BYTES = TC_BM * TC_BK * 4;            // 16 KB
constexpr uint32_t TC_B_BYTES = TC_BN * TC_BK * 4;            // 32 KB
constexpr uint32_t TC_STAGE_BYTES = TC_A_BYTES + TC_B_BYTES;  // 48 KB
constexpr uint32_t TC_SMEM_BYTES = TC_STAGES * TC_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t TC_TMEM_COLS = 512;

// instruction descriptor: D=f32, A=B=tf32, both K-major, N=256, M=128
constexpr uint32_t TC_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
// K-major, 128-byte swizzle: 8-row groups are 1024 B apart (SBO), LBO unused (=1), descriptor version 1
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr)
{
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t da_, uint64_t db_, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(da_), "l"(db_), "r"(idesc), "r"(accumulate) : "memory");
}
// kind::f16 with bf16 operands (split-plane stores: the upper 16-bit plane IS a truncated-bf16 matrix)
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da_, uint64_t db_, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(da_), "l"(db_), "r"(idesc), "r"(accumulate) : "memory");
}
// instruction descriptor: D=f32, A=B=bf16, both K-major, N=256, M=128 (K = 16 per instruction = 32 bytes, like tf32's 8)
constexpr uint32_t TC_IDESC_BF16 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);

__device__ __forceinline__ void umma_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// ---- CTA pairs (cta_group::2): two CTAs of one cluster on one TPC run each MMA together.  Every CTA stages its own
// 128 A rows and HALF of the B tile; the leader (cluster rank 0) issues the MMAs, whose completion is multicast to the
// barriers of both CTAs; both CTAs' TMA loads complete on the leader's barrier.
__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t saddr, uint32_t rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr)
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap *map, uint32_t bar_cluster, int c0, int c1)
{
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void umma_tf32_pair(uint32_t tmem_d, uint64_t da_, uint64_t db_, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(da_), "l"(db_), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t da_, uint64_t db_, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(da_), "l"(db_), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint32_t bar)      // arrives on `bar` of BOTH CTAs of the pair
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
// kind::i8, signed 8-bit operands, s32 accumulators (int8 sidecar stores): 32 elements of K per instruction
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t da_, uint64_t db_, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(da_), "l"(db_), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_i8_pair(uint32_t tmem_d, uint64_t da_, uint64_t db_, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(da_), "l"(db_), "r"(idesc), "r"(accumulate) : "memory");
}
// D = s32 (2 << 4), A = B = signed 8 bit (1 << 7, 1 << 10)
constexpr uint32_t TC_IDESC_I8 = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
constexpr uint32_t TC_IDESC_I8_PAIR = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)((2 * TC_BM) >> 4) << 24);
// pair instruction descriptors: M = 256 (128 rows of D in each CTA's TMEM), N = 256
constexpr uint32_t TC_IDESC_PAIR = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)((2 * TC_BM) >> 4) << 24);
constexpr uint32_t TC_IDESC_BF16_PAIR = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)((2 * TC_BM) >> 4) << 24);

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(TC_THREADS, 1)
tc_rowargmax_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                    uint32_t M, uint32_t N, int K, int n_parts, unsigned long long *__restrict__ out_packed)
{
    extern __shared__ unsigned char smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;          // SW128 tiles need 1024-byte alignment
    const uint32_t bars = base + TC_STAGES * TC_STAGE_BYTES;
    // barrier slots (8 B each): full[4] | empty[4] | tmem_full[2] | tmem_empty[2] | tmem base address
    const uint32_t bar_full = bars, bar_empty = bars + 8 * TC_STAGES;
    const uint32_t bar_tfull = bars + 16 * TC_STAGES, bar_tempty = bar_tfull + 16;
    const uint32_t tmem_slot = bar_tempty + 16;
    volatile uint32_t *tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t *>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t m_tile = blockIdx.x / n_parts, part = blockIdx.x % n_parts;
    const uint32_t m0 = m_tile * TC_BM;
    const int n_tiles_total = (int)((N + TC_BN - 1) / TC_BN);
    const int per_part = (n_tiles_total + n_parts - 1) / n_parts;
    const int t_begin = (int)part * per_part;
    const int t_end = min(n_tiles_total, t_begin + per_part);
    const int kb_count = (K + TC_BK - 1) / TC_BK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_STAGES; s++) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
        for (int s = 0; s < 2; s++) { mbar_init(bar_tfull + 8 * s, 1); mbar_init(bar_tempty + 8 * s, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TC_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int t = t_begin; t < t_end; t++) {
                for (int kb = 0; kb < kb_count; kb++) {
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    const uint32_t sa = base + stage * TC_STAGE_BYTES, sb = sa + TC_A_BYTES;
                    mbar_expect_tx(bar_full + 8 * stage, TC_STAGE_BYTES);
                    tma_load_2d(sa, &map_a, bar_full + 8 * stage, kb * TC_BK, (int)m0);
                    tma_load_2d(sb, &map_b, bar_full + 8 * stage, kb * TC_BK, t * TC_BN);
                    if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one elected lane) =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            int it = 0;
            for (int t = t_begin; t < t_end; t++, it++) {
                const int as = it & 1;
                const uint32_t aphase = (uint32_t)(it >> 1) & 1;
                mbar_wait(bar_tempty + 8 * as, aphase ^ 1);            // epilogue drained this accumulator
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_d = tmem_base + (uint32_t)as * TC_BN;
                for (int kb = 0; kb < kb_count; kb++) {
                    mbar_wait(bar_full + 8 * stage, phase);              // TMA bytes have landed
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = base + stage * TC_STAGE_BYTES, sb = sa + TC_A_BYTES;
                    const uint64_t da0 = smem_desc_sw128(sa), db0 = smem_desc_sw128(sb);
#pragma unroll
                    for (int k = 0; k < TC_BK / TC_UMMA_K; k++) {
                        // advancing K inside the 128-byte swizzle row: +32 bytes = +2 in the >>4 start address
                        umma_tf32(tmem_d, da0 + (uint64_t)(k * 2), db0 + (uint64_t)(k * 2), TC_IDESC, (uint32_t)((kb | k) != 0));
                    }
                    umma_commit(bar_empty + 8 * stage);                  // frees the smem slot when the MMAs retire
                    if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(bar_tfull + 8 * as);                         // accumulator complete -> epilogue
            }
        }
    } else {
        // ===== epilogue: warps 2..5, TMEM lane quadrant = warp % 4 =====
        const uint32_t quad = (uint32_t)warp & 3u;
        const uint32_t row = m0 + quad * 32 + (uint32_t)lane;
        float best = -3.0e38f;
        uint32_t best_n = 0;
        int it = 0;
        for (int t = t_begin; t < t_end; t++, it++) {
            const int as = it & 1;
            const uint32_t aphase = (uint32_t)(it >> 1) & 1;
            mbar_wait(bar_tfull + 8 * as, aphase);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + (uint32_t)as * TC_BN;
#pragma unroll 1
            for (int c = 0; c < TC_BN / 32; c++) {
                uint32_t r[32];
                tmem_ld32(taddr + (uint32_t)c * 32u, r);
                const uint32_t n0 = (uint32_t)t * TC_BN + (uint32_t)c * 32u;
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    const float v = __uint_as_float(r[j]);
                    if (n0 + j < N && v > best) { best = v; best_n = n0 + j; }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty + 8 * as);
        }
        if (row < M && t_begin < t_end) {
            const unsigned long long packed = ((unsigned long long)score_to_ordered(best) << 32) | (unsigned long long)(~best_n);
            atomicMax(out_packed + row, packed);
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TC_TMEM_COLS) : "memory");
    }
}

// ---------------------------------------------------------------------------
// Plain score tiles for the batched IVFFlat query -> centroid step: out[split][m][n] = sum over the split's slice
// of the dimension of A[m][k] * B[n][k], A = the queries of a batch, B = the centroids.  The problem is small
// (64 x 4096 x 3072 = 1.6 GFLOP, 50 MB of centroids) and short of tiles (16 for that shape), so the dimension is
// cut into `ks` slices, one CTA per (tile, slice): every SM streams a share of the centroid matrix, and the partial
// sums are added in a fixed order by the selection kernel (da_ivf.cu) -- deterministic, no atomics.
// Same pipeline as above (TMA -> 4-stage ring -> tcgen05.mma kind::tf32 -> TMEM), one accumulator per CTA.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(TC_THREADS, 1)
tc_scores_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 uint32_t M, uint32_t N, int kb_total, int kb_per_split, int n_tiles, float *__restrict__ out)
{
    extern __shared__ unsigned char smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = base + TC_STAGES * TC_STAGE_BYTES;
    const uint32_t bar_full = bars, bar_empty = bars + 8 * TC_STAGES, bar_tfull = bars + 16 * TC_STAGES;
    const uint32_t tmem_slot = bar_tfull + 8;
    volatile uint32_t *tmem_slot_ptr = reinterpret_cast<volatile uint32_t *>(smem_raw + (tmem_slot - smem_u32(smem_raw)));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t tile = blockIdx.x, split = blockIdx.y;
    const uint32_t m0 = (tile / (uint32_t)n_tiles) * TC_BM, n0 = (tile % (uint32_t)n_tiles) * TC_BN;
    const int kb_begin = (int)split * kb_per_split;
    const int kb_end = min(kb_total, kb_begin + kb_per_split);

    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_STAGES; s++) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
        mbar_init(bar_tfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)TC_BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int kb = kb_begin; kb < kb_end; kb++) {
                mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                const uint32_t sa = base + stage * TC_STAGE_BYTES, sb = sa + TC_A_BYTES;
                mbar_expect_tx(bar_full + 8 * stage, TC_STAGE_BYTES);
                tma_load_2d(sa, &map_a, bar_full + 8 * stage, kb * TC_BK, (int)m0);
                tma_load_2d(sb, &map_b, bar_full + 8 * stage, kb * TC_BK, (int)n0);
                if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int kb = kb_begin; kb < kb_end; kb++) {
                mbar_wait(bar_full + 8 * stage, phase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa = base + stage * TC_STAGE_BYTES, sb = sa + TC_A_BYTES;
                const uint64_t da0 = smem_desc_sw128(sa), db0 = smem_desc_sw128(sb);
#pragma unroll
                for (int k = 0; k < TC_BK / TC_UMMA_K; k++)
                    umma_tf32(tmem_base, da0 + (uint64_t)(k * 2), db0 + (uint64_t)(k * 2), TC_IDESC, (uint32_t)((kb > kb_begin) | (k != 0)));
                umma_commit(bar_empty + 8 * stage);
                if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
            }
            umma_commit(bar_tfull);
        }
    } else {
        // epilogue warps 2..5: thread = one row of A (one query), 256 scores in eight chunks of 32
        const uint32_t quad = (uint32_t)warp & 3u;
        const uint32_t m = m0 + quad * 32 + (uint32_t)lane;
        const bool any = kb_begin < kb_end;
        if (any) {
            mbar_wait(bar_tfull, 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        float *dst = out + ((size_t)split * M + m) * N;
        const bool vec = (N & 3u) == 0u;
#pragma unroll 1
        for (int c = 0; c < TC_BN / 32; c++) {
            uint32_t r[32];
            if (any) tmem_ld32(tmem_base + ((quad * 32u) << 16) + (uint32_t)c * 32u, r);
            else {
#pragma unroll
                for (int j = 0; j < 32; j++) r[j] = 0u;
            }
            const uint32_t nb = n0 + (uint32_t)c * 32u;
            if (m < M) {
                if (vec && nb + 32u <= N) {
#pragma unroll
                    for (int j = 0; j < 8; j++)
                        reinterpret_cast<float4 *>(dst + nb)[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                                                             __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
                } else {
#pragma unroll
                    for (int j = 0; j < 32; j++)
                        if (nb + (uint32_t)j < N) dst[nb + j] = __uint_as_float(r[j]);
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TC_BN) : "memory");
    }
}

__global__ void unpack_assign_kernel(const unsigned long long *packed, uint32_t n, uint32_t *assign)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        assign[i] = ~(uint32_t)(packed[i] & 0xFFFFFFFFull);
}

__global__ void gather_rows_kernel(const float *mat, int D, const uint32_t *rows, uint32_t n, float *out)
{
    const int lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = w; i < n; i += nw) {
        const float4 *src = reinterpret_cast<const float4 *>(mat + (size_t)rows[i] * D);
        float4 *dst = reinterpret_cast<float4 *>(out + (size_t)i * D);
        for (int c = lane; c < (D >> 2); c += 32) dst[c] = src[c];
    }
}

// ---------------------------------------------------------------------------
// Batched queries: scores = Q (batch x K) . X^T (N rows x K) on the tensor cores, fused with a
// per-query candidate filter, so the batch x N score matrix never leaves the SM.
//
// Roles swap relative to the assignment kernel: A = queries (an M tile = 128 queries), B = corpus
// rows (N tiles of 256 rows).  A CTA owns a contiguous range of row tiles and, for each of them,
// sweeps ALL query tiles back to back -- the row tile comes from HBM once and from L2 for the
// other query tiles, so DRAM traffic is the corpus once per pass whatever the batch.  Epilogue
// thread r owns query r of the current query tile and walks the 256 row scores of the accumulator.
//
// TF32 scores are only a candidate generator.  With eps >= |cos_tf32 - cos_fp32|, a row can be in
// the exact top-k only if  cos_tf32 >= T_k - 2*eps  and  cos_tf32 >= floor - eps, where T_k is any
// lower bound of the k-th best TF32 score.  T_k is GLOBAL here: every appended candidate is counted
// in a per-query two-level histogram of scores (64 x 64 bins); scanning it from the top until k
// entries are covered gives a bound that all CTAs share, lock-free (counters only grow).  Survivors
// go to one list per query in global memory; a second kernel drops those below the final gate and
// re-scores the rest in fp32 with the scan kernel's arithmetic; a third selects the top-k.
// ---------------------------------------------------------------------------
constexpr int TCB_MAX_MTILES = DA_TC_MAX_BATCH / TC_BM;   // 32
constexpr uint32_t TCB_SMEM_BYTES = TC_STAGES * TC_STAGE_BYTES + 1024 + 256 + 4 * TC_BN * 4 + TCB_MAX_MTILES * TC_BM * 4;

struct BatchArgs {
    uint32_t n_rows;            // rows resident on this shard
    int K, batch, k;
    int cap;                    // entries per query list
    int group;                  // CTAs that share one range of row tiles and split the query tiles among them
    float min_sim;              // similarity floor (exact scores)
    float eps;                  // bound of |approximate - exact| over rows (and queries, unless q_err adds the query's own part)
    const float *q_err;         // int8 operands: per query, its quantisation error bound (the gate slack becomes per query)
    const float *inv_norm;      // per row
    const uint32_t *doc_slot;   // per row (with doc_bitmap)
    const uint32_t *doc_bitmap; // nullptr = every document
    uint32_t bitmap_stride;     // words between consecutive queries' bitmaps; 0 = one shared bitmap
    const float *q_inv;         // per query
    unsigned long long *cand;   // [batch][cap]   ordered(score) << 32 | local row
    uint32_t *cand_count;       // [batch]
    uint32_t *hist;             // [batch][64 coarse + 4096 fine]
    uint32_t *overflow;         // set to 1 if a query list ran out of room
    uint32_t *group_sync;       // [n_groups] row tiles started by the members of a group (zeroed per launch)
};
constexpr int kHistWords = 64 + 64 * 64;

__device__ __forceinline__ void hist_bins(float v, int &c, int &f)
{
    const float x = (v + 1.0f) * 32.0f;                 // [-1, 1] -> [0, 64]
    c = min(63, max(0, (int)x));
    f = min(63, max(0, (int)((x - (float)c) * 64.0f)));
}
__device__ __forceinline__ float hist_edge(int c, int f) { return ((float)c + (float)f * (1.0f / 64.0f)) * (1.0f / 32.0f) - 1.0f; }

// A lower bound of the k-th best appended score of one query (-inf while fewer than k were appended).
// Two rounds of 16 independent 128-bit loads (coarse row, then one fine row): latency of two L2 trips.
__device__ float hist_kth_lower_bound(const uint32_t *h, int k)
{
    uint4 w[16];
#pragma unroll
    for (int i = 0; i < 16; i++) w[i] = __ldcg(reinterpret_cast<const uint4 *>(h) + i);
    uint32_t cum = 0;
    int cbin = -1;
#pragma unroll
    for (int i = 15; i >= 0; i--) {
        const uint32_t v[4] = {w[i].x, w[i].y, w[i].z, w[i].w};
#pragma unroll
        for (int j = 3; j >= 0; j--) {
            if (cbin < 0) {
                if (cum + v[j] >= (uint32_t)k) cbin = i * 4 + j; else cum += v[j];
            }
        }
    }
    if (cbin < 0) return -3.0e38f;
    const uint4 *fine = reinterpret_cast<const uint4 *>(h + 64 + cbin * 64);
#pragma unroll
    for (int i = 0; i < 16; i++) w[i] = __ldcg(fine + i);
    int fbin = -1;
#pragma unroll
    for (int i = 15; i >= 0; i--) {
        const uint32_t v[4] = {w[i].x, w[i].y, w[i].z, w[i].w};
#pragma unroll
        for (int j = 3; j >= 0; j--) {
            if (fbin < 0) {
                if (cum + v[j] >= (uint32_t)k) fbin = i * 4 + j; else cum += v[j];
            }
        }
    }
    return hist_edge(cbin, fbin < 0 ? 0 : fbin);   // fine counters may lag the coarse one: its edge is still valid
}

constexpr int OP_TF32 = 0, OP_BF16 = 1, OP_I8 = 2;

// Slack of query q: |approximate - exact| <= eps for every row.  int8 operands: the rows' bound a.eps (the largest
// quantisation error of a resident row) combined with the query's own (da_q8.cu: e_x + e_q + e_x e_q, + rounding).
__device__ __forceinline__ float batch_eps(const BatchArgs &a, uint32_t q)
{
    if (!a.q_err || q >= (uint32_t)a.batch) return a.eps;
    const float qe = a.q_err[q];
    return a.eps + qe + a.eps * qe + 2.0e-5f;
}
// accumulator word -> float (s32 accumulators with int8 operands)
template <bool I8> __device__ __forceinline__ float acc_value(uint32_t w)
{
    return I8 ? (float)(int)w : __uint_as_float(w);
}

// Warp roles of the batch kernel: warps 0..3 epilogue (TMEM lane quadrant = warp % 4), warp 4 TMA producer, warp 5
// TMEM owner + MMA issuer.
constexpr int kProdWarp = 4, kMmaWarp = 5, kEpiFirst = 0;

// OP: operand type of the contraction -- OP_TF32, OP_BF16 (kind::f16) or OP_I8 (kind::i8, s32 accumulators: the int8
// sidecar; the per-row factor is then scale/|row|, the per-query one scale/|q|) --, fixed at compile time (the issue loop has no
// run-time choice).  NCTA = 2: CTA pairs (cluster of two), each pair takes TWO query tiles against one row tile per
// accumulator; a CTA stages its own query tile and half of the row tile, which halves the shared-memory traffic per
// MMA (the limit of the single-CTA tf32 form) and the L2 traffic per query tile.
template <int OP, int NCTA>
__global__ void __launch_bounds__(TC_THREADS, 1)
tc_batch_candidates_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_x,
                           const BatchArgs a)
{
    constexpr uint32_t B_BYTES = TC_B_BYTES / NCTA;
    constexpr uint32_t STAGE_BYTES = TC_A_BYTES + B_BYTES;
    constexpr int STAGES = TC_STAGES * (int)TC_STAGE_BYTES / (int)STAGE_BYTES;      // 4 x 48 KB or 6 x 32 KB
    extern __shared__ unsigned char smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = base + STAGES * STAGE_BYTES;
    const uint32_t bar_full = bars, bar_empty = bars + 8 * STAGES;
    const uint32_t bar_tfull = bars + 16 * STAGES, bar_tempty = bar_tfull + 16;
    const uint32_t tmem_slot = bar_tempty + 16;
    volatile uint32_t *tmem_slot_ptr = reinterpret_cast<volatile uint32_t *>(smem_raw + (tmem_slot - smem_u32(smem_raw)));
    float *s_inv = reinterpret_cast<float *>(smem_raw + (bars + 256 - smem_u32(smem_raw)));   // [2][TC_BN]
    uint32_t *s_slot = reinterpret_cast<uint32_t *>(s_inv + 2 * TC_BN);                        // [2][TC_BN], per-query masks only
    float *gate_s = s_inv + 4 * TC_BN;                                                        // [m_tiles][128]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = NCTA == 2 ? (int)cluster_ctarank() : 0;        // 0 = leader of the pair
    const int m_tiles = (a.batch + TC_BM - 1) / TC_BM;
    const int m_units = (m_tiles + NCTA - 1) / NCTA;                // query tiles (pairs of them) dealt to a unit
    const int n_tiles_total = (int)((a.n_rows + TC_BN - 1) / TC_BN);
    // A unit is one CTA (or CTA pair).  `group` neighbouring units share one range of row tiles (the tile reaches
    // them through L2) and deal the query tiles out among themselves, so that the row tiles in flight across the
    // chip stay L2-resident.
    const int unit = (int)blockIdx.x / NCTA;
    const int n_groups = ((int)gridDim.x / NCTA) / a.group;
    const int grp = unit / a.group, member = unit % a.group;
    const int per_grp = (n_tiles_total + n_groups - 1) / n_groups;
    const int t_begin = grp * per_grp;
    const int t_end = min(n_tiles_total, t_begin + per_grp);
    constexpr bool BF16 = OP == OP_BF16, I8 = OP == OP_I8;
    const int bk = I8 ? 4 * TC_BK : BF16 ? 2 * TC_BK : TC_BK;     // elements per 128-byte stage row
    const int kb_count = (a.K + bk - 1) / bk;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
        for (int s = 0; s < 2; s++) { mbar_init(bar_tfull + 8 * s, 1); mbar_init(bar_tempty + 8 * s, 4 * NCTA); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kMmaWarp) {
        if (NCTA == 2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TC_TMEM_COLS) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TC_TMEM_COLS) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    if (NCTA == 2) cluster_sync_all(); else __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot_ptr;
    // barriers that live on the leader only: "stage full" (both CTAs' loads land on it) and "accumulator drained"
    const uint32_t lead_full = NCTA == 2 ? mapa_rank(bar_full, 0) : bar_full;
    const uint32_t lead_tempty = NCTA == 2 ? mapa_rank(bar_tempty, 0) : bar_tempty;

    if (warp == kProdWarp) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            bool lockstep = true;
            for (int t = t_begin; t < t_end; t++) {
                if (a.group > 1 && rank == 0) {
                    // Lockstep within the group: start row tile i only when every member has started tile i-1, so at
                    // most two row tiles per group are live in L2 (members that drift apart would each pull their own
                    // copy from HBM and push the others' out).  It is a locality measure, not a correctness one: a
                    // member that waits a millisecond (its neighbours are not resident yet, say, because another
                    // kernel holds their SMs) stops waiting for good, so the kernel needs no co-residency guarantee.
                    // The second CTA of a pair follows its leader through the stage barriers.
                    const uint32_t need = (uint32_t)a.group * (uint32_t)(t - t_begin);
                    volatile uint32_t *ctr = a.group_sync + grp;
                    for (int spins = 0; lockstep && *ctr < need; spins++) {
                        __nanosleep(200);
                        if (spins > 5000) lockstep = false;
                    }
                    atomicAdd(a.group_sync + grp, 1u);
                }
                for (int mu = member; mu < m_units; mu += a.group) {
                    const int mt = mu * NCTA + rank;
                    for (int kb = 0; kb < kb_count; kb++) {
                        mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                        const uint32_t sa = base + stage * STAGE_BYTES, sb = sa + TC_A_BYTES;
                        if (NCTA == 2) {
                            if (rank == 0) mbar_expect_tx(bar_full + 8 * stage, 2 * STAGE_BYTES);
                            tma_load_2d_pair(sa, &map_q, lead_full + 8 * stage, kb * bk, mt * TC_BM);
                            tma_load_2d_pair(sb, &map_x, lead_full + 8 * stage, kb * bk, t * TC_BN + rank * (TC_BN / 2));
                        } else {
                            mbar_expect_tx(bar_full + 8 * stage, STAGE_BYTES);
                            tma_load_2d(sa, &map_q, bar_full + 8 * stage, kb * bk, mt * TC_BM);
                            tma_load_2d(sb, &map_x, bar_full + 8 * stage, kb * bk, t * TC_BN);
                        }
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == kMmaWarp) {
        if (lane == 0 && rank == 0) {
            int stage = 0; uint32_t phase = 0;
            int it = 0;
            for (int t = t_begin; t < t_end; t++) {
                for (int mu = member; mu < m_units; mu += a.group, it++) {
                    const int as = it & 1;
                    const uint32_t aphase = (uint32_t)(it >> 1) & 1;
                    mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t tmem_d = tmem_base + (uint32_t)as * TC_BN;
                    for (int kb = 0; kb < kb_count; kb++) {
                        mbar_wait(bar_full + 8 * stage, phase);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint32_t sa = base + stage * STAGE_BYTES, sb = sa + TC_A_BYTES;
                        const uint64_t da0 = smem_desc_sw128(sa), db0 = smem_desc_sw128(sb);
#pragma unroll
                        for (int k = 0; k < TC_BK / TC_UMMA_K; k++) {     // 4 MMAs of 32 bytes of K each, either kind
                            const uint64_t da_ = da0 + (uint64_t)(k * 2), db_ = db0 + (uint64_t)(k * 2);
                            const uint32_t acc = (uint32_t)((kb | k) != 0);
                            if (NCTA == 2) {
                                if (I8) umma_i8_pair(tmem_d, da_, db_, TC_IDESC_I8_PAIR, acc);
                                else if (BF16) umma_bf16_pair(tmem_d, da_, db_, TC_IDESC_BF16_PAIR, acc);
                                else umma_tf32_pair(tmem_d, da_, db_, TC_IDESC_PAIR, acc);
                            } else {
                                if (I8) umma_i8(tmem_d, da_, db_, TC_IDESC_I8, acc);
                                else if (BF16) umma_bf16(tmem_d, da_, db_, TC_IDESC_BF16, acc);
                                else umma_tf32(tmem_d, da_, db_, TC_IDESC, acc);
                            }
                        }
                        if (NCTA == 2) umma_commit_pair(bar_empty + 8 * stage); else umma_commit(bar_empty + 8 * stage);
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                    }
                    if (NCTA == 2) umma_commit_pair(bar_tfull + 8 * as); else umma_commit(bar_tfull + 8 * as);
                }
            }
        }
    } else {
        // ===== epilogue warps 0..3: thread = query `qlocal` of every query tile =====
        const uint32_t quad = (uint32_t)warp & 3u;
        const int etid = (int)threadIdx.x - kEpiFirst * 32;                      // 0..127 among the epilogue threads
        const uint32_t qlocal = quad * 32u + (uint32_t)lane;
        for (int mu = member; mu < m_units; mu += a.group) {
            const uint32_t q = (uint32_t)(mu * NCTA + rank) * TC_BM + qlocal;
            gate_s[(mu * NCTA + rank) * TC_BM + qlocal] = a.min_sim - batch_eps(a, q);
        }
        int it = 0, visited = 0, next_refresh = 1;
        for (int t = t_begin; t < t_end; t++, visited++) {
            const uint32_t n0 = (uint32_t)t * TC_BN;
            // per-row factor for this row tile: 1/|row|, or 0 when the row is absent, zero or filtered out
            float *inv_t = s_inv + (t & 1) * TC_BN;
            uint32_t *slot_t = s_slot + (t & 1) * TC_BN;
            const bool per_query = a.doc_bitmap && a.bitmap_stride;     // each query has its own document set
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int j = etid + h * 128;
                const uint32_t row = n0 + (uint32_t)j;
                float f = 0.f;
                uint32_t slot = 0;
                if (row < a.n_rows) {
                    f = a.inv_norm[row];
                    if (a.doc_bitmap) {
                        slot = a.doc_slot[row];
                        if (!per_query && !((a.doc_bitmap[slot >> 5] >> (slot & 31)) & 1u)) f = 0.f;
                    }
                }
                inv_t[j] = f > 0.f ? f : __int_as_float(0x7fc00000);   // NaN: never passes a gate
                slot_t[j] = slot;
            }
            // tighten this thread's gates from the shared histograms: often at first, then every 16 row tiles
            if (visited == next_refresh) {
                next_refresh = visited < 4 ? visited + 1 : (visited < 16 ? visited + 2 : visited + 16);
                for (int mu = member; mu < m_units; mu += a.group) {
                    const int mt = mu * NCTA + rank;
                    const uint32_t q = (uint32_t)mt * TC_BM + qlocal;
                    if (q < (uint32_t)a.batch) {
                        const float lb = hist_kth_lower_bound(a.hist + (size_t)q * kHistWords, a.k);
                        const float eps = batch_eps(a, q);
                        gate_s[mt * TC_BM + qlocal] = fmaxf(a.min_sim - eps, lb - 2.f * eps);
                    }
                }
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            for (int mu = member; mu < m_units; mu += a.group, it++) {
                const int mt = mu * NCTA + rank;
                const int as = it & 1;
                const uint32_t aphase = (uint32_t)(it >> 1) & 1;
                const uint32_t q = (uint32_t)mt * TC_BM + qlocal;
                const bool q_valid = q < (uint32_t)a.batch;
                const float qinv = q_valid ? a.q_inv[q] : 0.f;
                // gate on the un-normalised product: acc * (1/|row|) >= gate / (1/|q|), one multiply and one compare per
                // row.  The factor of an unwanted row is NaN, which fails every comparison; a zero query and an
                // absent one (padding of the last query tile) take nothing.
                const float gate = gate_s[mt * TC_BM + qlocal];
                float thr = __int_as_float(0x7f800000);
                if (q_valid) {
                    if (qinv > 0.f) { thr = gate / qinv; thr -= fabsf(thr) * 9.5367431640625e-7f; }      // (a zero query takes nothing)
                }
                // per-query document sets are tested lazily, only for the few rows that pass the gate
                const uint32_t *qmask = (per_query && q_valid) ? a.doc_bitmap + (size_t)q * a.bitmap_stride : nullptr;
                unsigned long long *lst = a.cand + (size_t)q * a.cap;
                uint32_t *hq = a.hist + (size_t)q * kHistWords;
                mbar_wait(bar_tfull + 8 * as, aphase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + (uint32_t)as * TC_BN;
                // pass 1, branch-free: how many of the 256 rows pass this query's gate, and in which 32-row chunks
                uint32_t n_pass = 0, hits = 0;
#pragma unroll 1
                for (int c = 0; c < TC_BN / 32; c++) {
                    uint32_t r[32];
                    tmem_ld32(taddr + (uint32_t)c * 32u, r);
                    const float4 *f4 = reinterpret_cast<const float4 *>(inv_t + c * 32);
                    uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const float4 f = f4[j];
                        c0 += (acc_value<I8>(r[4 * j + 0]) * f.x >= thr) ? 1u : 0u;
                        c1 += (acc_value<I8>(r[4 * j + 1]) * f.y >= thr) ? 1u : 0u;
                        c2 += (acc_value<I8>(r[4 * j + 2]) * f.z >= thr) ? 1u : 0u;
                        c3 += (acc_value<I8>(r[4 * j + 3]) * f.w >= thr) ? 1u : 0u;
                    }
                    const uint32_t cc = (c0 + c1) + (c2 + c3);
                    n_pass += cc;
                    hits |= (cc ? 1u : 0u) << c;
                }
                // per-query document sets: recount exactly, visiting only the chunks that had gate survivors
                if (per_query) {
                    uint32_t wh = __reduce_or_sync(0xFFFFFFFFu, hits), kept = 0;
                    n_pass = 0;
                    while (wh) {
                        const int c = __ffs(wh) - 1;
                        wh &= wh - 1;
                        uint32_t r[32];
                        tmem_ld32(taddr + (uint32_t)c * 32u, r);
                        if ((hits >> c) & 1u) {
#pragma unroll
                            for (int j = 0; j < 32; j++) {
                                if (acc_value<I8>(r[j]) * inv_t[c * 32 + j] >= thr) {
                                    const uint32_t sl = slot_t[c * 32 + j];
                                    if ((qmask[sl >> 5] >> (sl & 31)) & 1u) { n_pass++; kept |= 1u << c; }
                                }
                            }
                        }
                        __syncwarp();
                    }
                    hits = kept;
                }
                // pass 2 (chunks with survivors only): one slot reservation per query, then plain stores and
                // fire-and-forget histogram increments
                uint32_t wh = __reduce_or_sync(0xFFFFFFFFu, hits);
                if (wh) {
                    uint32_t pos = n_pass ? atomicAdd(a.cand_count + q, n_pass) : 0u;
                    while (wh) {
                        const int c = __ffs(wh) - 1;
                        wh &= wh - 1;
                        uint32_t r[32];
                        tmem_ld32(taddr + (uint32_t)c * 32u, r);
                        if ((hits >> c) & 1u) {
#pragma unroll
                            for (int j = 0; j < 32; j++) {
                                const float pv = acc_value<I8>(r[j]) * inv_t[c * 32 + j];
                                if (pv >= thr) {
                                    if (qmask) {
                                        const uint32_t sl = slot_t[c * 32 + j];
                                        if (!((qmask[sl >> 5] >> (sl & 31)) & 1u)) continue;
                                    }
                                    if (pos < (uint32_t)a.cap) {
                                        const float v = pv * qinv;
                                        lst[pos] = ((unsigned long long)score_to_ordered(v) << 32) | (unsigned long long)(n0 + c * 32 + j);
                                        int hc, hf;
                                        hist_bins(fminf(1.f, fmaxf(-1.f, v)), hc, hf);
                                        atomicAdd(hq + 64 + hc * 64 + hf, 1u);
                                        atomicAdd(hq + hc, 1u);
                                    } else {
                                        *a.overflow = 1u;
                                    }
                                    pos++;
                                }
                            }
                        }
                        __syncwarp();       // the next TMEM load is warp-wide
                    }
                }
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
                    if (NCTA == 2) mbar_arrive_cluster(lead_tempty + 8 * as); else mbar_arrive(bar_tempty + 8 * as);
                }
            }
        }
    }

    // a pair leaves together: the leader's MMAs read the other CTA's shared memory and write its TMEM
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncwarp();
    if (NCTA == 2) cluster_sync_all(); else __syncthreads();
    if (warp == kMmaWarp) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (NCTA == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TC_TMEM_COLS) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TC_TMEM_COLS) : "memory");
    }
}

// ---------------------------------------------------------------------------
// The same contraction for FEW queries (batch <= 64, BASELINE configs[2]'s b = 64), with the roles the other way round:
// A = a tile of 128 corpus rows, B = the queries (N = 32 or 64 columns), accumulator = 128 lanes (rows) x N columns.
// Why a second orientation: with a handful of queries the kernel above still multiplies a full 128-query tile (half
// or more of it zero padding) against every row -- 128 tensor cycles per 8 KB of rows, a tensor pipe 40-75 % busy next
// to a saturated HBM, SM clocks pulled to 1.36 GHz by the power cap, and a third of every 48 KB pipeline stage spent on
// the query tile: 19.6 ms per pass at 10M x 3072 where the SIMT scan streams the same rows in 16.6.  Here an MMA costs
// 128*N/256 cycles (16-32 per 4 KB of rows), a stage is 16 KB of rows + N x 128 B of queries, and 8-10 stages are in
// flight.  Epilogue thread = one ROW: its N scores against the queries' gates (shared memory, broadcast reads).
// Candidates are rare, so each takes its slot with its own atomicAdd; everything downstream (lists, histograms,
// final gate, exact re-score, selection) is shared with the kernel above.
// ---------------------------------------------------------------------------
constexpr int TCR_MAX_N = 64;
constexpr uint32_t TCR_SMEM_BUDGET = 200 * 1024;                 // pipeline stages
constexpr uint32_t TCR_SMEM_BYTES = TCR_SMEM_BUDGET + 1024 + 256 + 2 * TCR_MAX_N * 4;

// NCTA = 2: CTA pairs (tcgen05 cta_group::2, M = 256 rows): each CTA stages its own 128 rows and HALF of the queries, so the
// query operand costs N/2 x 128 B of L2 traffic per 16 KB of rows -- at N = 64 the single-CTA form moves 1.5 bytes from L2
// for every byte from HBM, which is what holds it at 19.7 ms when the power cap pulls the L2 clock down (N = 32: 17.6).
template <int OP, int NCTA>
__global__ void __launch_bounds__(TC_THREADS, 1)
tc_rows_candidates_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_x,
                          const BatchArgs a, const int N /* 32 or 64 */)
{
    constexpr bool BF16 = OP == OP_BF16, I8 = OP == OP_I8;
    const int q_rows = N / NCTA;                                  // rows of the query operand this CTA stages
    const uint32_t stage_bytes = TC_A_BYTES + (uint32_t)q_rows * 128u;
    const int stages = (int)(TCR_SMEM_BUDGET / stage_bytes);
    extern __shared__ unsigned char smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = base + TCR_SMEM_BUDGET;
    const uint32_t bar_full = bars, bar_empty = bars + 8 * 12;          // up to 12 stages
    const uint32_t bar_tfull = bars + 16 * 12, bar_tempty = bar_tfull + 16;
    const uint32_t tmem_slot = bar_tempty + 16;
    volatile uint32_t *tmem_slot_ptr = reinterpret_cast<volatile uint32_t *>(smem_raw + (tmem_slot - smem_u32(smem_raw)));
    float *thr_s = reinterpret_cast<float *>(smem_raw + (bars + 256 - smem_u32(smem_raw)));     // [TCR_MAX_N] gate / (1/|q|)
    float *qinv_s = thr_s + TCR_MAX_N;                                                            // [TCR_MAX_N]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = NCTA == 2 ? (int)cluster_ctarank() : 0;        // 0 = leader of the pair
    const int unit = (int)blockIdx.x / NCTA, n_units = (int)gridDim.x / NCTA;
    const int n_tiles_total = (int)((a.n_rows + TC_BM * NCTA - 1) / (TC_BM * NCTA));      // tiles of 128 (256) rows
    const int per_unit = (n_tiles_total + n_units - 1) / n_units;
    const int t_begin = unit * per_unit;
    const int t_end = min(n_tiles_total, t_begin + per_unit);
    const int bk = I8 ? 4 * TC_BK : BF16 ? 2 * TC_BK : TC_BK;     // elements per 128-byte stage row
    const int kb_count = (a.K + bk - 1) / bk;
    const uint32_t tmem_cols = 2u * (uint32_t)N;                   // two accumulators: 64 or 128 columns
    const uint32_t idesc = (I8 ? ((2u << 4) | (1u << 7) | (1u << 10)) : BF16 ? ((1u << 4) | (1u << 7) | (1u << 10)) : ((1u << 4) | (2u << 7) | (2u << 10))) |
                           ((uint32_t)(N >> 3) << 17) | ((uint32_t)((NCTA * TC_BM) >> 4) << 24);

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; s++) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
        for (int s = 0; s < 2; s++) { mbar_init(bar_tfull + 8 * s, 1); mbar_init(bar_tempty + 8 * s, 4 * NCTA); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kMmaWarp) {
        if (NCTA == 2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tmem_cols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tmem_cols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    if (NCTA == 2) cluster_sync_all(); else __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot_ptr;
    // barriers that live on the leader only: "stage full" (both CTAs' loads land on it) and "accumulator drained"
    const uint32_t lead_full = NCTA == 2 ? mapa_rank(bar_full, 0) : bar_full;
    const uint32_t lead_tempty = NCTA == 2 ? mapa_rank(bar_tempty, 0) : bar_tempty;

    if (warp == kProdWarp) {
        if (lane == 0) {
            // (Issuing the stages in bursts of 2 / 4 / 8, so that consecutive 128-byte pieces of the same rows reach the
            // memory controllers together, measured 17.6 / 18.1 / 21.2 ms against 17.6: depth in flight matters, not that.)
            int stage = 0; uint32_t phase = 0;
            for (int t = t_begin; t < t_end; t++) {
                for (int kb = 0; kb < kb_count; kb++) {
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    const uint32_t sa = base + stage * stage_bytes, sb = sa + TC_A_BYTES;
                    if (NCTA == 2) {
                        if (rank == 0) mbar_expect_tx(bar_full + 8 * stage, 2 * stage_bytes);
                        tma_load_2d_pair(sa, &map_x, lead_full + 8 * stage, kb * bk, (t * 2 + rank) * TC_BM);
                        tma_load_2d_pair(sb, &map_q, lead_full + 8 * stage, kb * bk, rank * q_rows);
                    } else {
                        mbar_expect_tx(bar_full + 8 * stage, stage_bytes);
                        tma_load_2d(sa, &map_x, bar_full + 8 * stage, kb * bk, t * TC_BM);
                        tma_load_2d(sb, &map_q, bar_full + 8 * stage, kb * bk, 0);
                    }
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == kMmaWarp) {
        if (lane == 0 && rank == 0) {
            int stage = 0; uint32_t phase = 0;
            int it = 0;
            for (int t = t_begin; t < t_end; t++, it++) {
                const int as = it & 1;
                const uint32_t aphase = (uint32_t)(it >> 1) & 1;
                mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_d = tmem_base + (uint32_t)as * (uint32_t)N;
                for (int kb = 0; kb < kb_count; kb++) {
                    mbar_wait(bar_full + 8 * stage, phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = base + stage * stage_bytes, sb = sa + TC_A_BYTES;
                    const uint64_t da0 = smem_desc_sw128(sa), db0 = smem_desc_sw128(sb);
#pragma unroll
                    for (int k = 0; k < TC_BK / TC_UMMA_K; k++) {     // 4 MMAs of 32 bytes of K each, any kind
                        const uint64_t da_ = da0 + (uint64_t)(k * 2), db_ = db0 + (uint64_t)(k * 2);
                        const uint32_t acc = (uint32_t)((kb | k) != 0);
                        if (NCTA == 2) {
                            if (I8) umma_i8_pair(tmem_d, da_, db_, idesc, acc);
                            else if (BF16) umma_bf16_pair(tmem_d, da_, db_, idesc, acc);
                            else umma_tf32_pair(tmem_d, da_, db_, idesc, acc);
                        } else {
                            if (I8) umma_i8(tmem_d, da_, db_, idesc, acc);
                            else if (BF16) umma_bf16(tmem_d, da_, db_, idesc, acc);
                            else umma_tf32(tmem_d, da_, db_, idesc, acc);
                        }
                    }
                    if (NCTA == 2) umma_commit_pair(bar_empty + 8 * stage); else umma_commit(bar_empty + 8 * stage);
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
                if (NCTA == 2) umma_commit_pair(bar_tfull + 8 * as); else umma_commit(bar_tfull + 8 * as);
            }
        }
    } else {
        // ===== epilogue warps 0..3: thread = row `rlocal` of every row tile; thread q < N also keeps query q's gate =====
        const uint32_t quad = (uint32_t)warp & 3u;
        const uint32_t rlocal = quad * 32u + (uint32_t)lane;
        const bool per_query = a.doc_bitmap && a.bitmap_stride;
        const bool my_query = rlocal < (uint32_t)N;
        const bool q_valid = my_query && rlocal < (uint32_t)a.batch;
        const float my_qinv = q_valid ? a.q_inv[rlocal] : 0.f;
        const float my_eps = my_query ? batch_eps(a, rlocal) : 0.f;
        // gate -> threshold on the un-normalised product (see the kernel above): acc * (1/|row|) >= gate / (1/|q|)
        auto thr_of = [&](float gate) {
            float thr = __int_as_float(0x7f800000);                 // an absent query (padding) takes nothing
            if (q_valid) {
                if (my_qinv > 0.f) { thr = gate / my_qinv; thr -= fabsf(thr) * 9.5367431640625e-7f; }      // (a zero query takes nothing)
            }
            return thr;
        };
        if (my_query) { thr_s[rlocal] = thr_of(a.min_sim - my_eps); qinv_s[rlocal] = my_qinv; }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        int it = 0, visited = 0, next_refresh = 1;
        for (int t = t_begin; t < t_end; t++, it++, visited++) {
            const uint32_t row = (uint32_t)(t * NCTA + rank) * TC_BM + rlocal;
            float inv = __int_as_float(0x7fc00000);                 // NaN: an absent, zero or filtered row passes no gate
            uint32_t slot = 0;
            if (row < a.n_rows) {
                const float f = a.inv_norm[row];
                if (f > 0.f) inv = f;
                if (a.doc_bitmap) {
                    slot = a.doc_slot[row];
                    if (!per_query && !((a.doc_bitmap[slot >> 5] >> (slot & 31)) & 1u)) inv = __int_as_float(0x7fc00000);
                }
            }
            if (visited == next_refresh) {                          // tighten the gates: often at first, then every 32 row tiles
                next_refresh = visited < 8 ? visited + 1 : (visited < 32 ? visited + 4 : visited + 32);
                asm volatile("bar.sync 1, 128;" ::: "memory");     // everybody has read the old thresholds
                if (q_valid) {
                    const float lb = hist_kth_lower_bound(a.hist + (size_t)rlocal * kHistWords, a.k);
                    thr_s[rlocal] = thr_of(fmaxf(a.min_sim - my_eps, lb - 2.f * my_eps));
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
            }
            const int as = it & 1;
            const uint32_t aphase = (uint32_t)(it >> 1) & 1;
            mbar_wait(bar_tfull + 8 * as, aphase);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + (uint32_t)as * (uint32_t)N;
            for (int c = 0; c < N / 32; c++) {
                uint32_t r[32];
                tmem_ld32(taddr + (uint32_t)c * 32u, r);
                const float4 *t4 = reinterpret_cast<const float4 *>(thr_s + c * 32);
                uint32_t pass = 0;
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const float4 th = t4[j];
                    pass |= (acc_value<I8>(r[4 * j + 0]) * inv >= th.x ? 1u : 0u) << (4 * j + 0);
                    pass |= (acc_value<I8>(r[4 * j + 1]) * inv >= th.y ? 1u : 0u) << (4 * j + 1);
                    pass |= (acc_value<I8>(r[4 * j + 2]) * inv >= th.z ? 1u : 0u) << (4 * j + 2);
                    pass |= (acc_value<I8>(r[4 * j + 3]) * inv >= th.w ? 1u : 0u) << (4 * j + 3);
                }
                if (pass) {
#pragma unroll
                    for (int j = 0; j < 32; j++) {
                        if (!((pass >> j) & 1u)) continue;
                        const uint32_t q = (uint32_t)(c * 32 + j);
                        if (per_query) {
                            const uint32_t *qmask = a.doc_bitmap + (size_t)q * a.bitmap_stride;
                            if (!((qmask[slot >> 5] >> (slot & 31)) & 1u)) continue;
                        }
                        const uint32_t pos = atomicAdd(a.cand_count + q, 1u);
                        if (pos < (uint32_t)a.cap) {
                            const float v = acc_value<I8>(r[j]) * inv * qinv_s[q];
                            a.cand[(size_t)q * a.cap + pos] = ((unsigned long long)score_to_ordered(v) << 32) | (unsigned long long)row;
                            uint32_t *hq = a.hist + (size_t)q * kHistWords;
                            int hc, hf;
                            hist_bins(fminf(1.f, fmaxf(-1.f, v)), hc, hf);
                            atomicAdd(hq + 64 + hc * 64 + hf, 1u);
                            atomicAdd(hq + hc, 1u);
                        } else {
                            *a.overflow = 1u;
                        }
                    }
                }
                __syncwarp();                                       // the next TMEM load is warp-wide
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                if (NCTA == 2) mbar_arrive_cluster(lead_tempty + 8 * as); else mbar_arrive(bar_tempty + 8 * as);
            }
        }
    }

    // a pair leaves together: the leader's MMAs read the other CTA's shared memory and write its TMEM
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncwarp();
    if (NCTA == 2) cluster_sync_all(); else __syncthreads();
    if (warp == kMmaWarp) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (NCTA == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
    }
}

// Final gate per query, read by the re-score kernel.
__global__ void batch_final_gate_kernel(const uint32_t *hist, int batch, int k, const BatchArgs a, uint32_t *gate_ordered)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= batch) return;
    const float lb = hist_kth_lower_bound(hist + (size_t)q * kHistWords, k);
    const float eps = batch_eps(a, (uint32_t)q);
    gate_ordered[q] = score_to_ordered(fmaxf(a.min_sim - eps, lb - 2.f * eps));
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// rows x K fp32 row-major -> boxes of [box_rows][32 floats], 128-byte swizzle, zero fill out of bounds
bool make_map_bf16(CUtensorMap *map, const uint16_t *ptr, uint64_t rows, uint64_t K, uint32_t box_rows)
{
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {K, rows};
    cuuint64_t strides[1] = {K * 2};
    cuuint32_t box[2] = {(cuuint32_t)(2 * TC_BK), box_rows};        // 64 bf16 = 128 bytes
    cuuint32_t estr[2] = {1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<uint16_t *>(ptr), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool make_map(CUtensorMap *map, const float *ptr, uint64_t rows, uint64_t K, uint32_t box_rows)
{
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {K, rows};
    cuuint64_t strides[1] = {K * 4};
    cuuint32_t box[2] = {(cuuint32_t)TC_BK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(ptr), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

size_t tc_assign_scratch_bytes(uint32_t M) { return (size_t)M * sizeof(unsigned long long); }

cudaError_t tc_launch_assign(const float *mat, uint32_t M, int D, const float *cent, int L,
                             unsigned long long *packed_scratch, uint32_t *out_assign, cudaStream_t stream)
{
    if (M == 0) return cudaSuccess;
    if ((reinterpret_cast<uintptr_t>(mat) & 15) || (reinterpret_cast<uintptr_t>(cent) & 15) || (D & 3)) return cudaErrorInvalidValue;
    CUtensorMap map_a, map_b;
    if (!make_map(&map_a, mat, M, (uint64_t)D, TC_BM) || !make_map(&map_b, cent, (uint64_t)L, (uint64_t)D, TC_BN))
        return cudaErrorInvalidValue;
    cudaError_t e = cudaMemsetAsync(packed_scratch, 0, tc_assign_scratch_bytes(M), stream);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(tc_rowargmax_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_BYTES);
    if (e != cudaSuccess) return e;
    const int n_tiles = (L + TC_BN - 1) / TC_BN;
    // N partitions: sibling CTAs share one A tile through L2 and each sweeps at most 4 B tiles
    int n_parts = (n_tiles + 3) / 4;
    if (n_parts < 1) n_parts = 1;
    const uint32_t m_tiles = (M + TC_BM - 1) / TC_BM;
    tc_rowargmax_kernel<<<m_tiles * (uint32_t)n_parts, TC_THREADS, TC_SMEM_BYTES, stream>>>(map_a, map_b, M, (uint32_t)L, D, n_parts, packed_scratch);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    unpack_assign_kernel<<<148 * 4, 256, 0, stream>>>(packed_scratch, M, out_assign);
    return cudaGetLastError();
}

// How many slices of the dimension the score GEMM is cut into: enough (tile, slice) CTAs to occupy the chip, at most 16.
int tc_scores_splits(int M, int N, int D, int sm_count)
{
    const int tiles = ((M + TC_BM - 1) / TC_BM) * ((N + TC_BN - 1) / TC_BN);
    const int kb_total = (D + TC_BK - 1) / TC_BK;
    int ks = sm_count / (tiles > 0 ? tiles : 1);
    if (ks > 16) ks = 16;
    if (ks > kb_total) ks = kb_total;
    if (ks < 1) ks = 1;
    const int per = (kb_total + ks - 1) / ks;
    return (kb_total + per - 1) / per;                       // slices that actually receive work
}

cudaError_t tc_launch_scores(const float *A, int M, const float *B, int N, int D, int ks, float *out, cudaStream_t stream)
{
    if (M <= 0 || N <= 0) return cudaSuccess;
    if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) || (reinterpret_cast<uintptr_t>(out) & 15) || (D & 3))
        return cudaErrorInvalidValue;
    CUtensorMap map_a, map_b;
    if (!make_map(&map_a, A, (uint64_t)M, (uint64_t)D, TC_BM) || !make_map(&map_b, B, (uint64_t)N, (uint64_t)D, TC_BN))
        return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(tc_scores_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_BYTES);
    if (e != cudaSuccess) return e;
    const int n_tiles = (N + TC_BN - 1) / TC_BN, m_tiles = (M + TC_BM - 1) / TC_BM;
    const int kb_total = (D + TC_BK - 1) / TC_BK;
    const int per = (kb_total + ks - 1) / ks;
    tc_scores_kernel<<<dim3((unsigned)(m_tiles * n_tiles), (unsigned)ks), TC_THREADS, TC_SMEM_BYTES, stream>>>(
        map_a, map_b, (uint32_t)M, (uint32_t)N, kb_total, per, n_tiles, out);
    return cudaGetLastError();
}

// Entries per query list.  The first row tile of every CTA is appended unfiltered (no gate exists yet), so a
// list must hold at least grid x 256 entries; later appends pass the shared gate.
int tc_batch_cap(int k) { (void)k; return 65536; }
int tc_batch_m_tiles(int batch) { return (batch + TC_BM - 1) / TC_BM; }
size_t tc_batch_hist_words(int batch) { return (size_t)batch * kHistWords; }

bool make_map_i8(CUtensorMap *map, const int8_t *ptr, uint64_t rows, uint64_t K, uint32_t box_rows)
{
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {K, rows};
    cuuint64_t strides[1] = {K};
    cuuint32_t box[2] = {(cuuint32_t)(4 * TC_BK), box_rows};        // 128 int8 = 128 bytes
    cuuint32_t estr[2] = {1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<int8_t *>(ptr), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
cudaError_t tc_launch_batch_candidates(const float *queries, int batch, const float *q_inv, const float *mat,
                                       uint32_t n_rows, int D, const float *inv_norm, const uint32_t *doc_slot,
                                       const uint32_t *doc_bitmap, uint32_t bitmap_stride, int k, float min_sim, float eps, int sm_count,
                                       unsigned long long *cand, uint32_t *cand_count, uint32_t *hist,
                                       uint32_t *gate_ordered, uint32_t *overflow, uint32_t *group_sync,
                                       cudaStream_t stream, const uint16_t *q_hi, const uint16_t *x_hi, const TcInt8Operands *i8)
{
    const bool bf16 = q_hi && x_hi;       // split-plane store: contract the upper planes as bf16 (kind::f16)
    // int8 sidecar: contract the quantised queries and rows (kind::i8); `inv_norm` / `q_inv` then carry scale/|row| and
    // scale/|q|, `eps` the rows' quantisation-error bound, i8->q_err the queries' own
    if (i8 && (bf16 || !i8->queries || !i8->rows || !i8->q_err)) return cudaErrorInvalidValue;
    if (batch > DA_TC_MAX_BATCH) return cudaErrorInvalidValue;
    // Few queries: rows as the M operand, the queries as N = 32 or 64 columns (tc_rows_candidates_kernel);
    // DA_TC_QUERY_TILES=1 keeps the query-tile form below for comparisons.
    static const bool query_tiles_only = [] { const char *v = getenv("DA_TC_QUERY_TILES"); return v && v[0] == '1'; }();
    if (batch <= TCR_MAX_N && !query_tiles_only) {
        const int N = batch <= 32 ? 32 : 64;
        // CTA pairs halve the query operand's L2 traffic (see the kernel); DA_TCR_PAIR = 0 never, 1 at N = 64 (default), 2 always
        static const int pair_env = [] { const char *v = getenv("DA_TCR_PAIR"); return v ? atoi(v) : 1; }();
        const int ncta = (pair_env == 2 || (pair_env == 1 && N == 64)) && sm_count >= 2 ? 2 : 1;
        const uint32_t q_box = (uint32_t)(N / ncta);
        CUtensorMap map_q, map_x;
        bool ok;
        if (i8) {
            if ((reinterpret_cast<uintptr_t>(i8->rows) & 15) || (reinterpret_cast<uintptr_t>(i8->queries) & 15) || (D & 15)) return cudaErrorInvalidValue;
            ok = make_map_i8(&map_q, i8->queries, (uint64_t)batch, (uint64_t)D, q_box) && make_map_i8(&map_x, i8->rows, (uint64_t)n_rows, (uint64_t)D, TC_BM);
        } else if (bf16) {
            if ((reinterpret_cast<uintptr_t>(x_hi) & 15) || (reinterpret_cast<uintptr_t>(q_hi) & 15) || (D & 7)) return cudaErrorInvalidValue;
            ok = make_map_bf16(&map_q, q_hi, (uint64_t)batch, (uint64_t)D, q_box) && make_map_bf16(&map_x, x_hi, (uint64_t)n_rows, (uint64_t)D, TC_BM);
        } else {
            if ((reinterpret_cast<uintptr_t>(mat) & 15) || (reinterpret_cast<uintptr_t>(queries) & 15) || (D & 3)) return cudaErrorInvalidValue;
            ok = make_map(&map_q, queries, (uint64_t)batch, (uint64_t)D, q_box) && make_map(&map_x, mat, (uint64_t)n_rows, (uint64_t)D, TC_BM);
        }
        if (!ok) return cudaErrorInvalidValue;
        cudaError_t e = cudaMemsetAsync(cand_count, 0, (size_t)batch * 4, stream);
        if (e != cudaSuccess) return e;
        e = cudaMemsetAsync(hist, 0, tc_batch_hist_words(batch) * 4, stream);
        if (e != cudaSuccess) return e;
        BatchArgs a{};
        a.n_rows = n_rows; a.K = D; a.batch = batch; a.k = k; a.cap = tc_batch_cap(k); a.group = 1;
        a.min_sim = min_sim; a.eps = eps; a.q_err = i8 ? i8->q_err : nullptr;
        a.inv_norm = inv_norm; a.doc_slot = doc_slot; a.doc_bitmap = doc_bitmap; a.bitmap_stride = bitmap_stride; a.q_inv = q_inv;
        a.cand = cand; a.cand_count = cand_count; a.hist = hist; a.overflow = overflow; a.group_sync = group_sync;
        const void *kernel = ncta == 2
            ? (i8 ? (const void *)tc_rows_candidates_kernel<OP_I8, 2> : bf16 ? (const void *)tc_rows_candidates_kernel<OP_BF16, 2>
                                                                              : (const void *)tc_rows_candidates_kernel<OP_TF32, 2>)
            : (i8 ? (const void *)tc_rows_candidates_kernel<OP_I8, 1> : bf16 ? (const void *)tc_rows_candidates_kernel<OP_BF16, 1>
                                                                              : (const void *)tc_rows_candidates_kernel<OP_TF32, 1>);
        e = ensure_dynamic_smem(kernel, TCR_SMEM_BYTES);
        if (e != cudaSuccess) return e;
        cudaLaunchAttribute attrs[1];
        cudaLaunchConfig_t cfg{};
        cfg.blockDim = dim3(TC_THREADS); cfg.dynamicSmemBytes = TCR_SMEM_BYTES; cfg.stream = stream;
        cfg.attrs = attrs; cfg.numAttrs = 0;
        int n_units = sm_count / ncta;
        if (ncta == 2) {
            attrs[0].id = cudaLaunchAttributeClusterDimension;
            attrs[0].val.clusterDim.x = 2; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
            cfg.numAttrs = 1;
            cfg.gridDim = dim3((unsigned)(2 * n_units));
            int max_pairs = 0;
            e = cudaOccupancyMaxActiveClusters(&max_pairs, kernel, &cfg);
            if (e != cudaSuccess) return e;
            if (max_pairs < 1) return cudaErrorLaunchOutOfResources;
            if (n_units > max_pairs) n_units = max_pairs;
        }
        const uint32_t n_tiles = (n_rows + TC_BM * ncta - 1) / (TC_BM * ncta);
        if ((uint32_t)n_units > n_tiles) n_units = n_tiles ? (int)n_tiles : 1;
        cfg.gridDim = dim3((unsigned)(n_units * ncta));
        int n_arg = N;
        void *args[] = {(void *)&map_q, (void *)&map_x, (void *)&a, (void *)&n_arg};
        e = cudaLaunchKernelExC(&cfg, kernel, args);
        if (e != cudaSuccess) return e;
        batch_final_gate_kernel<<<(batch + 127) / 128, 128, 0, stream>>>(hist, batch, k, a, gate_ordered);
        return cudaGetLastError();
    }
    const int m_tiles = tc_batch_m_tiles(batch);
    // CTA pairs need two query tiles to work on; DA_TC_SINGLE_CTA=1 keeps the single-CTA form (for comparisons)
    static const bool single_only = [] { const char *v = getenv("DA_TC_SINGLE_CTA"); return v && v[0] == '1'; }();
    const int ncta = (m_tiles >= 2 && !single_only) ? 2 : 1;
    CUtensorMap map_q, map_x;
    if (i8) {
        if ((reinterpret_cast<uintptr_t>(i8->rows) & 15) || (reinterpret_cast<uintptr_t>(i8->queries) & 15) || (D & 15)) return cudaErrorInvalidValue;
        if (!make_map_i8(&map_q, i8->queries, (uint64_t)batch, (uint64_t)D, TC_BM) ||
            !make_map_i8(&map_x, i8->rows, (uint64_t)n_rows, (uint64_t)D, TC_BN / ncta))
            return cudaErrorInvalidValue;
    } else if (bf16) {
        if ((reinterpret_cast<uintptr_t>(x_hi) & 15) || (reinterpret_cast<uintptr_t>(q_hi) & 15) || (D & 7)) return cudaErrorInvalidValue;
        if (!make_map_bf16(&map_q, q_hi, (uint64_t)batch, (uint64_t)D, TC_BM) ||
            !make_map_bf16(&map_x, x_hi, (uint64_t)n_rows, (uint64_t)D, TC_BN / ncta))
            return cudaErrorInvalidValue;
    } else {
        if ((reinterpret_cast<uintptr_t>(mat) & 15) || (reinterpret_cast<uintptr_t>(queries) & 15) || (D & 3)) return cudaErrorInvalidValue;
        if (!make_map(&map_q, queries, (uint64_t)batch, (uint64_t)D, TC_BM) ||
            !make_map(&map_x, mat, (uint64_t)n_rows, (uint64_t)D, TC_BN / ncta))
            return cudaErrorInvalidValue;
    }
    cudaError_t e = cudaMemsetAsync(cand_count, 0, (size_t)batch * 4, stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(hist, 0, tc_batch_hist_words(batch) * 4, stream);
    if (e != cudaSuccess) return e;
    BatchArgs a{};
    a.n_rows = n_rows; a.K = D; a.batch = batch; a.k = k; a.cap = tc_batch_cap(k);
    a.min_sim = min_sim; a.eps = eps; a.q_err = i8 ? i8->q_err : nullptr;
    a.inv_norm = inv_norm; a.doc_slot = doc_slot; a.doc_bitmap = doc_bitmap; a.bitmap_stride = bitmap_stride; a.q_inv = q_inv;
    a.cand = cand; a.cand_count = cand_count; a.hist = hist; a.overflow = overflow;
    const void *kernel = ncta == 2
        ? (i8 ? (const void *)tc_batch_candidates_kernel<OP_I8, 2> : bf16 ? (const void *)tc_batch_candidates_kernel<OP_BF16, 2>
                                                                          : (const void *)tc_batch_candidates_kernel<OP_TF32, 2>)
        : (i8 ? (const void *)tc_batch_candidates_kernel<OP_I8, 1> : bf16 ? (const void *)tc_batch_candidates_kernel<OP_BF16, 1>
                                                                          : (const void *)tc_batch_candidates_kernel<OP_TF32, 1>);
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TCB_SMEM_BYTES);
    if (e != cudaSuccess) return e;

    cudaLaunchAttribute attrs[1];
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(TC_THREADS); cfg.dynamicSmemBytes = TCB_SMEM_BYTES; cfg.stream = stream;
    cfg.attrs = attrs; cfg.numAttrs = 0;
    int n_units = sm_count / ncta;           // CTAs, or CTA pairs, that can be resident at once
    if (ncta == 2) {
        attrs[cfg.numAttrs].id = cudaLaunchAttributeClusterDimension;
        attrs[cfg.numAttrs].val.clusterDim.x = 2; attrs[cfg.numAttrs].val.clusterDim.y = 1; attrs[cfg.numAttrs].val.clusterDim.z = 1;
        cfg.numAttrs++;
        cfg.gridDim = dim3((unsigned)(2 * n_units));
        int max_pairs = 0;
        e = cudaOccupancyMaxActiveClusters(&max_pairs, kernel, &cfg);
        if (e != cudaSuccess) return e;
        if (max_pairs < 1) return cudaErrorLaunchOutOfResources;
        if (n_units > max_pairs) n_units = max_pairs;
    }
    const uint32_t n_tiles = (n_rows + TC_BN - 1) / TC_BN;
    // Row tiles in flight across the chip (one per group) plus the queries should sit in a fraction of the
    // 126 MB L2; otherwise every query tile pulls its row tile from HBM again.  Measured on cfg4 (4096 queries,
    // 1.25M x 1536 rows, free-running groups of 1/4/16/32 CTAs): 274/188/152/102 GB of DRAM reads for a 7.7 GB shard;
    // with the members of a group in lockstep (below): 44 GB at group 4, 22.7 GB at group 8 (profiles/README.md).
    // The budget is 55 MB, not the 126 MB on the data sheet: the L2 is two halves of 63 MB and a line that both dies
    // touch lives in both, so a live set that every SM shares sees about half of it (round 2 sweep: budget 90 MB ->
    // group 4 -> 28.6 ms per configs[3] shard, 55 MB -> group 8 -> 26.3 ms).
    const int m_units = (m_tiles + ncta - 1) / ncta;
    const double esz = i8 ? 1.0 : bf16 ? 2.0 : 4.0;
    const double tile_mb = (double)TC_BN * D * esz / 1e6, q_mb = (double)batch * D * esz / 1e6;
    // DA_TC_L2_MB / DA_TC_GROUP: measurement knobs (the budget for the live set, or the group size itself)
    static const double l2_budget_mb = [] { const char *v = getenv("DA_TC_L2_MB"); return v ? atof(v) : 55.0; }();
    static const int group_env = [] { const char *v = getenv("DA_TC_GROUP"); return v ? atoi(v) : 0; }();
    int group = 1;
    while (group < m_units && group < n_units && 2.0 * (n_units / group) * tile_mb + q_mb > l2_budget_mb) group++;   // two live row tiles per group (lockstep window)
    // the query tiles are dealt round-robin to the members: a group size that divides their number keeps the members level
    // (measured on a configs[3] shard, CTA pairs, 16 tile pairs: group 4 / 8 / 12 / 16 = 28.6 / 26.3 / 34.6 / 26.6 ms)
    while (group < m_units && m_units % group != 0) group++;
    if (group_env > 0) group = group_env;
    if (group > m_units) group = m_units;
    if (group > n_units) group = n_units;
    int n_groups = n_units / group;
    if ((uint32_t)n_groups > n_tiles) n_groups = (int)n_tiles;
    if (n_groups < 1) n_groups = 1;
    a.group = group;
    a.group_sync = group_sync;
    cfg.gridDim = dim3((unsigned)(n_groups * group * ncta));
    if (group > 1) {
        e = cudaMemsetAsync(group_sync, 0, (size_t)n_groups * 4, stream);
        if (e != cudaSuccess) return e;
    }
    void *args[] = {(void *)&map_q, (void *)&map_x, (void *)&a};
    e = cudaLaunchKernelExC(&cfg, kernel, args);
    if (e != cudaSuccess) return e;
    batch_final_gate_kernel<<<(batch + 127) / 128, 128, 0, stream>>>(hist, batch, k, a, gate_ordered);
    return cudaGetLastError();
}

namespace {
__global__ void truncate_bf16_kernel(const float *src, size_t n, uint16_t *dst)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = (uint16_t)(__float_as_uint(src[i]) >> 16);
}
}  // namespace

cudaError_t tc_launch_truncate_bf16(const float *src, size_t n, uint16_t *dst, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    truncate_bf16_kernel<<<148, 256, 0, stream>>>(src, n, dst);
    return cudaGetLastError();
}

cudaError_t tc_launch_gather_rows(const float *mat, int D, const uint32_t *rows, uint32_t n, float *out, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    gather_rows_kernel<<<148 * 4, 256, 0, stream>>>(mat, D, rows, n, out);
    return cudaGetLastError();
}

}  // namespace da
