// Device helpers shared by the scan kernels (da_scan.cu, da_split.cu): streaming loads, the per-warp bounded
// list with cached minimum, the CTA-wide bitonic sort.  Internal to the library.
#pragma once

#include "da_internal.h"

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
__device__ __forceinline__ void warp_list_insert(key_t64 *list, int K, int &cnt, key_t64 &wmin,
                                                 key_t64 key, int lane)
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
        return;
    }
    if (key <= wmin) return;
    // keys are unique (they embed the row), so exactly one slot holds wmin
    for (int i = lane; i < K; i += 32)
        if (list[i] == wmin) list[i] = key;
    __syncwarp();
    key_t64 m = ~0ull;
    for (int i = lane; i < K; i += 32) { key_t64 t = list[i]; m = t < m ? t : m; }
    wmin = warp_min_key(m);
}

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

inline int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

}  // namespace
}  // namespace da
