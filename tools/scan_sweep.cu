// Design-space probe for the streaming part of the fused scan (not part of the library):
// rows x 3072 fp32 dot one query, warp-per-R-rows, variants of threads / R / U / cache hints / grid.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o scan_sweep tools/scan_sweep.cu && ./scan_sweep
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

template <int HINT>
__device__ __forceinline__ float4 ldg(const float4 *p)
{
    float4 v;
    if constexpr (HINT == 0)
        asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    else if constexpr (HINT == 1)
        asm volatile("ld.global.nc.L1::no_allocate.L2::256B.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    else if constexpr (HINT == 3)
        asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    else
        asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

template <int THREADS, int R, int U, int HINT>
__global__ void __launch_bounds__(THREADS) sweep_kernel(const float *mat, const float *q, uint32_t n_rows, int D4, float *out)
{
    extern __shared__ float4 q_s[];
    for (int i = threadIdx.x; i < D4; i += THREADS) q_s[i] = reinterpret_cast<const float4 *>(q)[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int WARPS = THREADS / 32;
    const uint32_t n_groups = n_rows / R, gstride = gridDim.x * WARPS;
    float best = -1e30f;
    for (uint32_t g = blockIdx.x * WARPS + warp; g < n_groups; g += gstride) {
        const float4 *rp = reinterpret_cast<const float4 *>(mat) + (size_t)g * R * D4;
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; r++) acc[r] = 0.f;
        for (int c0 = 0; c0 < D4; c0 += 32 * U) {
            float4 v[R][U];
#pragma unroll
            for (int r = 0; r < R; r++)
#pragma unroll
                for (int u = 0; u < U; u++) v[r][u] = ldg<HINT>(rp + (size_t)r * D4 + c0 + u * 32 + lane);
#pragma unroll
            for (int u = 0; u < U; u++) {
                float4 w = q_s[c0 + u * 32 + lane];
#pragma unroll
                for (int r = 0; r < R; r++) {
                    acc[r] = fmaf(v[r][u].x, w.x, acc[r]); acc[r] = fmaf(v[r][u].y, w.y, acc[r]);
                    acc[r] = fmaf(v[r][u].z, w.z, acc[r]); acc[r] = fmaf(v[r][u].w, w.w, acc[r]);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            float s = acc[r];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
            best = fmaxf(best, s);
        }
    }
    if (lane == 0) out[blockIdx.x * WARPS + warp] = best;
}

struct f8 { float4 a, b; };
template <int HINT>
__device__ __forceinline__ f8 ldg8(const float *p)
{
    f8 v;
    if constexpr (HINT == 0)
        asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(v.a.x), "=f"(v.a.y), "=f"(v.a.z), "=f"(v.a.w), "=f"(v.b.x), "=f"(v.b.y), "=f"(v.b.z), "=f"(v.b.w) : "l"(p));
    else if constexpr (HINT == 1)
        asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(v.a.x), "=f"(v.a.y), "=f"(v.a.z), "=f"(v.a.w), "=f"(v.b.x), "=f"(v.b.y), "=f"(v.b.z), "=f"(v.b.w) : "l"(p));
    else
        asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.L2::256B.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(v.a.x), "=f"(v.a.y), "=f"(v.a.z), "=f"(v.a.w), "=f"(v.b.x), "=f"(v.b.y), "=f"(v.b.z), "=f"(v.b.w) : "l"(p));
    return v;
}

// 256-bit loads: a lane owns 8 consecutive floats; a warp instruction covers 1 KB of a row
template <int THREADS, int R, int U, int HINT>
__global__ void __launch_bounds__(THREADS) sweep8_kernel(const float *mat, const float *q, uint32_t n_rows, int D8, float *out)
{
    extern __shared__ float4 q_s[];
    for (int i = threadIdx.x; i < 2 * D8; i += THREADS) q_s[i] = reinterpret_cast<const float4 *>(q)[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int WARPS = THREADS / 32;
    const uint32_t n_groups = n_rows / R, gstride = gridDim.x * WARPS;
    float best = -1e30f;
    for (uint32_t g = blockIdx.x * WARPS + warp; g < n_groups; g += gstride) {
        const float *rp = mat + (size_t)g * R * D8 * 8;
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; r++) acc[r] = 0.f;
        for (int c0 = 0; c0 < D8; c0 += 32 * U) {
            f8 v[R][U];
#pragma unroll
            for (int r = 0; r < R; r++)
#pragma unroll
                for (int u = 0; u < U; u++) v[r][u] = ldg8<HINT>(rp + ((size_t)r * D8 + c0 + u * 32 + lane) * 8);
#pragma unroll
            for (int u = 0; u < U; u++) {
                float4 w0 = q_s[2 * (c0 + u * 32 + lane)], w1 = q_s[2 * (c0 + u * 32 + lane) + 1];
#pragma unroll
                for (int r = 0; r < R; r++) {
                    acc[r] = fmaf(v[r][u].a.x, w0.x, acc[r]); acc[r] = fmaf(v[r][u].a.y, w0.y, acc[r]);
                    acc[r] = fmaf(v[r][u].a.z, w0.z, acc[r]); acc[r] = fmaf(v[r][u].a.w, w0.w, acc[r]);
                    acc[r] = fmaf(v[r][u].b.x, w1.x, acc[r]); acc[r] = fmaf(v[r][u].b.y, w1.y, acc[r]);
                    acc[r] = fmaf(v[r][u].b.z, w1.z, acc[r]); acc[r] = fmaf(v[r][u].b.w, w1.w, acc[r]);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            float s = acc[r];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
            best = fmaxf(best, s);
        }
    }
    if (lane == 0) out[blockIdx.x * WARPS + warp] = best;
}

template <int THREADS, int R, int U, int HINT>
void run8(const char *name, const float *mat, const float *q, uint32_t n_rows, int D, float *out, int grid_mult, int sms)
{
    auto kern = sweep8_kernel<THREADS, R, U, HINT>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, D * 4);
    int grid = sms * grid_mult;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best_ms = 1e9f;
    for (int it = 0; it < 6; it++) {
        cudaEventRecord(e0);
        kern<<<grid, THREADS, D * 4>>>(mat, q, n_rows, D / 8, out);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (it >= 1 && ms < best_ms) best_ms = ms;
    }
    cudaError_t e = cudaGetLastError();
    printf("%-44s grid=%4d  %8.3f ms  %8.1f GB/s  %s\n", name, grid, best_ms, (double)n_rows * D * 4 / best_ms / 1e6,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
}

template <int THREADS, int R, int U, int HINT>
void run(const char *name, const float *mat, const float *q, uint32_t n_rows, int D, float *out, int grid_mult, int sms)
{
    auto kern = sweep_kernel<THREADS, R, U, HINT>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, D * 4);
    int grid = sms * grid_mult;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best_ms = 1e9f;
    for (int it = 0; it < 6; it++) {
        cudaEventRecord(e0);
        kern<<<grid, THREADS, D * 4>>>(mat, q, n_rows, D / 4, out);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (it >= 1 && ms < best_ms) best_ms = ms;
    }
    cudaError_t e = cudaGetLastError();
    printf("%-44s grid=%4d  %8.3f ms  %8.1f GB/s  %s\n", name, grid, best_ms, (double)n_rows * D * 4 / best_ms / 1e6,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main(int argc, char **argv)
{
    const int D = 3072;
    uint32_t n_rows = argc > 1 ? (uint32_t)atol(argv[1]) : 4000000u;
    n_rows -= n_rows % 64;
    float *mat, *q, *out;
    cudaMalloc(&mat, (size_t)n_rows * D * 4);
    cudaMalloc(&q, D * 4);
    cudaMalloc(&out, 1 << 22);
    cudaMemset(mat, 0, (size_t)n_rows * D * 4);
    cudaMemset(q, 0, D * 4);
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    printf("rows=%u (%.1f GB), %d SMs\n", n_rows, (double)n_rows * D * 4 / 1e9, sms);
#define RUN(T, R, U, H, G) run<T, R, U, H>("threads=" #T " R=" #R " U=" #U " hint=" #H, mat, q, n_rows, D, out, G, sms)
    RUN(512, 4, 4, 0, 1);
    RUN(512, 8, 2, 0, 1);
    RUN(512, 4, 4, 0, 2);
    RUN(512, 8, 3, 0, 1);
    RUN(1024, 4, 2, 0, 1);
#define RUN8(T, R, U, H, G) run8<T, R, U, H>("v8 threads=" #T " R=" #R " U=" #U " hint=" #H, mat, q, n_rows, D, out, G, sms)
    RUN8(512, 2, 6, 1, 1);
    RUN8(512, 2, 6, 0, 1);
    RUN8(512, 2, 6, 2, 1);
    RUN8(1024, 2, 2, 1, 1);
    RUN8(1024, 2, 2, 0, 1);
    RUN8(1024, 2, 3, 1, 1);
    RUN8(1024, 1, 6, 1, 1);
    RUN8(1024, 1, 4, 1, 1);
    RUN8(768, 2, 4, 1, 1);
    RUN8(768, 2, 3, 1, 1);
    RUN8(512, 1, 12, 1, 1);
    RUN8(512, 1, 6, 1, 1);
    RUN8(512, 2, 6, 1, 2);
    RUN8(256, 2, 6, 1, 2);
    RUN8(256, 2, 6, 1, 3);
    RUN8(256, 2, 6, 1, 4);
    RUN8(640, 2, 6, 1, 1);
    RUN8(512, 3, 4, 1, 1);
    return 0;
}
