// Probe: TMEM -> register read throughput (tcgen05.ld.32x32b.x32) per SM with 4 / 8 / 16 warps, and FMNMX3 (3-input min)
// throughput.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -o tmem_ld_probe tmem_ld_probe.cu ; prints cycles.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ float min3(float a, float b, float c) {
    float d;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// mode 0: loads only; mode 1: loads + min3 tree over the 32 values; mode 2: min3 only (register values)
template <int MODE>
__global__ void probe(int iters, float* out, long long* cyc) {
    __shared__ uint32_t s_base;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_base)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = s_base + ((uint32_t)((warp & 3) * 32) << 16);
    float acc = 3.0e38f;
    uint32_t r[32], q[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) { r[j] = __float_as_uint(1.0f + threadIdx.x + j); q[j] = r[j]; }
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        const uint32_t col = (uint32_t)((i * 32 + (warp >> 2) * 64) & 511) & ~31u;
        if (MODE != 2) {
            ld32(base + col, r);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        }
        if (MODE != 0) {
            float m[10];
#pragma unroll
            for (int j = 0; j < 10; ++j) m[j] = min3(__uint_as_float(r[3 * j]), __uint_as_float(r[3 * j + 1]), __uint_as_float(r[3 * j + 2]));
            const float a = min3(m[0], m[1], m[2]), b = min3(m[3], m[4], m[5]), c = min3(m[6], m[7], m[8]);
            const float d = min3(m[9], __uint_as_float(r[30]), __uint_as_float(r[31]));
            acc = min3(acc, min3(a, b, c), d);
            if (MODE == 2) {
#pragma unroll
                for (int j = 0; j < 32; ++j) r[j] = q[j] ^ (uint32_t)i;  // keeps the compiler from hoisting; LOP3 = ALU pipe too
            }
        }
    }
    const long long t1 = clock64();
    if (MODE == 0) {
#pragma unroll
        for (int j = 0; j < 32; ++j) acc = fminf(acc, __uint_as_float(r[j]));
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(s_base), "r"(512u) : "memory");
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    const int iters = 4096;
    for (int mode = 0; mode < 3; ++mode)
        for (int warps = 4; warps <= 16; warps *= 2) {
            for (int rep = 0; rep < 2; ++rep) {
                if (mode == 0) probe<0><<<148, warps * 32>>>(iters, out, cyc);
                if (mode == 1) probe<1><<<148, warps * 32>>>(iters, out, cyc);
                if (mode == 2) probe<2><<<148, warps * 32>>>(iters, out, cyc);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("mode %d warps %d: %s\n", mode, warps, cudaGetErrorString(e)); return 1; }
            }
            long long h[148];
            cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
            const double c = (double)h[0] / iters;
            printf("mode %d (%s) warps %2d: %.1f cycles per iteration per warp; per SM: %.1f B/clk TMEM, %.2f values/clk\n", mode,
                   mode == 0 ? "ld only" : (mode == 1 ? "ld + min3 tree" : "min3 tree + 32 LOP3"), warps, c, mode == 2 ? 0.0 : warps * 4096.0 / c,
                   warps * 1024.0 / c);
        }
    return 0;
}
