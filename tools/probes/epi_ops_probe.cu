// Probe: per-SM throughput of the chain kernel's epilogue instructions: F2FP (cvt.rn.bf16x2.f32), HMUL2.BF16, HMNMX2.BF16, FADD2.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o epi_ops_probe epi_ops_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int KIND>
__global__ void __launch_bounds__(512) probe(int iters, uint32_t* out, long long* cyc) {
    uint32_t h[8];
    float f[16];
    unsigned long long v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        h[i] = 0x3f803f80u + threadIdx.x + i;
        f[2 * i] = 1.f + threadIdx.x + i; f[2 * i + 1] = 2.f + i;
        v[i] = ((unsigned long long)__float_as_uint(1.f + i) << 32) | __float_as_uint(0.5f + threadIdx.x);
    }
    const unsigned long long t2 = ((unsigned long long)__float_as_uint(1e-7f) << 32) | __float_as_uint(1e-7f);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (KIND == 0) { asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(f[2 * i]), "f"(__uint_as_float(h[i]))); }
                if (KIND == 1) { asm volatile("mul.rn.bf16x2 %0, %0, %1;" : "+r"(h[i]) : "r"(0x3e4d3e4du)); }
                if (KIND == 2) { asm volatile("max.bf16x2 %0, %0, %1;" : "+r"(h[i]) : "r"(h[(i + 1) & 7])); }
                if (KIND == 3) { asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(v[i]) : "l"(t2)); }
                if (KIND == 4) {  // the epilogue's mix: one of each per column pair
                    asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(v[i]) : "l"(t2));
                    asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(f[2 * i]), "f"(__uint_as_float(h[i])));
                    uint32_t t;
                    asm volatile("mul.rn.bf16x2 %0, %1, %2;" : "=r"(t) : "r"(h[i]), "r"(0x3e4d3e4du));
                    asm volatile("max.bf16x2 %0, %0, %1;" : "+r"(h[i]) : "r"(t));
                }
                if (KIND == 5) {  // FADD2 + F2FP only
                    asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(v[i]) : "l"(t2));
                    asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(f[2 * i]), "f"(__uint_as_float(h[i])));
                }
                if (KIND == 6) {  // HMUL2 + HMNMX2 only
                    uint32_t t;
                    asm volatile("mul.rn.bf16x2 %0, %1, %2;" : "=r"(t) : "r"(h[i]), "r"(0x3e4d3e4du));
                    asm volatile("max.bf16x2 %0, %0, %1;" : "+r"(h[i]) : "r"(t));
                }
                if (KIND == 7) {  // F2FP + HMNMX2
                    asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(f[2 * i]), "f"(__uint_as_float(h[i])));
                    asm volatile("max.bf16x2 %0, %0, %1;" : "+r"(h[(i + 3) & 7]) : "r"(0x3e4d3e4du));
                }
            }
        }
    }
    const long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r ^= h[i] ^ (uint32_t)v[i] ^ __float_as_uint(f[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 148 * 8);
    const int iters = 2000;
    const char* names[8] = {"F2FP cvt.rn.bf16x2.f32", "HMUL2.BF16", "HMNMX2.BF16", "FADD2", "mix of the four", "FADD2 + F2FP", "HMUL2 + HMNMX2", "F2FP + HMNMX2"};
    const int per[8] = {1, 1, 1, 1, 4, 2, 2, 2};
    for (int k = 0; k < 8; ++k) {
        for (int rep = 0; rep < 2; ++rep) {
            if (k == 0) probe<0><<<148, 512>>>(iters, out, cyc);
            if (k == 1) probe<1><<<148, 512>>>(iters, out, cyc);
            if (k == 2) probe<2><<<148, 512>>>(iters, out, cyc);
            if (k == 3) probe<3><<<148, 512>>>(iters, out, cyc);
            if (k == 4) probe<4><<<148, 512>>>(iters, out, cyc);
            if (k == 5) probe<5><<<148, 512>>>(iters, out, cyc);
            if (k == 6) probe<6><<<148, 512>>>(iters, out, cyc);
            if (k == 7) probe<7><<<148, 512>>>(iters, out, cyc);
            cudaDeviceSynchronize();
        }
        long long h[148];
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        const double inst = (double)iters * 64 * 16 * per[k];  // warp instructions per SM (16 warps)
        printf("%-24s %.2f warp-instructions / clk / SM  (%.1f threads / clk)\n", names[k], inst / h[0], 32 * inst / h[0]);
    }
    return 0;
}
