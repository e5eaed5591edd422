// Fused linear layer  C = epilogue(A · Wᵀ)  shared by the fp32 CUDA-core kernel (linear_simt.cu) and the bf16
// tcgen05 kernel (linear_umma.cu).  The epilogue is what turns a GEMM into a layer of the reference's ResBlock
// (DexGanGrasp/models/networks.py:43-56):
//     v = acc
//     v += base[((row0 + m) / rows_per_obj) * ldbase + n]     per-object hoisted BPS projection (SURVEY §8a note H)
//     v += bias[n]
//     v  = v * scale[n] + shift[n]                    BatchNorm (eval) folded to an affine
//     v += residual[m * ldr + n]                      ResBlock skip branch
//     v  = act ? leaky_relu(v, 0.2) : v
//     C[m * ldc + n] = v (+ beta * C)
#pragma once
#include "common.cuh"

namespace dexgg {

struct Epilogue {
    const float* base = nullptr;
    int64_t ldbase = 0;
    int rows_per_obj = 1;
    int64_t row0 = 0;  // global row index of row 0 of this call (chunked calls): object = (row0 + m) / rows_per_obj
    const float* bias = nullptr;
    const float* scale = nullptr;
    const float* shift = nullptr;
    const float* residual = nullptr;
    int64_t ldr = 0;
    int act = 0;         // 1 = LeakyReLU(0.2)
    int accumulate = 0;  // 1: C += v (used by weight gradients)
};

struct LinearF32 {
    const float* A = nullptr;  // [M,K]: A[m*lda + k]   (trans_a: A[k*lda + m])
    int64_t lda = 0;
    const float* W = nullptr;  // [N,K]: W[n*ldw + k]   (trans_w: W[k*ldw + n])
    int64_t ldw = 0;
    int M = 0, N = 0, K = 0;
    int trans_a = 0, trans_w = 0;
    float* C = nullptr;
    int64_t ldc = 0;
    Epilogue ep;
};

int linear_f32(const LinearF32& p, cudaStream_t st);  // linear_simt.cu

struct LinearB16 {  // bf16 operands (K-major, row pitch a multiple of 16 bytes), fp32 accumulation on the tensor cores
    const void* A = nullptr;  // bf16 [M,K]
    int64_t lda = 0;
    const void* W = nullptr;  // bf16 [N,K]
    int64_t ldw = 0;
    int M = 0, N = 0, K = 0;
    Epilogue ep;                       // fp32 side inputs; ep.residual is ignored
    const void* residual16 = nullptr;  // bf16 [M, ldr]
    int64_t ldr = 0;
    void* C16 = nullptr;   // bf16 output [M, ldc] and / or
    float* C32 = nullptr;  // fp32 output [M, ldc]
    int64_t ldc = 0;
    // split-bf16 ("fp32 on tensor cores") plumbing: a value v is carried as hi = bf16(v), lo = bf16(v - hi)
    int split_seg = 0;               // > 0: C16 gets hi at column n, lo at n + split_seg, hi again at n + 2 * split_seg (the [hi|lo|hi]
                                     // operand layout that pairs with [w_hi|w_hi|w_lo] weights: hi*hi + lo*hi + hi*lo in one GEMM)
    int c32_from = 1 << 30;          // columns >= c32_from go to C32 ONLY, at column n - c32_from, row pitch ldc32
    int64_t ldc32 = 0;               // row pitch of C32 when split / c32_from are used (0: ldc)
    const float* residual32 = nullptr;  // fp32 residual [M, ldr32]
    int64_t ldr32 = 0;
};
int linear_bf16(const LinearB16& p, cudaStream_t st);  // linear_umma.cu

#define DEXGG_LRELU_SLOPE 0.2f

__device__ __forceinline__ float apply_epilogue(const Epilogue& ep, float v, int m, int n) {
    if (ep.base) v += ep.base[((ep.row0 + m) / ep.rows_per_obj) * ep.ldbase + n];
    if (ep.bias) v += ep.bias[n];
    if (ep.scale) v = fmaf(v, ep.scale[n], ep.shift[n]);
    if (ep.residual) v += ep.residual[(int64_t)m * ep.ldr + n];
    if (ep.act) v = v > 0.f ? v : DEXGG_LRELU_SLOPE * v;
    return v;
}

}  // namespace dexgg
