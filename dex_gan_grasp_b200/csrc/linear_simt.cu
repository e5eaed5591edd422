// fp32 CUDA-core fused linear layer (the "fp32" precision mode: outputs within 1e-4 of the reference's torch fp32 path).
// Classic register-tiled SGEMM: CTA tile 128 x BN x 16, 256 threads, 8 x (BN/16) outputs per thread, register-prefetch
// double buffering, epilogue of linear.cuh.  Supports transposed operands so the same kernel serves the backward pass
// (dX = dY·W, dW = dYᵀ·X).  Roofline: FP32 pipe.
#include "linear.cuh"

namespace dexgg {

constexpr int SG_BM = 128, SG_BK = 16, SG_THREADS = 256;

template <int BN, bool TA, bool TW>
__global__ void __launch_bounds__(SG_THREADS, 2) linear_simt_kernel(const LinearF32 p, const int vec_a, const int vec_w,
                                                                   const int vec_c) {
    constexpr int EW = BN / 16;  // W elements per thread per k-tile, and output columns per thread
    __shared__ __align__(16) float As[2][SG_BK][SG_BM];
    __shared__ __align__(16) float Ws[2][SG_BK][BN];
    const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
    const int m0 = blockIdx.x * SG_BM, n0 = blockIdx.y * BN;
    const int M = p.M, N = p.N, K = p.K;

    float ra[8], rw[EW];
    auto load_a = [&](int kt) {
        if (!TA) {
            const int row = t & 127, k0 = (t >> 7) * 8, m = m0 + row;
            const float* src = p.A + (int64_t)m * p.lda + kt + k0;
            if (m < M && vec_a && kt + k0 + 8 <= K) {
                const float4 u = *reinterpret_cast<const float4*>(src), v = *reinterpret_cast<const float4*>(src + 4);
                ra[0] = u.x; ra[1] = u.y; ra[2] = u.z; ra[3] = u.w; ra[4] = v.x; ra[5] = v.y; ra[6] = v.z; ra[7] = v.w;
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) ra[i] = (m < M && kt + k0 + i < K) ? src[i] : 0.f;
            }
        } else {
            const int k = kt + (t >> 4);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int m = m0 + (t & 15) + 16 * i;
                ra[i] = (k < K && m < M) ? p.A[(int64_t)k * p.lda + m] : 0.f;
            }
        }
    };
    auto store_a = [&](int buf) {
        if (!TA) {
            const int row = t & 127, k0 = (t >> 7) * 8;
#pragma unroll
            for (int i = 0; i < 8; ++i) As[buf][k0 + i][row] = ra[i];
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) As[buf][t >> 4][(t & 15) + 16 * i] = ra[i];
        }
    };
    auto load_w = [&](int kt) {
        if (!TW) {
            const int row = t % BN, k0 = (t / BN) * EW, n = n0 + row;
            const float* src = p.W + (int64_t)n * p.ldw + kt + k0;
            bool done = false;
            if constexpr (EW == 8) {
                if (n < N && vec_w && kt + k0 + 8 <= K) {
                    const float4 u = *reinterpret_cast<const float4*>(src), v = *reinterpret_cast<const float4*>(src + 4);
                    rw[0] = u.x; rw[1] = u.y; rw[2] = u.z; rw[3] = u.w; rw[4] = v.x; rw[5] = v.y; rw[6] = v.z; rw[7] = v.w;
                    done = true;
                }
            }
            if (!done) {
#pragma unroll
                for (int i = 0; i < EW; ++i) rw[i] = (n < N && kt + k0 + i < K) ? src[i] : 0.f;
            }
        } else {
            const int k = kt + (t >> 4);
#pragma unroll
            for (int i = 0; i < EW; ++i) {
                const int n = n0 + (t & 15) + 16 * i;
                rw[i] = (k < K && n < N) ? p.W[(int64_t)k * p.ldw + n] : 0.f;
            }
        }
    };
    auto store_w = [&](int buf) {
        if (!TW) {
            const int row = t % BN, k0 = (t / BN) * EW;
#pragma unroll
            for (int i = 0; i < EW; ++i) Ws[buf][k0 + i][row] = rw[i];
        } else {
#pragma unroll
            for (int i = 0; i < EW; ++i) Ws[buf][t >> 4][(t & 15) + 16 * i] = rw[i];
        }
    };

    float acc[8][EW];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < EW; ++j) acc[i][j] = 0.f;

    const int nk = (K + SG_BK - 1) / SG_BK;
    load_a(0);
    load_w(0);
    store_a(0);
    store_w(0);
    __syncthreads();
    for (int kt = 0; kt < nk; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < nk) {
            load_a((kt + 1) * SG_BK);
            load_w((kt + 1) * SG_BK);
        }
#pragma unroll
        for (int k = 0; k < SG_BK; ++k) {
            float a[8], w[EW];
            const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][64 + ty * 4]);
            a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
            if constexpr (EW == 8) {
                const float4 w0 = *reinterpret_cast<const float4*>(&Ws[buf][k][tx * 4]);
                const float4 w1 = *reinterpret_cast<const float4*>(&Ws[buf][k][(BN / 2) + tx * 4]);
                w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w; w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
            } else {
                const float2 w0 = *reinterpret_cast<const float2*>(&Ws[buf][k][tx * 2]);
                w[0] = w0.x; w[1] = w0.y;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < EW; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
        }
        if (kt + 1 < nk) {
            store_a(buf ^ 1);
            store_w(buf ^ 1);
        }
        __syncthreads();
    }

    // ---- epilogue ----
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (m >= M) continue;
        if constexpr (EW == 8) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int n = n0 + h * (BN / 2) + tx * 4;
                float v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] = (n + j < N) ? apply_epilogue(p.ep, acc[i][h * 4 + j], m, n + j) : 0.f;
                float* dst = p.C + (int64_t)m * p.ldc + n;
                if (vec_c && n + 3 < N) {
                    if (p.ep.accumulate) {
                        const float4 o = *reinterpret_cast<const float4*>(dst);
                        v[0] += o.x; v[1] += o.y; v[2] += o.z; v[3] += o.w;
                    }
                    *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (n + j < N) dst[j] = p.ep.accumulate ? dst[j] + v[j] : v[j];
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < EW; ++j) {
                const int n = n0 + tx * 2 + j;
                if (n < N) {
                    const float v = apply_epilogue(p.ep, acc[i][j], m, n);
                    float* dst = p.C + (int64_t)m * p.ldc + n;
                    *dst = p.ep.accumulate ? *dst + v : v;
                }
            }
        }
    }
}

template <int BN>
static int launch_simt(const LinearF32& p, cudaStream_t st) {
    const int va = (!p.trans_a && p.lda % 4 == 0 && (reinterpret_cast<uintptr_t>(p.A) & 15) == 0) ? 1 : 0;
    const int vw = (!p.trans_w && p.ldw % 4 == 0 && (reinterpret_cast<uintptr_t>(p.W) & 15) == 0) ? 1 : 0;
    const int vc = (p.ldc % 4 == 0 && (reinterpret_cast<uintptr_t>(p.C) & 15) == 0) ? 1 : 0;
    dim3 grid((p.M + SG_BM - 1) / SG_BM, (p.N + BN - 1) / BN);
    if (!p.trans_a && !p.trans_w) { linear_simt_kernel<BN, false, false><<<grid, SG_THREADS, 0, st>>>(p, va, vw, vc); ::dexgg::count_launches(1); }
    else if (!p.trans_a && p.trans_w) { linear_simt_kernel<BN, false, true><<<grid, SG_THREADS, 0, st>>>(p, va, vw, vc); ::dexgg::count_launches(1); }
    else if (p.trans_a && !p.trans_w) { linear_simt_kernel<BN, true, false><<<grid, SG_THREADS, 0, st>>>(p, va, vw, vc); ::dexgg::count_launches(1); }
    else { linear_simt_kernel<BN, true, true><<<grid, SG_THREADS, 0, st>>>(p, va, vw, vc); ::dexgg::count_launches(1); }
    DEXGG_LAUNCH_OK("linear_simt_kernel");
    return DEXGG_OK;
}

int linear_f32(const LinearF32& p, cudaStream_t st) {
    DEXGG_REQUIRE(p.M >= 0 && p.N >= 1 && p.K >= 1 && p.A && p.W && p.C, "linear_f32: bad arguments (M=%d N=%d K=%d)", p.M, p.N, p.K);
    DEXGG_REQUIRE(!p.ep.base || p.ep.rows_per_obj >= 1, "linear_f32: rows_per_obj must be >= 1");
    if (p.M == 0) return DEXGG_OK;
    return p.N <= 32 ? launch_simt<32>(p, st) : launch_simt<128>(p, st);
}

}  // namespace dexgg

// C ABI: a single fused linear layer (building block of the training step; also used by the tests to check the kernel
// in isolation).  Replaces nn.Linear (+ folded BatchNorm1d + LeakyReLU + residual) of networks.py:26-56.
extern "C" int dexgg_linear_f32(const float* A, int64_t lda, int trans_a, const float* W, int64_t ldw, int trans_w, int M,
                                int N, int K, const float* bias, const float* scale, const float* shift,
                                const float* residual, int64_t ldr, int act, int accumulate, float* C, int64_t ldc,
                                void* stream) {
    DEXGG_ARCH_OK();
    dexgg::LinearF32 p;
    p.A = A; p.lda = lda; p.trans_a = trans_a; p.W = W; p.ldw = ldw; p.trans_w = trans_w;
    p.M = M; p.N = N; p.K = K; p.C = C; p.ldc = ldc;
    p.ep.bias = bias; p.ep.scale = scale; p.ep.shift = shift; p.ep.residual = residual; p.ep.ldr = ldr;
    p.ep.act = act; p.ep.accumulate = accumulate;
    DEXGG_REQUIRE(!scale == !shift, "linear_f32: scale and shift come together");
    return dexgg::linear_f32(p, static_cast<cudaStream_t>(stream));
}
