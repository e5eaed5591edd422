// Training-step kernels that are not GEMMs: train-mode BatchNorm1d (batch statistics, forward and backward), LeakyReLU
// backward, column sums (bias gradients) and a fused multi-tensor Adam.
// Reference: nn.BatchNorm1d / nn.LeakyReLU inside ResBlock (DexGanGrasp/models/networks.py:26-56) as exercised by
// DexGanGrasp.update_dexgangrasp / update_dexevaluator (models/dexgangrasp.py:739-823); torch.optim.Adam at :97-112.
// All of these are HBM-bound streaming kernels over [rows, features] fp32 activations.
#include "common.cuh"

namespace dexgg {

constexpr int CR_TX = 32, CR_TY = 8;  // column-reduce CTA: 32 columns x 8 row lanes

// mode 0: out0[n] += sum_m x[m,n]
// mode 1: out0[n] += sum_m (x[m,n] - mean[n])^2
// mode 2: out0[n] += sum_m g[m,n];  out1[n] += sum_m g[m,n] * (x[m,n] - mean[n]) * invstd[n]      (x = second input)
template <int MODE>
__global__ void __launch_bounds__(CR_TX * CR_TY)
col_reduce_kernel(const float* __restrict__ a, int64_t lda, const float* __restrict__ x, int64_t ldx, const float* __restrict__ mean,
                  const float* __restrict__ invstd, int64_t M, int N, int rows_per_cta, float* __restrict__ out0,
                  float* __restrict__ out1) {
    __shared__ float s0[CR_TY][CR_TX + 1], s1[CR_TY][CR_TX + 1];
    const int tx = threadIdx.x % CR_TX, ty = threadIdx.x / CR_TX;
    const int n = blockIdx.x * CR_TX + tx;
    const int64_t m_lo = (int64_t)blockIdx.y * rows_per_cta;
    const int64_t m_hi = (m_lo + rows_per_cta < M) ? m_lo + rows_per_cta : M;
    float acc0 = 0.f, acc1 = 0.f;
    if (n < N) {
        const float mu = (MODE >= 1) ? mean[n] : 0.f;
        const float is = (MODE == 2) ? invstd[n] : 0.f;
        for (int64_t m = m_lo + ty; m < m_hi; m += CR_TY) {
            const float v = a[m * lda + n];
            if (MODE == 0) acc0 += v;
            if (MODE == 1) acc0 = fmaf(v - mu, v - mu, acc0);
            if (MODE == 2) {
                acc0 += v;
                acc1 = fmaf(v, (x[m * ldx + n] - mu) * is, acc1);
            }
        }
    }
    s0[ty][tx] = acc0;
    s1[ty][tx] = acc1;
    __syncthreads();
    if (ty == 0 && n < N) {
#pragma unroll
        for (int j = 1; j < CR_TY; ++j) {
            acc0 += s0[j][tx];
            acc1 += s1[j][tx];
        }
        atomicAdd(out0 + n, acc0);
        if (MODE == 2) atomicAdd(out1 + n, acc1);
    }
}

// mean = sum / count; (second pass) var = sumsq / count (biased); invstd = rsqrt(var + eps);
// running_mean = (1-mom)*running_mean + mom*mean; running_var = (1-mom)*running_var + mom*var*count/(count-1)
__global__ void bn_finalize_mean_kernel(const float* sum, float count, int N, float* mean) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < N) mean[n] = sum[n] / count;
}
__global__ void bn_finalize_var_kernel(const float* sumsq, const float* mean, float count, int N, float eps, float momentum,
                                       float* invstd, float* running_mean, float* running_var) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float var = sumsq[n] / count;
    invstd[n] = 1.0f / sqrtf(var + eps);
    if (running_mean) {
        running_mean[n] = (1.f - momentum) * running_mean[n] + momentum * mean[n];
        const float unbiased = count > 1.f ? var * (count / (count - 1.f)) : var;
        running_var[n] = (1.f - momentum) * running_var[n] + momentum * unbiased;
    }
}

// y = act(gamma * (x - mean) * invstd + beta + residual)
__global__ void bn_apply_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ mean,
                                const float* __restrict__ invstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                                const float* __restrict__ residual, int64_t ldr, int act, int64_t M, int N, float* __restrict__ y,
                                int64_t ldy) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M * N) return;
    const int64_t m = i / N;
    const int n = (int)(i % N);
    float v = fmaf((x[m * ldx + n] - mean[n]) * invstd[n], gamma[n], beta[n]);
    if (residual) v += residual[m * ldr + n];
    if (act) v = v > 0.f ? v : 0.2f * v;
    y[m * ldy + n] = v;
}

// g_out = g * (y > 0 ? 1 : 0.2)       (LeakyReLU backward from the saved OUTPUT; sign(y) == sign(pre-activation))
__global__ void lrelu_backward_kernel(const float* __restrict__ g, int64_t ldg, const float* __restrict__ y, int64_t ldy, int64_t M,
                                      int N, float* __restrict__ out, int64_t ldo) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M * N) return;
    const int64_t m = i / N;
    const int n = (int)(i % N);
    const float v = g[m * ldg + n];
    out[m * ldo + n] = y[m * ldy + n] > 0.f ? v : 0.2f * v;
}

// gx = gamma * invstd * (g - sum_g / count - xhat * sum_gx / count)
__global__ void bn_backward_apply_kernel(const float* __restrict__ g, int64_t ldg, const float* __restrict__ x, int64_t ldx,
                                         const float* __restrict__ mean, const float* __restrict__ invstd,
                                         const float* __restrict__ gamma, const float* __restrict__ sum_g,
                                         const float* __restrict__ sum_gx, float count, int64_t M, int N, float* __restrict__ gx,
                                         int64_t ldo) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M * N) return;
    const int64_t m = i / N;
    const int n = (int)(i % N);
    const float xhat = (x[m * ldx + n] - mean[n]) * invstd[n];
    gx[m * ldo + n] = gamma[n] * invstd[n] * (g[m * ldg + n] - sum_g[n] / count - xhat * sum_gx[n] / count);
}

// torch.optim.Adam (coupled L2 weight decay, bias correction, eps outside the square root), many tensors in one launch.
struct AdamChunk {
    float *w, *g, *m, *v;
    int n;
};
__global__ void adam_multi_kernel(const AdamChunk* __restrict__ chunks, float lr, float beta1, float beta2, float eps, float wd,
                                  float bc1, float bc2_sqrt) {
    const AdamChunk c = chunks[blockIdx.x];
    for (int i = threadIdx.x; i < c.n; i += blockDim.x) {
        const float w = c.w[i];
        const float g = fmaf(wd, w, c.g[i]);
        const float m = beta1 * c.m[i] + (1.f - beta1) * g;
        const float v = beta2 * c.v[i] + (1.f - beta2) * g * g;
        c.m[i] = m;
        c.v[i] = v;
        const float denom = sqrtf(v) / bc2_sqrt + eps;
        c.w[i] = w - (lr / bc1) * (m / denom);
    }
}

// Variant for a step captured in a CUDA graph: the step count lives in device memory (a kernel argument would be frozen at
// capture time).  adam_step_inc_kernel bumps it, then every CTA derives the bias corrections from it in double precision.
__global__ void adam_step_inc_kernel(int* step) { *step += 1; }

// one element of torch.optim.Adam (coupled L2 weight decay, bias correction, eps outside the square root)
__device__ __forceinline__ void adam_one(float& w, float g, float& m, float& v, float lr_bc1, float beta1, float beta2, float eps, float wd,
                                         float bc2_sqrt) {
    g = fmaf(wd, w, g);
    m = beta1 * m + (1.f - beta1) * g;
    v = beta2 * v + (1.f - beta2) * g * g;
    w = w - lr_bc1 * (m / (sqrtf(v) / bc2_sqrt + eps));
}

__global__ void __launch_bounds__(256) adam_multi_dev_kernel(const AdamChunk* __restrict__ chunks, float lr, float beta1, float beta2, float eps,
                                                             float wd, const int* __restrict__ step) {
    __shared__ float s_bc[2];
    if (threadIdx.x == 0) {
        const int t = *step;
        s_bc[0] = (float)(1.0 - pow((double)beta1, (double)t));
        s_bc[1] = (float)sqrt(1.0 - pow((double)beta2, (double)t));
    }
    __syncthreads();
    const float lr_bc1 = lr / s_bc[0], bc2_sqrt = s_bc[1];
    const AdamChunk c = chunks[blockIdx.x];
    const bool vec = ((reinterpret_cast<uintptr_t>(c.w) | reinterpret_cast<uintptr_t>(c.g) | reinterpret_cast<uintptr_t>(c.m) |
                       reinterpret_cast<uintptr_t>(c.v)) & 15) == 0;
    const int n4 = vec ? c.n / 4 : 0;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {  // 16-byte accesses: 28 B of HBM traffic per parameter is all this kernel is
        float4 w = reinterpret_cast<float4*>(c.w)[i], m = reinterpret_cast<float4*>(c.m)[i], v = reinterpret_cast<float4*>(c.v)[i];
        const float4 g = reinterpret_cast<const float4*>(c.g)[i];
        adam_one(w.x, g.x, m.x, v.x, lr_bc1, beta1, beta2, eps, wd, bc2_sqrt);
        adam_one(w.y, g.y, m.y, v.y, lr_bc1, beta1, beta2, eps, wd, bc2_sqrt);
        adam_one(w.z, g.z, m.z, v.z, lr_bc1, beta1, beta2, eps, wd, bc2_sqrt);
        adam_one(w.w, g.w, m.w, v.w, lr_bc1, beta1, beta2, eps, wd, bc2_sqrt);
        reinterpret_cast<float4*>(c.w)[i] = w;
        reinterpret_cast<float4*>(c.m)[i] = m;
        reinterpret_cast<float4*>(c.v)[i] = v;
    }
    for (int i = 4 * n4 + threadIdx.x; i < c.n; i += blockDim.x) {
        float w = c.w[i], m = c.m[i], v = c.v[i];
        adam_one(w, c.g[i], m, v, lr_bc1, beta1, beta2, eps, wd, bc2_sqrt);
        c.w[i] = w;
        c.m[i] = m;
        c.v[i] = v;
    }
}

}  // namespace dexgg

using namespace dexgg;

static inline unsigned nb(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }

static void reduce_grid(int64_t M, int N, dim3& grid, int& rows_per_cta) {
    const int gx = (N + CR_TX - 1) / CR_TX;
    int64_t want = (4ll * sm_count() + gx - 1) / gx;  // ~4 CTAs per SM in total
    if (want < 1) want = 1;
    int64_t rpc = (M + want - 1) / want;
    if (rpc < 64) rpc = 64;
    rows_per_cta = (int)rpc;
    grid = dim3(gx, (unsigned)((M + rpc - 1) / rpc));
}

// sums[N] += column sums of x[M,N]   (bias gradient; also pass 1 of the BatchNorm statistics).  `sums` must be zeroed by the caller.
extern "C" int dexgg_col_sum(const float* x, int64_t ldx, int64_t M, int N, float* sums, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && N >= 1 && sums && (M == 0 || x), "col_sum: bad arguments");
    if (M == 0) return DEXGG_OK;
    dim3 grid;
    int rpc;
    reduce_grid(M, N, grid, rpc);
    col_reduce_kernel<0><<<grid, CR_TX * CR_TY, 0, static_cast<cudaStream_t>(stream)>>>(x, ldx, nullptr, 0, nullptr, nullptr, M, N, rpc, sums, nullptr); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("col_reduce_kernel<0>");
    return DEXGG_OK;
}

// sumsq[N] += sum_m (x - mean)^2   (pass 2 of the BatchNorm statistics)
extern "C" int dexgg_col_centered_sumsq(const float* x, int64_t ldx, const float* mean, int64_t M, int N, float* sumsq, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && N >= 1 && sumsq && mean && (M == 0 || x), "col_centered_sumsq: bad arguments");
    if (M == 0) return DEXGG_OK;
    dim3 grid;
    int rpc;
    reduce_grid(M, N, grid, rpc);
    col_reduce_kernel<1><<<grid, CR_TX * CR_TY, 0, static_cast<cudaStream_t>(stream)>>>(x, ldx, nullptr, 0, mean, nullptr, M, N, rpc, sumsq, nullptr); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("col_reduce_kernel<1>");
    return DEXGG_OK;
}

extern "C" int dexgg_bn_finalize(const float* sum_or_sumsq, int pass, float count, int N, float eps, float momentum, float* mean,
                                 float* invstd, float* running_mean, float* running_var, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(N >= 1 && count >= 1.f && sum_or_sumsq && mean && (pass == 0 || invstd), "bn_finalize: bad arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (pass == 0) {
        bn_finalize_mean_kernel<<<nb(N), 256, 0, st>>>(sum_or_sumsq, count, N, mean); ::dexgg::count_launches(1);
    } else {
        bn_finalize_var_kernel<<<nb(N), 256, 0, st>>>(sum_or_sumsq, mean, count, N, eps, momentum, invstd, running_mean, running_var); ::dexgg::count_launches(1);
    }
    DEXGG_LAUNCH_OK("bn_finalize kernels");
    return DEXGG_OK;
}

extern "C" int dexgg_bn_apply(const float* x, int64_t ldx, const float* mean, const float* invstd, const float* gamma,
                              const float* beta, const float* residual, int64_t ldr, int act, int64_t M, int N, float* y,
                              int64_t ldy, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && N >= 1 && mean && invstd && gamma && beta && (M == 0 || (x && y)), "bn_apply: bad arguments");
    if (M == 0) return DEXGG_OK;
    bn_apply_kernel<<<nb(M * N), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, ldx, mean, invstd, gamma, beta, residual, ldr, act, M, N, y, ldy); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("bn_apply_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_lrelu_backward(const float* g, int64_t ldg, const float* y, int64_t ldy, int64_t M, int N, float* out,
                                    int64_t ldo, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && N >= 1 && (M == 0 || (g && y && out)), "lrelu_backward: bad arguments");
    if (M == 0) return DEXGG_OK;
    lrelu_backward_kernel<<<nb(M * N), 256, 0, static_cast<cudaStream_t>(stream)>>>(g, ldg, y, ldy, M, N, out, ldo); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("lrelu_backward_kernel");
    return DEXGG_OK;
}

// sum_g[N] += sum_m g;  sum_gx[N] += sum_m g * xhat    (both zeroed by the caller; they are also the beta / gamma gradients)
extern "C" int dexgg_bn_backward_reduce(const float* g, int64_t ldg, const float* x, int64_t ldx, const float* mean,
                                        const float* invstd, int64_t M, int N, float* sum_g, float* sum_gx, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && N >= 1 && mean && invstd && sum_g && sum_gx && (M == 0 || (g && x)), "bn_backward_reduce: bad arguments");
    if (M == 0) return DEXGG_OK;
    dim3 grid;
    int rpc;
    reduce_grid(M, N, grid, rpc);
    col_reduce_kernel<2><<<grid, CR_TX * CR_TY, 0, static_cast<cudaStream_t>(stream)>>>(g, ldg, x, ldx, mean, invstd, M, N, rpc, sum_g, sum_gx); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("col_reduce_kernel<2>");
    return DEXGG_OK;
}

extern "C" int dexgg_bn_backward_apply(const float* g, int64_t ldg, const float* x, int64_t ldx, const float* mean,
                                       const float* invstd, const float* gamma, const float* sum_g, const float* sum_gx,
                                       float count, int64_t M, int N, float* gx, int64_t ldo, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && N >= 1 && count >= 1.f && mean && invstd && gamma && sum_g && sum_gx && (M == 0 || (g && x && gx)),
                  "bn_backward_apply: bad arguments");
    if (M == 0) return DEXGG_OK;
    bn_backward_apply_kernel<<<nb(M * N), 256, 0, static_cast<cudaStream_t>(stream)>>>(g, ldg, x, ldx, mean, invstd, gamma, sum_g, sum_gx, count, M, N, gx, ldo); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("bn_backward_apply_kernel");
    return DEXGG_OK;
}

// Graph-capturable form of dexgg_adam_step: `step_dev` (device int, the number of steps taken so far) is incremented first and
// the bias corrections are computed from it on the device.
extern "C" int dexgg_adam_step_dev(const void* chunks_dev, int n_chunks, float lr, float beta1, float beta2, float eps,
                                   float weight_decay, int* step_dev, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(n_chunks >= 0 && step_dev && (n_chunks == 0 || chunks_dev), "adam_step_dev: bad arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    adam_step_inc_kernel<<<1, 1, 0, st>>>(step_dev); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("adam_step_inc_kernel");
    if (n_chunks == 0) return DEXGG_OK;
    adam_multi_dev_kernel<<<n_chunks, 256, 0, st>>>(static_cast<const AdamChunk*>(chunks_dev), lr, beta1, beta2, eps, weight_decay, step_dev); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("adam_multi_dev_kernel");
    return DEXGG_OK;
}

// One Adam step over `n_chunks` (tensor slices of at most 65536 elements) described by a DEVICE array of
// {w, g, m, v, n} records (5 x 8 bytes each: four pointers and an int padded to 8 bytes).  step >= 1.
extern "C" int dexgg_adam_step(const void* chunks_dev, int n_chunks, float lr, float beta1, float beta2, float eps,
                               float weight_decay, int step, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(n_chunks >= 0 && step >= 1 && (n_chunks == 0 || chunks_dev), "adam_step: bad arguments");
    if (n_chunks == 0) return DEXGG_OK;
    static_assert(sizeof(AdamChunk) == 40, "AdamChunk layout is part of the ABI");
    const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
    adam_multi_kernel<<<n_chunks, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const AdamChunk*>(chunks_dev), lr, beta1, beta2, eps, weight_decay, (float)bc1, (float)sqrt(bc2)); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("adam_multi_kernel");
    return DEXGG_OK;
}
