// Stages 2-3 — DexGenerator, Discriminator and DexEvaluator forwards (eval-mode BatchNorm) as chains of fused linear
// layers, with the per-object hoisting of the shared BPS projection (SURVEY.md §8a note H).
//
// Reference: DexGanGrasp/models/networks.py — ResBlock :26-56, Generator.forward :88-103, Discriminator.forward
// :148-159, DexEvaluator.forward :315-337; utils.rot_matrix_from_ortho6d (utils/utils.py:380-394) and
// utils.translate_along_axis (utils/utils.py:19-48) in the generator epilogue.
//
// A net handle owns PACKED weights derived from the caller's fp32 parameters:
//   Wb  [768*n_blocks, 4096]  the BPS columns of every block's fc1 (256 rows) and fc3 (512 rows), input BatchNorm folded
//   bb  [768*n_blocks]        their biases plus the folded BatchNorm shift
//   per block: Wt [768, KT]   the remaining columns in the order [previous X (512) | tail | zero pad], KT = 512+tailw or tailw
//              sc/sh [768]    bn1 affine for the fc1 rows, identity for the fc3 rows
//              W2 [512,256], sc2/sh2 [512]   fc2 with bn2 (and fc2's bias) folded into the affine
//   heads Wh [nh,512], bh [nh]
// "tail" = the per-row inputs that are not BPS: z (generator), bn1(rot9,transl3) (evaluator), bn1(rot9,transl3,joints)
// (discriminator), zero-padded to tailw = 16 or 32 columns.
#include <cuda_bf16.h>

#include "chain.cuh"
#include "linear.cuh"
#include <stdlib.h>

#include <new>

namespace dexgg {

constexpr int IN_BPS = 4096, HID = 512, INNER = 256, NCAT = INNER + HID;  // 768 = fc1 ‖ fc3
constexpr float BN_EPS = 1e-5f;
constexpr int64_t CHUNK_ROWS = 32768;

// ---------------------------------------------------------------- packing kernels
__global__ void bn_affine_kernel(const float* w, const float* b, const float* rm, const float* rv, int n, float* a, float* c) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const float s = w[i] / sqrtf(rv[i] + BN_EPS);
        a[i] = s;
        c[i] = b[i] - s * rm[i];
    }
}
// dst[r, dcol0 + j] = src[r, scol0 + j] * (colscale ? colscale[j] : 1)
__global__ void pack_cols_kernel(float* dst, int64_t ldd, int dcol0, const float* src, int64_t lds, int scol0, int ncols,
                                 int nrows, const float* colscale) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)ncols * nrows) return;
    const int r = (int)(i / ncols), j = (int)(i % ncols);
    dst[r * ldd + dcol0 + j] = src[r * lds + scol0 + j] * (colscale ? colscale[j] : 1.f);
}
// dst[r] = bias[r] + sum_j src[r, scol0 + j] * c[j]      (one warp per row)
__global__ void fold_bias_kernel(float* dst, const float* bias, const float* src, int64_t lds, int scol0, int ncols,
                                 int nrows, const float* c) {
    const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= nrows) return;
    float s = 0.f;
    if (c)
        for (int j = lane; j < ncols; j += 32) s = fmaf(src[r * lds + scol0 + j], c[j], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) dst[r] = bias[r] + s;
}
__global__ void fill_kernel(float* p, float v, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
// sc2 = a; sh2 = a * bias + c
__global__ void affine_bias_kernel(const float* a, const float* c, const float* bias, int n, float* sc, float* sh) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        sc[i] = a[i];
        sh[i] = fmaf(a[i], bias[i], c[i]);
    }
}

// ---------------------------------------------------------------- per-row helper kernels
// Concatenate up to three per-row sources, apply the input-BatchNorm affine, zero-pad to `width` and write the tail
// columns of up to three activation buffers.
struct TailDst {
    float* p[3];
    int64_t ld[3];
};
__global__ void pack_tail_kernel(const float* s0, int n0, const float* s1, int n1, const float* s2, int n2, const float* a,
                                 const float* c, int width, int64_t M, TailDst d) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M * width) return;
    const int64_t m = i / width;
    const int j = (int)(i % width);
    float v = 0.f;
    if (j < n0) v = s0[m * n0 + j];
    else if (j < n0 + n1) v = s1[m * n1 + (j - n0)];
    else if (j < n0 + n1 + n2) v = s2[m * n2 + (j - n0 - n1)];
    if (a && j < n0 + n1 + n2) v = fmaf(a[j], v, c[j]);
#pragma unroll
    for (int q = 0; q < 3; ++q)
        if (d.p[q]) d.p[q][m * d.ld[q] + j] = v;
}

// Generator epilogue: heads [M, ldh] = (joint_conf nJ | rot6d 6 | transl 3) -> rot_matrix (Gram-Schmidt, columns x,y,z),
// transl - z_offset * R[:,0] evaluated in float64 like utils.translate_along_axis, joint_conf.
__global__ void pose_finish_kernel(const float* heads, int64_t ldh, int nJ, int64_t M, float z_offset, float* rot,
                                   float* transl, float* joint, float* rot6d) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    const float* h = heads + m * ldh;
    for (int j = 0; j < nJ; ++j) joint[m * nJ + j] = h[j];
    const float a0 = h[nJ], a1 = h[nJ + 1], a2 = h[nJ + 2], b0 = h[nJ + 3], b1 = h[nJ + 4], b2 = h[nJ + 5];
    if (rot6d) {
        float* r6 = rot6d + m * 6;
        r6[0] = a0; r6[1] = a1; r6[2] = a2; r6[3] = b0; r6[4] = b1; r6[5] = b2;
    }
    const float na = fmaxf(sqrtf(a0 * a0 + a1 * a1 + a2 * a2), 1e-8f);
    const float x0 = a0 / na, x1 = a1 / na, x2 = a2 / na;
    float z0 = x1 * b2 - x2 * b1, z1 = x2 * b0 - x0 * b2, z2 = x0 * b1 - x1 * b0;
    const float nz = fmaxf(sqrtf(z0 * z0 + z1 * z1 + z2 * z2), 1e-8f);
    z0 /= nz; z1 /= nz; z2 /= nz;
    const float y0 = z1 * x2 - z2 * x1, y1 = z2 * x0 - z0 * x2, y2 = z0 * x1 - z1 * x0;
    float* r = rot + m * 9;
    r[0] = x0; r[1] = y0; r[2] = z0;
    r[3] = x1; r[4] = y1; r[5] = z1;
    r[6] = x2; r[7] = y2; r[8] = z2;
    const double off = (double)z_offset;
    transl[m * 3 + 0] = (float)((double)h[nJ + 6] - (double)x0 * off);
    transl[m * 3 + 1] = (float)((double)h[nJ + 7] - (double)x1 * off);
    transl[m * 3 + 2] = (float)((double)h[nJ + 8] - (double)x2 * off);
}

// p[m] = sigmoid(X[m,:] . w + b)     (one warp per row, K = 512)
__global__ void rowdot_sigmoid_kernel(const float* X, int64_t ldx, const float* w, const float* b, int K, int64_t M, float* p) {
    const int64_t m = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (m >= M) return;
    const float* x = X + m * ldx;
    float s = 0.f;
    for (int k = lane * 4; k < K; k += 128) {
        const float4 xv = *reinterpret_cast<const float4*>(x + k), wv = *reinterpret_cast<const float4*>(w + k);
        s = fmaf(xv.x, wv.x, s);
        s = fmaf(xv.y, wv.y, s);
        s = fmaf(xv.z, wv.z, s);
        s = fmaf(xv.w, wv.w, s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) p[m] = 1.f / (1.f + expf(-(s + b[0])));
}

// ---------------------------------------------------------------- bf16 tensor-core mode helpers
// A value x is carried as two bf16 numbers hi = bf16(x), lo = bf16(x - hi); with operands laid out along K as
// A' = [a_hi | a_lo | a_hi] and W' = [w_hi | w_hi | w_lo] one ordinary bf16 GEMM accumulates a_hi*w_hi + a_lo*w_hi + a_hi*w_lo,
// i.e. the product to ~2^-17 relative.  Used for the inputs that are NOT hidden activations (BPS encoding, z, pose).
__device__ __forceinline__ void split_bf16(float x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
    hi = __float2bfloat16_rn(x);
    lo = __float2bfloat16_rn(x - __bfloat162float(hi));
}
// dst[r, c_hi1 + j] = dst[r, c_hi2 + j] = hi(src[r, scol0 + j]);  dst[r, c_lo + j] = lo(...)
__global__ void pack_split_cols_kernel(__nv_bfloat16* dst, int64_t ldd, int c_hi1, int c_hi2, int c_lo, const float* src,
                                       int64_t lds, int scol0, int ncols, int nrows) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)ncols * nrows) return;
    const int r = (int)(i / ncols), j = (int)(i % ncols);
    __nv_bfloat16 hi, lo;
    split_bf16(src[r * lds + scol0 + j], hi, lo);
    dst[r * ldd + c_hi1 + j] = hi;
    dst[r * ldd + c_hi2 + j] = hi;
    dst[r * ldd + c_lo + j] = lo;
}
__global__ void pack_cols_bf16_kernel(__nv_bfloat16* dst, int64_t ldd, int dcol0, const float* src, int64_t lds, int scol0,
                                      int ncols, int nrows) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)ncols * nrows) return;
    const int r = (int)(i / ncols), j = (int)(i % ncols);
    dst[r * ldd + dcol0 + j] = __float2bfloat16_rn(src[r * lds + scol0 + j]);
}
// out[o, :] = [hi | lo | hi] of a[k]*bps[o,k] + c[k]   (input BatchNorm applied in fp32 before the split)
__global__ void bps_bn_split_kernel(const float* __restrict__ bps, const float* __restrict__ a, const float* __restrict__ c,
                                    __nv_bfloat16* __restrict__ out, int64_t n_obj) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obj * IN_BPS) return;
    const int64_t o = i / IN_BPS;
    const int k = (int)(i % IN_BPS);
    __nv_bfloat16 hi, lo;
    split_bf16(fmaf(a[k], bps[i], c[k]), hi, lo);
    __nv_bfloat16* row = out + o * (3 * IN_BPS);
    row[k] = hi;
    row[IN_BPS + k] = lo;
    row[2 * IN_BPS + k] = hi;
}
struct TailDst16 {
    __nv_bfloat16* p[3];
    int64_t ld[3];
};
// like pack_tail_kernel, output [hi (width) | lo (width) | hi (width)] in bf16
__global__ void pack_tail16_kernel(const float* s0, int n0, const float* s1, int n1, const float* s2, int n2, const float* a,
                                   const float* c, int width, int64_t M, TailDst16 d) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M * width) return;
    const int64_t m = i / width;
    const int j = (int)(i % width);
    float v = 0.f;
    if (j < n0) v = s0[m * n0 + j];
    else if (j < n0 + n1) v = s1[m * n1 + (j - n0)];
    else if (j < n0 + n1 + n2) v = s2[m * n2 + (j - n0 - n1)];
    if (a && j < n0 + n1 + n2) v = fmaf(a[j], v, c[j]);
    __nv_bfloat16 hi, lo;
    split_bf16(v, hi, lo);
#pragma unroll
    for (int q = 0; q < 3; ++q)
        if (d.p[q]) {
            __nv_bfloat16* row = d.p[q] + m * d.ld[q];
            row[j] = hi;
            row[width + j] = lo;
            row[2 * width + j] = hi;
        }
}
__global__ void rowdot_sigmoid_bf16_kernel(const __nv_bfloat16* X, int64_t ldx, const float* w, const float* b, int K, int64_t M,
                                           float* p) {
    const int64_t m = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (m >= M) return;
    const __nv_bfloat16* x = X + m * ldx;
    float s = 0.f;
    for (int k = lane * 8; k < K; k += 256) {
        const uint4 u = *reinterpret_cast<const uint4*>(x + k);
        const uint32_t q[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            s = fmaf(__uint_as_float(q[h] << 16), w[k + 2 * h], s);
            s = fmaf(__uint_as_float(q[h] & 0xffff0000u), w[k + 2 * h + 1], s);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) p[m] = 1.f / (1.f + expf(-(s + b[0])));
}

}  // namespace dexgg

using namespace dexgg;

// ---------------------------------------------------------------- net handle
struct BlockPack {
    float *Wt, *sc, *sh, *W2, *sc2, *sh2;
    int KT, has_x;
    // bf16 tensor-core mode: Wt16 [768, KT16] = [X (512) | tail hi | tail hi | tail lo], W2_16 [512,256]
    __nv_bfloat16 *Wt16, *W2_16;
    int KT16;
};
struct dexgg_net {
    int kind, precision, latentD, nJ;
    int n_blocks, n_tail, tailw, nbase, nh, n_params;
    float *Wb, *bb, *in_a, *in_c, *tail_a, *tail_c, *Wh, *bh, *tmp_a, *tmp_c;
    BlockPack blk[3];
    // bf16 mode: Wb16 [nbase, 3*4096] = [hi | hi | lo] of the RAW BPS columns (BatchNorm applied to the input instead),
    // bb_raw [nbase] raw biases, Wh16 [32,512] generator heads
    __nv_bfloat16 *Wb16, *Wh16;
    float* bb_raw;
    // bf16 mode, generator / evaluator: pre-swizzled weight tile stream of the fused whole-network kernel (chain_umma.cu)
    __nv_bfloat16* chain_stream;
    float *chain_sc, *chain_sh;  // [nbase] first-layer affine of every block, applied to the hoisted base by compute_base
    int chain_ok;
    void* arena;
    size_t arena_bytes;
};

namespace {
inline bool is_tc(const dexgg_net* n) { return n->precision != DEXGG_PRECISION_FP32; }        // bf16 tensor-core GEMMs
inline bool is_split(const dexgg_net* n) { return n->precision == DEXGG_PRECISION_FP32_TC; }  // hi + lo pairs everywhere
struct Layout {  // column ranges of one ResBlock's input
    int bps_off, tail_off, x_off;  // x_off < 0: no previous-X segment
};
int n_expected(int kind) { return kind == DEXGG_NET_GENERATOR ? 38 : kind == DEXGG_NET_DISCRIMINATOR ? 34 : 48; }

Layout block_layout(const dexgg_net* n, int b) {
    const int zD = n->latentD, np = n->n_tail;
    if (n->kind == DEXGG_NET_GENERATOR) return b == 0 ? Layout{zD, 0, -1} : Layout{zD, 0, zD + IN_BPS};  // cat[z,bps] / cat[X0,X]
    if (n->kind == DEXGG_NET_DISCRIMINATOR) return b == 0 ? Layout{0, IN_BPS, -1} : Layout{0, IN_BPS, IN_BPS + np};  // cat[X0,X]
    return b == 0 ? Layout{0, IN_BPS, -1} : Layout{HID, HID + IN_BPS, 0};  // evaluator: cat[X, X0]
}

inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }

int pack_net(dexgg_net* n, const float* const* P, cudaStream_t st) {
    // input BatchNorm -> affine over [bps | tail]
    const bool gen = n->kind == DEXGG_NET_GENERATOR;
    const int n_in = IN_BPS + (gen ? 0 : n->n_tail);
    bn_affine_kernel<<<nblk(n_in), 256, 0, st>>>(P[0], P[1], P[2], P[3], n_in, n->in_a, n->in_c); ::dexgg::count_launches(1);
    DEXGG_CUDA_OK(cudaMemsetAsync(n->tail_a, 0, 2 * n->tailw * sizeof(float), st));
    if (!gen) {
        DEXGG_CUDA_OK(cudaMemcpyAsync(n->tail_a, n->in_a + IN_BPS, n->n_tail * 4, cudaMemcpyDeviceToDevice, st));
        DEXGG_CUDA_OK(cudaMemcpyAsync(n->tail_c, n->in_c + IN_BPS, n->n_tail * 4, cudaMemcpyDeviceToDevice, st));
    }
    for (int b = 0; b < n->n_blocks; ++b) {
        const float* const* q = P + 4 + 14 * b;  // fc1.w fc1.b bn1.{w,b,rm,rv} fc2.w fc2.b bn2.{w,b,rm,rv} fc3.w fc3.b
        const Layout L = block_layout(n, b);
        const int fin = IN_BPS + n->n_tail + (L.x_off >= 0 ? HID : 0);
        BlockPack& B = n->blk[b];
        DEXGG_CUDA_OK(cudaMemsetAsync(B.Wt, 0, (size_t)NCAT * B.KT * 4, st));
        for (int part = 0; part < 2; ++part) {
            const float* w = part == 0 ? q[0] : q[12];
            const float* bias = part == 0 ? q[1] : q[13];
            const int rows = part == 0 ? INNER : HID, r0 = part == 0 ? 0 : INNER;
            float* wb = n->Wb + (size_t)(b * NCAT + r0) * IN_BPS;
            pack_cols_kernel<<<nblk((int64_t)rows * IN_BPS), 256, 0, st>>>(wb, IN_BPS, 0, w, fin, L.bps_off, IN_BPS, rows, n->in_a); ::dexgg::count_launches(1);
            fold_bias_kernel<<<nblk(rows, 8), 256, 0, st>>>(n->bb + b * NCAT + r0, bias, w, fin, L.bps_off, IN_BPS, rows, n->in_c); ::dexgg::count_launches(1);
            float* wt = B.Wt + (size_t)r0 * B.KT;
            int col = 0;
            if (L.x_off >= 0) {
                pack_cols_kernel<<<nblk((int64_t)rows * HID), 256, 0, st>>>(wt, B.KT, 0, w, fin, L.x_off, HID, rows, nullptr); ::dexgg::count_launches(1);
                col = HID;
            }
            pack_cols_kernel<<<nblk((int64_t)rows * n->n_tail), 256, 0, st>>>(wt, B.KT, col, w, fin, L.tail_off, n->n_tail, rows, nullptr); ::dexgg::count_launches(1);
        }
        bn_affine_kernel<<<nblk(INNER), 256, 0, st>>>(q[2], q[3], q[4], q[5], INNER, B.sc, B.sh); ::dexgg::count_launches(1);
        fill_kernel<<<nblk(HID), 256, 0, st>>>(B.sc + INNER, 1.f, HID); ::dexgg::count_launches(1);
        fill_kernel<<<nblk(HID), 256, 0, st>>>(B.sh + INNER, 0.f, HID); ::dexgg::count_launches(1);
        DEXGG_CUDA_OK(cudaMemcpyAsync(B.W2, q[6], (size_t)HID * INNER * 4, cudaMemcpyDeviceToDevice, st));
        bn_affine_kernel<<<nblk(HID), 256, 0, st>>>(q[8], q[9], q[10], q[11], HID, n->tmp_a, n->tmp_c); ::dexgg::count_launches(1);
        affine_bias_kernel<<<nblk(HID), 256, 0, st>>>(n->tmp_a, n->tmp_c, q[7], HID, B.sc2, B.sh2); ::dexgg::count_launches(1);
        if (is_tc(n)) {
            // bf16: [X (512) | t_hi | t_hi | t_lo];  split: [X_hi | X_hi | X_lo | t_hi | t_hi | t_lo] (pairs with [a_hi | a_lo | a_hi])
            const bool sp = is_split(n);
            const int tw = n->tailw, xw = B.has_x ? HID : 0, xo = sp ? 3 * xw : xw;
            DEXGG_CUDA_OK(cudaMemsetAsync(B.Wt16, 0, (size_t)NCAT * B.KT16 * 2, st));
            for (int part = 0; part < 2; ++part) {
                const float* w = part == 0 ? q[0] : q[12];
                const float* bias = part == 0 ? q[1] : q[13];
                const int rows = part == 0 ? INNER : HID, r0 = part == 0 ? 0 : INNER;
                __nv_bfloat16* wb = n->Wb16 + (size_t)(b * NCAT + r0) * (3 * IN_BPS);
                pack_split_cols_kernel<<<nblk((int64_t)rows * IN_BPS), 256, 0, st>>>(wb, 3 * IN_BPS, 0, IN_BPS, 2 * IN_BPS, w, fin, L.bps_off, IN_BPS, rows); ::dexgg::count_launches(1);
                DEXGG_CUDA_OK(cudaMemcpyAsync(n->bb_raw + b * NCAT + r0, bias, rows * 4, cudaMemcpyDeviceToDevice, st));
                __nv_bfloat16* wt = B.Wt16 + (size_t)r0 * B.KT16;
                if (B.has_x) {
                    if (sp) { pack_split_cols_kernel<<<nblk((int64_t)rows * HID), 256, 0, st>>>(wt, B.KT16, 0, HID, 2 * HID, w, fin, L.x_off, HID, rows); ::dexgg::count_launches(1); }
                    else { pack_cols_bf16_kernel<<<nblk((int64_t)rows * HID), 256, 0, st>>>(wt, B.KT16, 0, w, fin, L.x_off, HID, rows); ::dexgg::count_launches(1); }
                }
                pack_split_cols_kernel<<<nblk((int64_t)rows * n->n_tail), 256, 0, st>>>(wt, B.KT16, xo, xo + tw, xo + 2 * tw, w, fin, L.tail_off, n->n_tail, rows); ::dexgg::count_launches(1);
            }
            if (sp) { pack_split_cols_kernel<<<nblk((int64_t)HID * INNER), 256, 0, st>>>(B.W2_16, 3 * INNER, 0, INNER, 2 * INNER, q[6], INNER, 0, INNER, HID); ::dexgg::count_launches(1); }
            else { pack_cols_bf16_kernel<<<nblk((int64_t)HID * INNER), 256, 0, st>>>(B.W2_16, INNER, 0, q[6], INNER, 0, INNER, HID); ::dexgg::count_launches(1); }
        }
    }
    const float* const* h = P + 4 + 14 * n->n_blocks;
    if (gen) {  // rows: joint_conf (nJ) | rot6d (6) | transl (3)
        const int rows[3] = {n->nJ, 6, 3};
        int r = 0;
        for (int i = 0; i < 3; ++i) {
            DEXGG_CUDA_OK(cudaMemcpyAsync(n->Wh + (size_t)r * HID, h[2 * i], (size_t)rows[i] * HID * 4, cudaMemcpyDeviceToDevice, st));
            DEXGG_CUDA_OK(cudaMemcpyAsync(n->bh + r, h[2 * i + 1], rows[i] * 4, cudaMemcpyDeviceToDevice, st));
            r += rows[i];
        }
    } else {
        DEXGG_CUDA_OK(cudaMemcpyAsync(n->Wh, h[0], HID * 4, cudaMemcpyDeviceToDevice, st));
        DEXGG_CUDA_OK(cudaMemcpyAsync(n->bh, h[1], 4, cudaMemcpyDeviceToDevice, st));
    }
    if (gen && is_tc(n)) {
        const int hs = is_split(n) ? 3 : 1;
        DEXGG_CUDA_OK(cudaMemsetAsync(n->Wh16, 0, (size_t)32 * HID * hs * 2, st));
        if (is_split(n)) { pack_split_cols_kernel<<<nblk((int64_t)n->nh * HID), 256, 0, st>>>(n->Wh16, 3 * HID, 0, HID, 2 * HID, n->Wh, HID, 0, HID, n->nh); ::dexgg::count_launches(1); }
        else { pack_cols_bf16_kernel<<<nblk((int64_t)n->nh * HID), 256, 0, st>>>(n->Wh16, HID, 0, n->Wh, HID, 0, HID, n->nh); ::dexgg::count_launches(1); }
    }
    DEXGG_LAUNCH_OK("net packing kernels");
    if (n->chain_ok) {
        ChainPackSrc cs;
        cs.kind = n->kind; cs.n_blocks = n->n_blocks; cs.tailw = n->tailw; cs.n_tail = n->n_tail; cs.nh = n->nh; cs.Wh = n->Wh;
        for (int b = 0; b < CH_MAX_BLOCKS; ++b) {
            const bool on = b < n->n_blocks;
            cs.Wt[b] = on ? n->blk[b].Wt : nullptr; cs.KT[b] = on ? n->blk[b].KT : 0;
            cs.has_x[b] = on ? n->blk[b].has_x : 0; cs.W2[b] = on ? n->blk[b].W2 : nullptr;
            cs.sc[b] = on ? n->blk[b].sc : nullptr; cs.sc2[b] = on ? n->blk[b].sc2 : nullptr;
            if (on) {
                DEXGG_CUDA_OK(cudaMemcpyAsync(n->chain_sc + b * NCAT, n->blk[b].sc, NCAT * 4, cudaMemcpyDeviceToDevice, st));
                DEXGG_CUDA_OK(cudaMemcpyAsync(n->chain_sh + b * NCAT, n->blk[b].sh, NCAT * 4, cudaMemcpyDeviceToDevice, st));
            }
        }
        int rc = chain_pack(cs, n->chain_stream, st);
        if (rc) return rc;
    }
    return DEXGG_OK;
}

struct Carver {
    char* p;
    size_t used = 0;
    template <typename T>
    T* take(size_t n) {
        T* r = p ? reinterpret_cast<T*>(p + used) : nullptr;
        used += align_up(n * sizeof(T), 256);
        return r;
    }
};

void carve_net(dexgg_net* n, Carver& c) {
    n->Wb = c.take<float>((size_t)n->nbase * IN_BPS);
    n->bb = c.take<float>(n->nbase);
    n->in_a = c.take<float>(IN_BPS + 64);
    n->in_c = c.take<float>(IN_BPS + 64);
    n->tail_a = c.take<float>(2 * n->tailw);
    n->tail_c = n->tail_a ? n->tail_a + n->tailw : nullptr;
    n->tmp_a = c.take<float>(HID);
    n->tmp_c = c.take<float>(HID);
    n->Wh = c.take<float>((size_t)32 * HID);
    n->bh = c.take<float>(32);
    for (int b = 0; b < n->n_blocks; ++b) {
        BlockPack& B = n->blk[b];
        B.has_x = block_layout(n, b).x_off >= 0;
        B.KT = (B.has_x ? HID : 0) + n->tailw;
        B.Wt = c.take<float>((size_t)NCAT * B.KT);
        B.sc = c.take<float>(NCAT);
        B.sh = c.take<float>(NCAT);
        B.W2 = c.take<float>((size_t)HID * INNER);
        B.sc2 = c.take<float>(HID);
        B.sh2 = c.take<float>(HID);
        B.KT16 = (is_split(n) ? 3 : 1) * (B.has_x ? HID : 0) + 3 * n->tailw;
        B.Wt16 = nullptr;
        B.W2_16 = nullptr;
        if (is_tc(n)) {
            B.Wt16 = c.take<__nv_bfloat16>((size_t)NCAT * B.KT16);
            B.W2_16 = c.take<__nv_bfloat16>((size_t)HID * INNER * (is_split(n) ? 3 : 1));
        }
    }
    n->Wb16 = nullptr; n->Wh16 = nullptr; n->bb_raw = nullptr;
    if (is_tc(n)) {
        n->Wb16 = c.take<__nv_bfloat16>((size_t)n->nbase * 3 * IN_BPS);
        n->Wh16 = c.take<__nv_bfloat16>((size_t)32 * HID * (is_split(n) ? 3 : 1));
        n->bb_raw = c.take<float>(n->nbase);
    }
    n->chain_stream = nullptr;
    n->chain_sc = n->chain_sh = nullptr;
    n->chain_ok = 0;
    // the fused chain kernel: tails of at most 16 inputs (generator, evaluator) or the discriminator's pose + joint tail of at most 24
    if (n->precision == DEXGG_PRECISION_BF16 && (n->kind == DEXGG_NET_DISCRIMINATOR ? (n->tailw == 32 && n->n_tail <= 24) : n->tailw == 16)) {
        int hx[CH_MAX_BLOCKS] = {0, 0, 0};
        for (int b = 0; b < n->n_blocks; ++b) hx[b] = n->blk[b].has_x;
        n->chain_stream = c.take<__nv_bfloat16>(chain_stream_bytes(n->kind, n->n_blocks, hx) / 2);
        n->chain_sc = c.take<float>(n->nbase);
        n->chain_sh = c.take<float>(n->nbase);
        n->chain_ok = 1;
    }
}

// Debug builds only: DEXGG_NO_CHAIN=1 forces the layer-by-layer tensor-core path (A/B measurements, cross-checks)
bool chain_enabled() {
#ifdef DEXGG_DEBUG
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("DEXGG_NO_CHAIN");
        v = (e && e[0] == '1') ? 0 : 1;
    }
    return v == 1;
#else
    return true;
#endif
}

void chain_common(const dexgg_net* n, ChainArgs& a, const float* base, int64_t n_obj, int64_t rows_per_obj, int64_t M) {
    a.kind = n->kind; a.n_blocks = n->n_blocks; a.nJ = n->nJ; a.nh = n->nh;
    a.stream = n->chain_stream;
    for (int b = 0; b < CH_MAX_BLOCKS; ++b) {
        const bool on = b < n->n_blocks;
        a.has_x[b] = on ? n->blk[b].has_x : 0;
        a.sh2[b] = on ? n->blk[b].sh2 : nullptr;
    }
    a.base = base; a.n_obj = n_obj; a.rows_per_obj = rows_per_obj; a.M = M; a.bh = n->bh;
    a.z = nullptr; a.zD = n->latentD; a.z_offset = 0.f; a.rot = a.transl = a.joint = a.rot6d = nullptr;
    a.in_rot = a.in_transl = a.in_joint = nullptr; a.tail_a = n->tail_a; a.tail_c = n->tail_c; a.p = nullptr;
    a.debug_clock = nullptr; a.debug_no_stream = 0;
}

// workspace of one forward call.  Activations are fp32 (CUDA-core mode) or bf16 (tensor-core mode); `es` is the element
// size, `tw` the tail width in elements (tailw, or 3*tailw for the hi|lo|hi split), ldx = 512 + tw.
struct FwdWs {
    float *base, *H;
    char *tail, *T, *XT0, *XT1;
    __nv_bfloat16* Ab16;  // bf16 mode: [n_obj, 3*4096] split of bn(bps)
    float *R32, *X32;     // split mode: fp32 skip branch [mc, 512], fp32 copy of the last block's output [mc, 512]
    float* base_part;     // tensor-core modes: [base_split, n_obj, nbase] partial sums of the split base GEMM
    int64_t mc;
    int es, tw, ldx, xoff;  // xoff: element offset of the tail columns inside an XT row (512, or 3*512 in split mode)
};
int base_split(const dexgg_net* n, int64_t n_obj);
size_t carve_ws(const dexgg_net* n, int64_t M, int64_t n_obj, void* ws, FwdWs& w) {
    Carver c{static_cast<char*>(ws)};
    const bool b16 = is_tc(n), sp = is_split(n);
    // bf16: one chunk = 128 rows per SM, so every layer's tile count is a multiple of the SM count and the chunk's
    // activations (~60 MB) stay resident in the 126 MB L2 between layers
    const int64_t chunk = b16 ? (int64_t)sm_count() * 128 : CHUNK_ROWS;
    const int64_t mc = M < chunk ? M : chunk;
    w.mc = mc;
    w.es = b16 ? 2 : 4;
    w.tw = b16 ? 3 * n->tailw : n->tailw;
    w.xoff = sp ? 3 * HID : HID;
    w.ldx = w.xoff + w.tw;
    w.base = c.take<float>((size_t)n_obj * n->nbase);
    w.H = c.take<float>((size_t)mc * 32);
    w.tail = c.take<char>((size_t)mc * w.tw * w.es);
    w.T = c.take<char>((size_t)mc * NCAT * w.es);
    w.XT0 = c.take<char>((size_t)mc * w.ldx * w.es);
    w.XT1 = c.take<char>((size_t)mc * w.ldx * w.es);
    w.Ab16 = b16 ? c.take<__nv_bfloat16>((size_t)n_obj * 3 * IN_BPS) : nullptr;
    w.base_part = (b16 && base_split(n, n_obj) > 1) ? c.take<float>((size_t)base_split(n, n_obj) * n_obj * n->nbase) : nullptr;
    w.R32 = sp ? c.take<float>((size_t)mc * HID) : nullptr;
    w.X32 = sp ? c.take<float>((size_t)mc * HID) : nullptr;
    return c.used;
}

int check_fwd(const dexgg_net* n, int kind, const void* bps, int64_t n_obj, int64_t rows_per_obj, int64_t M, const void* ws,
              size_t ws_bytes) {
    DEXGG_REQUIRE(n && n->kind == kind, "forward: wrong or null net handle");
    DEXGG_REQUIRE(n_obj >= 0 && rows_per_obj >= 1 && M == n_obj * rows_per_obj, "forward: M (%lld) must equal n_obj*rows_per_obj (%lld*%lld)",
                  (long long)M, (long long)n_obj, (long long)rows_per_obj);
    DEXGG_REQUIRE(rows_per_obj < (1ll << 31) && M < (1ll << 40) && n_obj < (1ll << 31), "forward: sizes too large");
    DEXGG_REQUIRE(M == 0 || bps, "forward: null bps");
    FwdWs w;
    const size_t need = carve_ws(n, M, n_obj, nullptr, w);
    if (M > 0 && (!ws || ws_bytes < need)) {
        set_error("forward: workspace of %zu bytes needed, %zu given", need, ws_bytes);
        return DEXGG_E_WORKSPACE;
    }
    return DEXGG_OK;
}

// The hoisted projection is a [n_obj x nbase x 12288] GEMM: 6 - 9 column tiles x n_obj / 128 row tiles, each 192 k-blocks long.
// Up to a few thousand objects that is far less than one wave of CTAs, so the contraction is split over CTAs (each split stores
// its partial sums as its own slice, base_finish_kernel adds them with the bias / affine): 86 -> ~35 us at 1024 objects and
// ~100 -> ~15 us at 128 (the per-GPU share of the strong 8-GPU split).
int base_split(const dexgg_net* n, int64_t n_obj) {
    const int64_t tiles = ((n_obj + 127) / 128) * ((n->nbase + 255) / 256);
    int64_t s = sm_count() / tiles;  // the most splits that still fit one wave of CTAs
    if (s > 16) s = 16;
    return (int)(s < 1 ? 1 : s);
}
// out[o, c] = sc[c] * (sum_s part[s, o, c] + bias[c]) + sh[c]
__global__ void base_finish_kernel(const float* __restrict__ part, int split, int64_t slice, const float* __restrict__ bias,
                                   const float* __restrict__ sc, const float* __restrict__ sh, int64_t total, int nbase,
                                   float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int c = (int)(i % nbase);
    float v = bias[c];
    for (int s = 0; s < split; ++s) v += part[s * slice + i];
    out[i] = sc ? fmaf(v, sc[c], sh[c]) : v;
}

// base[n_obj, nbase] = bn(bps) · W_bpsᵀ + bias    (once per object; for row-wise inputs n_obj == M)
int compute_base(const dexgg_net* n, const float* bps, int64_t n_obj, const FwdWs& w, cudaStream_t st, bool chain = false) {
    if (is_tc(n)) {
        bps_bn_split_kernel<<<nblk(n_obj * IN_BPS), 256, 0, st>>>(bps, n->in_a, n->in_c, w.Ab16, n_obj); ::dexgg::count_launches(1);
        const int split = base_split(n, n_obj);
        if (split > 1) {
            dexgg_gemm g = {};
            g.M = (int)n_obj; g.N = n->nbase; g.n_seg = 1;
            g.seg[0].A = w.Ab16; g.seg[0].lda = 3 * IN_BPS; g.seg[0].B = n->Wb16; g.seg[0].ldb = 3 * IN_BPS; g.seg[0].K = 3 * IN_BPS;
            g.act_from = n->nbase; g.C32 = w.base_part; g.ldc = n->nbase; g.split_k = split; g.tile_n = 256;
            g.split_slice_stride = (int64_t)n_obj * n->nbase;
            int rc = dexgg_gemm_bf16(&g, st);
            if (rc) return rc;
            const int64_t total = (int64_t)n_obj * n->nbase;
            base_finish_kernel<<<nblk(total), 256, 0, st>>>(w.base_part, split, total, n->bb_raw, chain ? n->chain_sc : nullptr,
                                                            chain ? n->chain_sh : nullptr, total, n->nbase, w.base); ::dexgg::count_launches(1);
            return DEXGG_OK;
        }
        LinearB16 q;
        q.A = w.Ab16; q.lda = 3 * IN_BPS; q.W = n->Wb16; q.ldw = 3 * IN_BPS;
        q.M = (int)n_obj; q.N = n->nbase; q.K = 3 * IN_BPS; q.C32 = w.base; q.ldc = n->nbase;
        q.ep.bias = n->bb_raw;
        if (chain) { q.ep.scale = n->chain_sc; q.ep.shift = n->chain_sh; }  // base' = sc * (bps.W + b) + sh
        return linear_bf16(q, st);
    }
    LinearF32 p;
    p.A = bps; p.lda = IN_BPS; p.W = n->Wb; p.ldw = IN_BPS;
    p.M = (int)n_obj; p.N = n->nbase; p.K = IN_BPS; p.C = w.base; p.ldc = n->nbase;
    p.ep.bias = n->bb;
    return linear_f32(p, st);
}

// tail columns (z / normalised pose) of the chunk into up to three activation buffers
void pack_tail(const dexgg_net* n, const FwdWs& w, const float* s0, int n0, const float* s1, int n1, const float* s2, int n2,
               bool affine, int64_t mc, char* d0, int64_t ld0, char* d1, int64_t ld1, char* d2, int64_t ld2, cudaStream_t st) {
    const float* a = affine ? n->tail_a : nullptr;
    const float* c = affine ? n->tail_c : nullptr;
    const int tw = n->tailw;
    if (is_tc(n)) {
        TailDst16 d{{reinterpret_cast<__nv_bfloat16*>(d0), reinterpret_cast<__nv_bfloat16*>(d1), reinterpret_cast<__nv_bfloat16*>(d2)}, {ld0, ld1, ld2}};
        pack_tail16_kernel<<<nblk(mc * tw), 256, 0, st>>>(s0, n0, s1, n1, s2, n2, a, c, tw, mc, d); ::dexgg::count_launches(1);
    } else {
        TailDst d{{reinterpret_cast<float*>(d0), reinterpret_cast<float*>(d1), reinterpret_cast<float*>(d2)}, {ld0, ld1, ld2}};
        pack_tail_kernel<<<nblk(mc * tw), 256, 0, st>>>(s0, n0, s1, n1, s2, n2, a, c, tw, mc, d); ::dexgg::count_launches(1);
    }
}

// One ResBlock on a chunk: T = lrelu(sc*(A·Wtᵀ + base) + sh);  Xout = lrelu(T[:,256:] + sc2*(T[:,:256]·W2ᵀ) + sh2)
int run_block(const dexgg_net* n, int b, const FwdWs& w, const char* A, int64_t lda, int rows_per_obj, int64_t row0, int64_t mc,
              char* Xout, cudaStream_t st) {
    const BlockPack& B = n->blk[b];
    Epilogue e1;
    e1.base = w.base + b * NCAT; e1.ldbase = n->nbase; e1.rows_per_obj = rows_per_obj; e1.row0 = row0;
    e1.scale = B.sc; e1.shift = B.sh; e1.act = 1;
    Epilogue e2;
    e2.scale = B.sc2; e2.shift = B.sh2; e2.act = 1;
    if (is_split(n)) {
        // fp32 on the tensor cores: operands are [hi | lo | hi] x [w_hi | w_hi | w_lo].  First layer: the fc1 columns leave as the
        // split operand of fc2, the fc3 columns (the skip branch) leave in fp32; fc2 adds the fp32 skip and writes the next split
        // operand (and, for the last block, an fp32 copy for the score head)
        LinearB16 p;
        p.A = A; p.lda = lda; p.W = B.Wt16; p.ldw = B.KT16; p.M = (int)mc; p.N = NCAT; p.K = B.KT16; p.ep = e1;
        p.C16 = w.T; p.ldc = 3 * INNER; p.split_seg = INNER; p.c32_from = INNER; p.C32 = w.R32; p.ldc32 = HID;
        int rc = linear_bf16(p, st);
        if (rc) return rc;
        LinearB16 q;
        q.A = w.T; q.lda = 3 * INNER; q.W = B.W2_16; q.ldw = 3 * INNER; q.M = (int)mc; q.N = HID; q.K = 3 * INNER; q.ep = e2;
        q.residual32 = w.R32; q.ldr32 = HID;
        q.C16 = Xout; q.ldc = w.ldx; q.split_seg = HID; q.C32 = w.X32; q.ldc32 = HID;
        return linear_bf16(q, st);
    }
    if (n->precision == DEXGG_PRECISION_BF16) {
        LinearB16 p;
        p.A = A; p.lda = lda; p.W = B.Wt16; p.ldw = B.KT16; p.M = (int)mc; p.N = NCAT; p.K = B.KT16; p.C16 = w.T; p.ldc = NCAT; p.ep = e1;
        int rc = linear_bf16(p, st);
        if (rc) return rc;
        LinearB16 q;
        q.A = w.T; q.lda = NCAT; q.W = B.W2_16; q.ldw = INNER; q.M = (int)mc; q.N = HID; q.K = INNER; q.C16 = Xout; q.ldc = w.ldx; q.ep = e2;
        q.residual16 = w.T + INNER * 2; q.ldr = NCAT;
        return linear_bf16(q, st);
    }
    LinearF32 p;
    p.A = reinterpret_cast<const float*>(A); p.lda = lda; p.W = B.Wt; p.ldw = B.KT; p.M = (int)mc; p.N = NCAT; p.K = B.KT;
    p.C = reinterpret_cast<float*>(w.T); p.ldc = NCAT; p.ep = e1;
    int rc = linear_f32(p, st);
    if (rc) return rc;
    LinearF32 q;
    q.A = reinterpret_cast<const float*>(w.T); q.lda = NCAT; q.W = B.W2; q.ldw = INNER; q.M = (int)mc; q.N = HID; q.K = INNER;
    q.C = reinterpret_cast<float*>(Xout); q.ldc = w.ldx; q.ep = e2;
    q.ep.residual = reinterpret_cast<const float*>(w.T) + INNER; q.ep.ldr = NCAT;
    return linear_f32(q, st);
}
}  // namespace

extern "C" int dexgg_net_create(int kind, const float* const* params, int n_params, int latent_dim, int n_joints,
                                int precision, void* stream, dexgg_net** out) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(out && params, "net_create: null argument");
    DEXGG_REQUIRE(kind >= DEXGG_NET_GENERATOR && kind <= DEXGG_NET_EVALUATOR, "net_create: unknown kind %d", kind);
    DEXGG_REQUIRE(n_params == n_expected(kind), "net_create: kind %d takes %d parameter tensors, got %d", kind, n_expected(kind), n_params);
    DEXGG_REQUIRE(precision == DEXGG_PRECISION_FP32 || precision == DEXGG_PRECISION_BF16 || precision == DEXGG_PRECISION_FP32_TC,
                  "net_create: unsupported precision %d", precision);
    DEXGG_REQUIRE(latent_dim >= 1 && latent_dim <= 16 && n_joints >= 1 && n_joints <= 20, "net_create: latent_dim/n_joints out of range");
    for (int i = 0; i < n_params; ++i) DEXGG_REQUIRE(params[i], "net_create: parameter %d is null", i);
    dexgg_net* n = new (std::nothrow) dexgg_net();
    DEXGG_REQUIRE(n, "net_create: out of host memory");
    n->kind = kind; n->precision = precision; n->latentD = latent_dim; n->nJ = n_joints; n->n_params = n_params;
    n->n_blocks = kind == DEXGG_NET_EVALUATOR ? 3 : 2;
    n->n_tail = kind == DEXGG_NET_GENERATOR ? latent_dim : kind == DEXGG_NET_DISCRIMINATOR ? 12 + n_joints : 12;
    n->tailw = n->n_tail <= 16 ? 16 : 32;
    n->nbase = NCAT * n->n_blocks;
    n->nh = kind == DEXGG_NET_GENERATOR ? n_joints + 9 : 1;
    Carver sz{nullptr};
    carve_net(n, sz);
    n->arena_bytes = sz.used;
    cudaError_t e = cudaMalloc(&n->arena, n->arena_bytes);
    if (e != cudaSuccess) {
        set_error("net_create: cudaMalloc(%zu) failed: %s", n->arena_bytes, cudaGetErrorString(e));
        delete n;
        return DEXGG_E_CUDA;
    }
    Carver c{static_cast<char*>(n->arena)};
    carve_net(n, c);
    int rc = pack_net(n, params, static_cast<cudaStream_t>(stream));
    if (rc != DEXGG_OK) {
        cudaFree(n->arena);
        delete n;
        return rc;
    }
    *out = n;
    return DEXGG_OK;
}

extern "C" int dexgg_net_update(dexgg_net* net, const float* const* params, int n_params, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(net && params && n_params == net->n_params, "net_update: bad arguments");
    return pack_net(net, params, static_cast<cudaStream_t>(stream));
}

extern "C" void dexgg_net_destroy(dexgg_net* net) {
    if (!net) return;
    cudaFree(net->arena);
    delete net;
}

extern "C" size_t dexgg_net_workspace_bytes(const dexgg_net* net, int64_t M, int64_t n_obj) {
    if (!net || M < 0 || n_obj < 0) return 0;
    FwdWs w;
    return carve_ws(net, M, n_obj, nullptr, w);
}

extern "C" int dexgg_generator_forward(const dexgg_net* net, const float* bps, int64_t n_obj, int64_t rows_per_obj,
                                       const float* z, int64_t M, float z_offset, float* rot, float* transl, float* joint,
                                       float* rot6d, void* workspace, size_t workspace_bytes, void* stream) {
    DEXGG_ARCH_OK();
    int rc = check_fwd(net, DEXGG_NET_GENERATOR, bps, n_obj, rows_per_obj, M, workspace, workspace_bytes);
    if (rc) return rc;
    if (M == 0) return DEXGG_OK;
    DEXGG_REQUIRE(z && rot && transl && joint, "generator_forward: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    FwdWs w;
    carve_ws(net, M, n_obj, workspace, w);
    const bool use_chain = net->chain_ok && chain_enabled();
    if ((rc = compute_base(net, bps, n_obj, w, st, use_chain))) return rc;
    const int zD = net->latentD;
    const bool b16 = is_tc(net);
    if (use_chain) {
        ChainArgs a;
        chain_common(net, a, w.base, n_obj, rows_per_obj, M);
        a.z = z; a.z_offset = z_offset; a.rot = rot; a.transl = transl; a.joint = joint; a.rot6d = rot6d;
        return chain_forward(a, st);
    }
    for (int64_t r0 = 0; r0 < M; r0 += w.mc) {
        const int64_t mc = (M - r0 < w.mc) ? M - r0 : w.mc;
        pack_tail(net, w, z + r0 * zD, zD, nullptr, 0, nullptr, 0, false, mc, w.tail, w.tw, w.XT0 + w.xoff * w.es, w.ldx, nullptr, 0, st);
        if ((rc = run_block(net, 0, w, w.tail, w.tw, (int)rows_per_obj, r0, mc, w.XT0, st))) return rc;
        if ((rc = run_block(net, 1, w, w.XT0, w.ldx, (int)rows_per_obj, r0, mc, w.XT1, st))) return rc;
        if (b16) {
            LinearB16 h;
            const int hk = is_split(net) ? 3 * HID : HID;  // split: heads contract over [X_hi | X_lo | X_hi]
            h.A = w.XT1; h.lda = w.ldx; h.W = net->Wh16; h.ldw = hk; h.M = (int)mc; h.N = net->nh; h.K = hk; h.C32 = w.H; h.ldc = 32;
            h.ep.bias = net->bh;
            if ((rc = linear_bf16(h, st))) return rc;
        } else {
            LinearF32 h;
            h.A = reinterpret_cast<const float*>(w.XT1); h.lda = w.ldx; h.W = net->Wh; h.ldw = HID; h.M = (int)mc; h.N = net->nh;
            h.K = HID; h.C = w.H; h.ldc = 32;
            h.ep.bias = net->bh;
            if ((rc = linear_f32(h, st))) return rc;
        }
        pose_finish_kernel<<<nblk(mc, 128), 128, 0, st>>>(w.H, 32, net->nJ, mc, z_offset, rot + r0 * 9, transl + r0 * 3,
                                                         joint + r0 * net->nJ, rot6d ? rot6d + r0 * 6 : nullptr); ::dexgg::count_launches(1);
    }
    DEXGG_LAUNCH_OK("generator_forward kernels");
    return DEXGG_OK;
}

static int score_forward(const dexgg_net* net, int kind, const float* bps, int64_t n_obj, int64_t rows_per_obj,
                         const float* rot, const float* transl, const float* joint, int64_t M, float* p, void* workspace,
                         size_t workspace_bytes, void* stream) {
    DEXGG_ARCH_OK();
    int rc = check_fwd(net, kind, bps, n_obj, rows_per_obj, M, workspace, workspace_bytes);
    if (rc) return rc;
    if (M == 0) return DEXGG_OK;
    DEXGG_REQUIRE(rot && transl && p && (kind == DEXGG_NET_EVALUATOR || joint), "score forward: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    FwdWs w;
    carve_ws(net, M, n_obj, workspace, w);
    const bool use_chain = net->chain_ok && chain_enabled();
    if ((rc = compute_base(net, bps, n_obj, w, st, use_chain))) return rc;
    const int nJ = net->nJ;
    const bool disc = kind == DEXGG_NET_DISCRIMINATOR;
    const bool b16 = is_tc(net), sp = is_split(net);
    if (use_chain) {
        ChainArgs a;
        chain_common(net, a, w.base, n_obj, rows_per_obj, M);
        a.in_rot = rot; a.in_transl = transl; a.in_joint = disc ? joint : nullptr; a.p = p;
        return chain_forward(a, st);
    }
    for (int64_t r0 = 0; r0 < M; r0 += w.mc) {
        const int64_t mc = (M - r0 < w.mc) ? M - r0 : w.mc;
        pack_tail(net, w, rot + r0 * 9, 9, transl + r0 * 3, 3, disc ? joint + r0 * nJ : nullptr, disc ? nJ : 0, true, mc, w.tail, w.tw,
                  w.XT0 + w.xoff * w.es, w.ldx, w.XT1 + w.xoff * w.es, w.ldx, st);
        if ((rc = run_block(net, 0, w, w.tail, w.tw, (int)rows_per_obj, r0, mc, w.XT0, st))) return rc;
        if ((rc = run_block(net, 1, w, w.XT0, w.ldx, (int)rows_per_obj, r0, mc, w.XT1, st))) return rc;
        const char* X = w.XT1;
        if (!disc) {
            if ((rc = run_block(net, 2, w, w.XT1, w.ldx, (int)rows_per_obj, r0, mc, w.XT0, st))) return rc;
            X = w.XT0;
        }
        if (sp) {  // the last block's fp32 output copy
            rowdot_sigmoid_kernel<<<nblk(mc, 8), 256, 0, st>>>(w.X32, HID, net->Wh, net->bh, HID, mc, p + r0); ::dexgg::count_launches(1);
        } else if (b16) {
            rowdot_sigmoid_bf16_kernel<<<nblk(mc, 8), 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(X), w.ldx, net->Wh, net->bh, HID, mc, p + r0); ::dexgg::count_launches(1);
        } else {
            rowdot_sigmoid_kernel<<<nblk(mc, 8), 256, 0, st>>>(reinterpret_cast<const float*>(X), w.ldx, net->Wh, net->bh, HID, mc, p + r0); ::dexgg::count_launches(1);
        }
    }
    DEXGG_LAUNCH_OK("score forward kernels");
    return DEXGG_OK;
}

extern "C" int dexgg_evaluator_forward(const dexgg_net* net, const float* bps, int64_t n_obj, int64_t rows_per_obj,
                                       const float* rot, const float* transl, int64_t M, float* p, void* workspace,
                                       size_t workspace_bytes, void* stream) {
    return score_forward(net, DEXGG_NET_EVALUATOR, bps, n_obj, rows_per_obj, rot, transl, nullptr, M, p, workspace,
                         workspace_bytes, stream);
}

extern "C" int dexgg_discriminator_forward(const dexgg_net* net, const float* bps, int64_t n_obj, int64_t rows_per_obj,
                                           const float* rot, const float* transl, const float* joint, int64_t M, float* p,
                                           void* workspace, size_t workspace_bytes, void* stream) {
    return score_forward(net, DEXGG_NET_DISCRIMINATOR, bps, n_obj, rows_per_obj, rot, transl, joint, M, p, workspace,
                         workspace_bytes, stream);
}
