// bf16 tensor-core fused linear layer for sm_100a:  C = epilogue(A · Wᵀ),  A [M,K] and W [N,K] bf16 K-major in HBM,
// fp32 accumulation in tensor memory (TMEM), epilogue of linear.cuh applied from TMEM.
//
// Persistent, warp-specialised CTA (one per SM), 192 threads:
//   warp 0      TMA producer: 2-D tensor-map loads (128B swizzle) of a 128x64 A box and a BNx64 W box per k-block into a
//               4-stage shared-memory ring, completion on mbarriers (expect_tx)
//   warp 1      MMA issuer: one elected thread issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=BN, K=16) x4 per k-block
//               into one of two TMEM accumulators (2 x 256 columns); tcgen05.commit frees the smem stage / publishes the
//               accumulator; also owns TMEM alloc / dealloc
//   warps 2..5  epilogue: tcgen05.ld 32 lanes x 32 columns -> registers -> hoisted base + bias/BN affine + residual +
//               LeakyReLU -> bf16 (or fp32) -> global; overlaps the next tile's MMAs through the second accumulator
// Roofline: tensor pipe (bf16); see DESIGN.md for the L2-traffic bound of the unfused layer-by-layer form.
#include <cuda_bf16.h>

#include "linear.cuh"
#include "umma.cuh"

namespace dexgg {

constexpr int UM_BM = 128, UM_BK = 64, UM_STAGES = 4, UM_THREADS = 192;
constexpr int UM_A_BYTES = UM_BM * UM_BK * 2;   // 16 KB
constexpr int UM_W_BYTES = 256 * UM_BK * 2;     // 32 KB (BN <= 256)
constexpr int UM_STAGE_BYTES = UM_A_BYTES + UM_W_BYTES;
constexpr int UM_SMEM_BYTES = UM_STAGES * UM_STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr int UM_TMEM_COLS = 512;

struct UmmaParams {
    int M, N, K, BN;
    Epilogue ep;                     // fp32 side inputs (ep.residual unused here)
    const __nv_bfloat16* residual;   // bf16 residual [M, ldr] or null
    int64_t ldr;
    __nv_bfloat16* C16;              // bf16 output or null
    float* C32;                      // fp32 output or null
    int64_t ldc;
    int split_seg, c32_from;         // see LinearB16
    int64_t ldc32;
    const float* residual32;
    int64_t ldr32;
};

__global__ void __launch_bounds__(UM_THREADS, 1)
linear_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const UmmaParams p) {
    extern __shared__ unsigned char um_smem_raw[];
    unsigned char* smem = um_smem_raw + ((1024u - (smem_u32(um_smem_raw) & 1023u)) & 1023u);  // keeps the shared address space
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + UM_STAGES * UM_STAGE_BYTES);
    uint64_t* full = bars;                       // [STAGES]
    uint64_t* empty = bars + UM_STAGES;          // [STAGES]
    uint64_t* tfull = bars + 2 * UM_STAGES;      // [2]
    uint64_t* tempty = bars + 2 * UM_STAGES + 2; // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * UM_STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntn = (p.N + p.BN - 1) / p.BN;
    const int ntiles = ((p.M + UM_BM - 1) / UM_BM) * ntn;
    const int nkb = (p.K + UM_BK - 1) / UM_BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmW);
        for (int s = 0; s < UM_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&tfull[a], 1);
            mbar_init(&tempty[a], 4);  // one arrival per epilogue warp
        }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, UM_TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0, phase = 0;
            const uint32_t bytes = UM_A_BYTES + (uint32_t)p.BN * UM_BK * 2;
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
                const int m0 = (t / ntn) * UM_BM, n0 = (t % ntn) * p.BN;
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    unsigned char* sa = smem + stage * UM_STAGE_BYTES;
                    mbar_expect_tx(&full[stage], bytes);
                    tma_load_2d(sa, &tmA, kb * UM_BK, m0, &full[stage]);
                    tma_load_2d(sa + UM_A_BYTES, &tmW, kb * UM_BK, n0, &full[stage]);
                    if (++stage == UM_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(UM_BM, p.BN);
            int stage = 0, phase = 0, it = 0;
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
                const int acc = it & 1, aphase = (it >> 1) & 1;
                mbar_wait(&tempty[acc], aphase ^ 1);  // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)acc * 256;
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * UM_STAGE_BYTES);
                    const uint64_t adesc = umma_desc_sw128_kmajor(sa), bdesc = umma_desc_sw128_kmajor(sa + UM_A_BYTES);
#pragma unroll
                    for (int k = 0; k < UM_BK / 16; ++k)  // advance 32 B (= 2 x 16 B units) inside the swizzle atom per K=16 step
                        umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                    umma_commit(&empty[stage]);  // stage reusable once these MMAs have read it
                    if (++stage == UM_STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(&tfull[acc]);  // accumulator complete
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int quarter = warp & 3;  // TMEM lanes this warp may read: [32*quarter, 32*quarter + 32)
        const int row_in_tile = quarter * 32 + lane;
        const bool vec_ok = (p.ldc % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.C16) & 15) == 0) &&
                            ((reinterpret_cast<uintptr_t>(p.C32) & 15) == 0) && (p.ldr % 8 == 0) &&
                            ((reinterpret_cast<uintptr_t>(p.residual) & 15) == 0) && (p.ep.ldbase % 4 == 0) &&
                            ((reinterpret_cast<uintptr_t>(p.ep.base) & 15) == 0) && (p.ldc32 % 4 == 0) && (p.ldr32 % 4 == 0) &&
                            ((reinterpret_cast<uintptr_t>(p.residual32) & 15) == 0) && (p.split_seg % 8 == 0) && (p.c32_from % 32 == 0);
        int it = 0;
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
            const int acc = it & 1, aphase = (it >> 1) & 1;
            const int m0 = (t / ntn) * UM_BM, n0 = (t % ntn) * p.BN;
            const int m = m0 + row_in_tile;
            mbar_wait(&tfull[acc], aphase);
            tc_fence_after();
            const float* base_row = p.ep.base ? p.ep.base + ((p.ep.row0 + m) / p.ep.rows_per_obj) * p.ep.ldbase : nullptr;
            for (int c = 0; c < p.BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * 256 + c), r);
                tmem_ld_wait();
                const int n = n0 + c;
                if (m < p.M && n < p.N) {
                    const bool full32 = vec_ok && (n + 32 <= p.N);
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                    if (full32) {
                        if (base_row) {
#pragma unroll
                            for (int j = 0; j < 32; j += 4) {
                                const float4 b = *reinterpret_cast<const float4*>(base_row + n + j);
                                v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
                            }
                        }
                        if (p.ep.bias) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] += __ldg(p.ep.bias + n + j);
                        }
                        if (p.ep.scale) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = fmaf(v[j], __ldg(p.ep.scale + n + j), __ldg(p.ep.shift + n + j));
                        }
                        if (p.residual) {
                            const uint4* rp = reinterpret_cast<const uint4*>(p.residual + (int64_t)m * p.ldr + n);
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const uint4 u = rp[q];
                                const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                                for (int h = 0; h < 4; ++h) {
                                    v[q * 8 + 2 * h] += __uint_as_float(w[h] << 16);
                                    v[q * 8 + 2 * h + 1] += __uint_as_float(w[h] & 0xffff0000u);
                                }
                            }
                        }
                        if (p.residual32) {
                            const float4* rp = reinterpret_cast<const float4*>(p.residual32 + (int64_t)m * p.ldr32 + n);
#pragma unroll
                            for (int q = 0; q < 8; ++q) {
                                const float4 u = rp[q];
                                v[4 * q] += u.x; v[4 * q + 1] += u.y; v[4 * q + 2] += u.z; v[4 * q + 3] += u.w;
                            }
                        }
                        if (p.ep.act) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = v[j] > 0.f ? v[j] : DEXGG_LRELU_SLOPE * v[j];
                        }
                        if (n >= p.c32_from) {  // this column range leaves in fp32 only (the ResBlock skip branch of the split mode)
                            float4* dst = reinterpret_cast<float4*>(p.C32 + (int64_t)m * p.ldc32 + (n - p.c32_from));
#pragma unroll
                            for (int q = 0; q < 8; ++q) dst[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                            continue;
                        }
                        if (p.split_seg) {  // [hi | lo | hi] operand of the next split GEMM (+ optional fp32 copy)
                            uint32_t hi[16], lo[16];
#pragma unroll
                            for (int h = 0; h < 16; ++h) {
                                const __nv_bfloat162 b2 = __floats2bfloat162_rn(v[2 * h], v[2 * h + 1]);
                                hi[h] = *reinterpret_cast<const uint32_t*>(&b2);
                                const __nv_bfloat162 l2 = __floats2bfloat162_rn(v[2 * h] - __uint_as_float(hi[h] << 16),
                                                                                v[2 * h + 1] - __uint_as_float(hi[h] & 0xffff0000u));
                                lo[h] = *reinterpret_cast<const uint32_t*>(&l2);
                            }
                            __nv_bfloat16* row = p.C16 + (int64_t)m * p.ldc + n;
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const uint4 uh = make_uint4(hi[4 * q], hi[4 * q + 1], hi[4 * q + 2], hi[4 * q + 3]);
                                reinterpret_cast<uint4*>(row)[q] = uh;
                                reinterpret_cast<uint4*>(row + p.split_seg)[q] = make_uint4(lo[4 * q], lo[4 * q + 1], lo[4 * q + 2], lo[4 * q + 3]);
                                reinterpret_cast<uint4*>(row + 2 * p.split_seg)[q] = uh;
                            }
                            if (p.C32 && p.c32_from >= (1 << 30)) {
                                float4* dst = reinterpret_cast<float4*>(p.C32 + (int64_t)m * p.ldc32 + n);
#pragma unroll
                                for (int q = 0; q < 8; ++q) dst[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                            }
                            continue;
                        }
                        if (p.C16) {
                            uint4* dst = reinterpret_cast<uint4*>(p.C16 + (int64_t)m * p.ldc + n);
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                uint32_t w[4];
#pragma unroll
                                for (int h = 0; h < 4; ++h) {
                                    const __nv_bfloat162 b2 = __floats2bfloat162_rn(v[q * 8 + 2 * h], v[q * 8 + 2 * h + 1]);
                                    w[h] = *reinterpret_cast<const uint32_t*>(&b2);
                                }
                                dst[q] = make_uint4(w[0], w[1], w[2], w[3]);
                            }
                        }
                        if (p.C32) {
                            float4* dst = reinterpret_cast<float4*>(p.C32 + (int64_t)m * p.ldc32 + n);
#pragma unroll
                            for (int q = 0; q < 8; ++q) dst[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                        }
                    } else {  // ragged edge: scalar, guarded
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (n + j < p.N) {
                                float x = v[j];
                                if (base_row) x += base_row[n + j];
                                if (p.ep.bias) x += p.ep.bias[n + j];
                                if (p.ep.scale) x = fmaf(x, p.ep.scale[n + j], p.ep.shift[n + j]);
                                if (p.residual) x += __bfloat162float(p.residual[(int64_t)m * p.ldr + n + j]);
                                if (p.residual32) x += p.residual32[(int64_t)m * p.ldr32 + n + j];
                                if (p.ep.act) x = x > 0.f ? x : DEXGG_LRELU_SLOPE * x;
                                if (n + j >= p.c32_from) {
                                    p.C32[(int64_t)m * p.ldc32 + (n + j - p.c32_from)] = x;
                                } else if (p.split_seg) {
                                    const __nv_bfloat16 h = __float2bfloat16_rn(x);
                                    __nv_bfloat16* row = p.C16 + (int64_t)m * p.ldc + n + j;
                                    row[0] = h;
                                    row[p.split_seg] = __float2bfloat16_rn(x - __bfloat162float(h));
                                    row[2 * p.split_seg] = h;
                                    if (p.C32 && p.c32_from >= (1 << 30)) p.C32[(int64_t)m * p.ldc32 + n + j] = x;
                                } else {
                                    if (p.C16) p.C16[(int64_t)m * p.ldc + n + j] = __float2bfloat16_rn(x);
                                    if (p.C32) p.C32[(int64_t)m * p.ldc32 + n + j] = x;
                                }
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, UM_TMEM_COLS);
    }
}

}  // namespace dexgg

// ------------------------------------------------------------------------------------------------ host side
namespace dexgg {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled is not available from the driver");
        return DEXGG_E_CUDA;
    }
    DEXGG_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld * 2) % 16 == 0, "tensor map: base and row pitch must be 16-byte aligned (ld=%lld)", (long long)ld);
    DEXGG_REQUIRE(rows >= 1 && cols >= 1 && box_rows >= 1 && box_rows <= 256, "tensor map: bad extents");
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    const cuuint32_t box[2] = {(cuuint32_t)UM_BK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld ld=%lld box_rows=%d)", (int)r, (long long)rows,
                  (long long)cols, (long long)ld, box_rows);
        return DEXGG_E_CUDA;
    }
    return DEXGG_OK;
}

int linear_bf16(const LinearB16& q, cudaStream_t st) {
    DEXGG_REQUIRE(q.M >= 0 && q.N >= 1 && q.K >= 1 && q.A && q.W && (q.C16 || q.C32), "linear_bf16: bad arguments (M=%d N=%d K=%d)", q.M, q.N, q.K);
    // K itself may be any length: the tensor maps clip at K and the TMA zero-fills the rest of the last k-block; only the
    // row pitches must be 16-byte multiples (checked by make_tmap_bf16)
    if (q.M == 0) return DEXGG_OK;
    int BN = 256;
    if (q.N <= 32) BN = 32;
    else if (q.N <= 64) BN = 64;
    else if (q.N <= 128) BN = 128;
    CUtensorMap tmA, tmW;
    int rc = make_tmap_bf16(&tmA, q.A, q.M, q.K, q.lda, UM_BM);
    if (rc) return rc;
    if ((rc = make_tmap_bf16(&tmW, q.W, q.N, q.K, q.ldw, BN))) return rc;
    static std::atomic<unsigned long long> attr_mask{0};
    DEXGG_SMEM_OPT_IN(linear_umma_kernel, UM_SMEM_BYTES, attr_mask);
    UmmaParams p;
    p.M = q.M; p.N = q.N; p.K = q.K; p.BN = BN; p.ep = q.ep; p.ep.residual = nullptr;
    p.residual = reinterpret_cast<const __nv_bfloat16*>(q.residual16); p.ldr = q.ldr;
    p.C16 = reinterpret_cast<__nv_bfloat16*>(q.C16); p.C32 = q.C32; p.ldc = q.ldc;
    p.split_seg = q.split_seg; p.c32_from = q.c32_from; p.ldc32 = q.ldc32 ? q.ldc32 : q.ldc; p.residual32 = q.residual32; p.ldr32 = q.ldr32;
    DEXGG_REQUIRE(q.c32_from >= (1 << 30) || q.C32, "linear_bf16: c32_from needs an fp32 output");
    DEXGG_REQUIRE(!q.split_seg || q.C16, "linear_bf16: split output needs a bf16 output");
    const int ntiles = ((q.M + UM_BM - 1) / UM_BM) * ((q.N + BN - 1) / BN);
    const int grid = ntiles < sm_count() ? ntiles : sm_count();
    linear_umma_kernel<<<grid, UM_THREADS, UM_SMEM_BYTES, st>>>(tmA, tmW, p); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("linear_umma_kernel");
    return DEXGG_OK;
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ src, int64_t lds, __nv_bfloat16* __restrict__ dst, int64_t ldd,
                                   int64_t rows, int cols) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * cols) return;
    const int64_t r = i / cols;
    const int c = (int)(i % cols);
    dst[r * ldd + c] = __float2bfloat16_rn(src[r * lds + c]);
}

// dst[c, r] = bf16(src[r, c]): the K-major operand of a GEMM that contracts over the ROWS of src (wgrad: dY^T, X^T; dgrad: W^T)
__global__ void f32_to_bf16_t_kernel(const float* __restrict__ src, int64_t lds, __nv_bfloat16* __restrict__ dst, int64_t ldd,
                                     int64_t rows, int cols) {
    __shared__ float tile[32][33];
    const int64_t r0 = (int64_t)blockIdx.y * 32;
    const int c0 = blockIdx.x * 32;
    for (int i = threadIdx.y; i < 32; i += 8) {
        const int64_t r = r0 + i;
        const int c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < rows && c < cols) ? src[r * lds + c] : 0.f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += 8) {
        const int c = c0 + i;
        const int64_t r = r0 + threadIdx.x;
        if (c < cols && r < rows) dst[(int64_t)c * ldd + r] = __float2bfloat16_rn(tile[threadIdx.x][i]);
    }
}

}  // namespace dexgg

using namespace dexgg;

extern "C" int dexgg_f32_to_bf16_t(const float* src, int64_t lds, void* dst, int64_t ldd, int64_t rows, int cols, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(rows >= 0 && cols >= 1 && (rows == 0 || (src && dst)) && ldd >= rows, "f32_to_bf16_t: bad arguments");
    if (rows == 0) return DEXGG_OK;
    const dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32));
    DEXGG_REQUIRE(grid.y <= 65535, "f32_to_bf16_t: too many rows (%lld)", (long long)rows);
    f32_to_bf16_t_kernel<<<grid, dim3(32, 8), 0, static_cast<cudaStream_t>(stream)>>>(src, lds, reinterpret_cast<__nv_bfloat16*>(dst), ldd, rows,
                                                                                     cols);
    ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("f32_to_bf16_t_kernel");
    return DEXGG_OK;
}

// C ABI: one bf16 tensor-core layer on caller-provided bf16 operands (isolated test / building block).
extern "C" int dexgg_linear_bf16(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, const float* bias,
                                 const float* scale, const float* shift, const void* residual_bf16, int64_t ldr, int act,
                                 void* C_bf16, float* C_f32, int64_t ldc, void* stream) {
    DEXGG_ARCH_OK();
    LinearB16 q;
    q.A = A; q.lda = lda; q.W = W; q.ldw = ldw; q.M = M; q.N = N; q.K = K;
    q.ep.bias = bias; q.ep.scale = scale; q.ep.shift = shift; q.ep.act = act;
    q.residual16 = residual_bf16; q.ldr = ldr; q.C16 = C_bf16; q.C32 = C_f32; q.ldc = ldc;
    DEXGG_REQUIRE(!scale == !shift, "linear_bf16: scale and shift come together");
    return linear_bf16(q, static_cast<cudaStream_t>(stream));
}

extern "C" int dexgg_f32_to_bf16(const float* src, int64_t lds, void* dst, int64_t ldd, int64_t rows, int cols, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(rows >= 0 && cols >= 1 && (rows == 0 || (src && dst)), "f32_to_bf16: bad arguments");
    if (rows == 0) return DEXGG_OK;
    f32_to_bf16_kernel<<<(unsigned)((rows * cols + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        src, lds, reinterpret_cast<__nv_bfloat16*>(dst), ldd, rows, cols); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("f32_to_bf16_kernel");
    return DEXGG_OK;
}
