// General bf16 tensor-core GEMM of the TRAINING step (sm_100a, tcgen05 / TMEM / TMA):
//
//     C[M,N] (+)= sum over segments s of  op(A_s)[M,K_s] . op(B_s)[N,K_s]^T      (+ bias, LeakyReLU on a column range)
//
// Every GEMM of DexGanGrasp.update_dexgangrasp / update_dexevaluator (DexGanGrasp/models/dexgangrasp.py:739-823) is one call:
//   forward   Y  = [xhat_bps | X,tail] . [W_bps | W_xt]^T      two K SEGMENTS accumulate into one TMEM tile: the reference's
//                                                               torch.cat([X0, X]) (networks.py:97,153,325) is never materialised
//   dgrad     dX = dY . W          B operand MN-major: W is read in its natural [N_out, K_in] layout, no transposed copy
//   wgrad     dW = dY^T . X        A and B operands MN-major: dY [rows, N_out] and X [rows, K_in] as they lie in memory
// Operands are bf16 in HBM.  K-major operand: element (i, k) at p[i*ld + k]; MN-major operand: element (i, k) at p[k*ld + i].
// Both are fetched by 2-D TMA tensor maps with the 128-byte swizzle: a K-major tile is one [rows x 64 k] box, an MN-major tile
// is a row of [64 k x 64 i] boxes (8 KB each, 64-element blocks of the MN index 8 KB apart: the canonical UMMA MN-major layout,
// leading byte offset 8192, stride byte offset 1024).
//
// Persistent warp-specialised CTA, 192 threads: warp 0 TMA producer (4-stage ring), warp 1 single-thread tcgen05.mma issuer
// (M = 128, N = BN <= 256, K = 16) into one of two TMEM accumulators, warps 2..5 epilogue (tcgen05.ld -> bias / LeakyReLU ->
// bf16 and / or fp32 stores, or fp32 atomic adds when the contraction is split over CTAs for outputs with few tiles).
#include <cuda_bf16.h>

#include "umma.cuh"

namespace dexgg {

constexpr int GM_BM = 128, GM_BK = 64, GM_STAGES = 4, GM_THREADS = 192, GM_MAX_SEG = 3;
constexpr int GM_A_BYTES = GM_BM * GM_BK * 2;  // 16 KB
constexpr int GM_B_BYTES = 256 * GM_BK * 2;    // 32 KB (BN <= 256)
constexpr int GM_STAGE_BYTES = GM_A_BYTES + GM_B_BYTES;
constexpr int GM_SMEM_BYTES = GM_STAGES * GM_STAGE_BYTES + 1024 + 256;

struct GemmParams {
    int M, N, BN, n_seg, split;
    int kb[GM_MAX_SEG];        // k-blocks of each segment
    int a_mn[GM_MAX_SEG], b_mn[GM_MAX_SEG];
    const float* bias;
    int act_from;              // columns >= act_from: LeakyReLU(0.2)
    __nv_bfloat16* C16;
    float* C32;
    int64_t ldc;
    int accumulate;            // 1: C32 += (plain read-modify-write, split == 1)
    int64_t slice_stride;      // split > 1: 0 = atomic accumulation into C32; > 0 = split ks stores its partial sums at C32 + ks * slice_stride
};

struct GemmMaps {
    CUtensorMap a[GM_MAX_SEG], b[GM_MAX_SEG];
};

// MN-major operand tile under the 128-byte swizzle: 64-element blocks of the MN index 8192 B apart (LBO), 8-k-row groups 1024 B
// apart (SBO); one K = 16 step = 2048 B
__device__ __forceinline__ uint64_t umma_desc_sw128_mnmajor(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3ffffu) >> 4);
    d |= (uint64_t)(8192 >> 4) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__host__ __device__ __forceinline__ uint32_t umma_idesc_bf16_major(int M, int N, int a_mn, int b_mn) {
    return umma_idesc_bf16(M, N) | ((uint32_t)(a_mn ? 1 : 0) << 15) | ((uint32_t)(b_mn ? 1 : 0) << 16);
}

__global__ void __launch_bounds__(GM_THREADS, 1) gemm_umma_kernel(const __grid_constant__ GemmMaps maps, const GemmParams p) {
    extern __shared__ unsigned char gm_smem_raw[];
    unsigned char* smem = gm_smem_raw + ((1024u - (smem_u32(gm_smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + GM_STAGES * GM_STAGE_BYTES);
    uint64_t* full = bars;
    uint64_t* empty = bars + GM_STAGES;
    uint64_t* tfull = bars + 2 * GM_STAGES;
    uint64_t* tempty = bars + 2 * GM_STAGES + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * GM_STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntn = (p.N + p.BN - 1) / p.BN;
    const int ntiles = ((p.M + GM_BM - 1) / GM_BM) * ntn;
    int nkb = 0;
    for (int s = 0; s < p.n_seg; ++s) nkb += p.kb[s];
    const int nwork = ntiles * p.split;                      // work item = (tile, k-range)
    const int kb_per = (nkb + p.split - 1) / p.split;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < p.n_seg; ++s) {
            tma_prefetch_desc(&maps.a[s]);
            tma_prefetch_desc(&maps.b[s]);
        }
        for (int s = 0; s < GM_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&tfull[a], 1);
            mbar_init(&tempty[a], 4);
        }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0, phase = 0;
            const uint32_t bytes = GM_A_BYTES + (uint32_t)p.BN * GM_BK * 2;
            for (int w = blockIdx.x; w < nwork; w += gridDim.x) {
                const int t = w / p.split, ks = w % p.split;
                const int m0 = (t / ntn) * GM_BM, n0 = (t % ntn) * p.BN;
                const int kb0 = ks * kb_per, kb1 = (kb0 + kb_per < nkb) ? kb0 + kb_per : nkb;
                for (int kb = kb0; kb < kb1; ++kb) {
                    int s = 0, k = kb;
                    while (k >= p.kb[s]) { k -= p.kb[s]; ++s; }
                    mbar_wait(&empty[stage], phase ^ 1);
                    unsigned char* sa = smem + stage * GM_STAGE_BYTES;
                    unsigned char* sb = sa + GM_A_BYTES;
                    mbar_expect_tx(&full[stage], bytes);
                    if (p.a_mn[s]) {
                        for (int j = 0; j < GM_BM / 64; ++j) tma_load_2d(sa + j * 8192, &maps.a[s], m0 + 64 * j, k * GM_BK, &full[stage]);
                    } else {
                        tma_load_2d(sa, &maps.a[s], k * GM_BK, m0, &full[stage]);
                    }
                    if (p.b_mn[s]) {
                        for (int j = 0; j < p.BN / 64; ++j) tma_load_2d(sb + j * 8192, &maps.b[s], n0 + 64 * j, k * GM_BK, &full[stage]);
                    } else {
                        tma_load_2d(sb, &maps.b[s], k * GM_BK, n0, &full[stage]);
                    }
                    if (++stage == GM_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            int stage = 0, phase = 0, it = 0;
            for (int w = blockIdx.x; w < nwork; w += gridDim.x, ++it) {
                const int ks = w % p.split;
                const int kb0 = ks * kb_per, kb1 = (kb0 + kb_per < nkb) ? kb0 + kb_per : nkb;
                const int acc = it & 1, aphase = (it >> 1) & 1;
                mbar_wait(&tempty[acc], aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)acc * 256;
                for (int kb = kb0; kb < kb1; ++kb) {
                    int s = 0, k = kb;
                    while (k >= p.kb[s]) { k -= p.kb[s]; ++s; }
                    const uint32_t idesc = umma_idesc_bf16_major(GM_BM, p.BN, p.a_mn[s], p.b_mn[s]);
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * GM_STAGE_BYTES), sb = sa + GM_A_BYTES;
                    const uint64_t ad = p.a_mn[s] ? umma_desc_sw128_mnmajor(sa) : umma_desc_sw128_kmajor(sa);
                    const uint64_t bd = p.b_mn[s] ? umma_desc_sw128_mnmajor(sb) : umma_desc_sw128_kmajor(sb);
                    const uint64_t astep = p.a_mn[s] ? (2048 >> 4) : 2, bstep = p.b_mn[s] ? (2048 >> 4) : 2;
#pragma unroll
                    for (int q = 0; q < GM_BK / 16; ++q) umma_bf16(d_tmem, ad + astep * q, bd + bstep * q, idesc, (kb > kb0) || q);
                    umma_commit(&empty[stage]);
                    if (++stage == GM_STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(&tfull[acc]);
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int quarter = warp & 3;
        const int row_in_tile = quarter * 32 + lane;
        const bool vec_ok = (p.ldc % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.C16) & 15) == 0) &&
                            ((reinterpret_cast<uintptr_t>(p.C32) & 15) == 0);
        int it = 0;
        for (int w = blockIdx.x; w < nwork; w += gridDim.x, ++it) {
            const int t = w / p.split, ks = w % p.split;
            const int kb0 = ks * kb_per;
            const int acc = it & 1, aphase = (it >> 1) & 1;
            const int m0 = (t / ntn) * GM_BM, n0 = (t % ntn) * p.BN;
            const int m = m0 + row_in_tile;
            mbar_wait(&tfull[acc], aphase);
            tc_fence_after();
            const bool empty_range = kb0 >= nkb;  // a split with no k-blocks: the accumulator holds nothing valid
            for (int c = 0; c < p.BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * 256 + c), r);
                tmem_ld_wait();
                const int n = n0 + c;
                if (m >= p.M || n >= p.N || (empty_range && !p.slice_stride)) continue;
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = empty_range ? 0.f : __uint_as_float(r[j]);
                if (p.split > 1 && !p.slice_stride) {  // partial sum of a split contraction: fp32 atomic accumulation (C32 zeroed by the caller)
                    float* dst = p.C32 + (int64_t)m * p.ldc + n;
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (n + j < p.N) atomicAdd(dst + j, v[j]);
                    continue;
                }
                const bool full32 = n + 32 <= p.N;
                if (p.bias) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (full32 || n + j < p.N) v[j] += __ldg(p.bias + n + j);
                }
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (n + j >= p.act_from) v[j] = v[j] > 0.f ? v[j] : 0.2f * v[j];
                if (full32 && vec_ok) {
                    if (p.C16) {
                        uint4* dst = reinterpret_cast<uint4*>(p.C16 + (int64_t)m * p.ldc + n);
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            uint32_t wv[4];
#pragma unroll
                            for (int h = 0; h < 4; ++h) {
                                const __nv_bfloat162 b2 = __floats2bfloat162_rn(v[q * 8 + 2 * h], v[q * 8 + 2 * h + 1]);
                                wv[h] = *reinterpret_cast<const uint32_t*>(&b2);
                            }
                            dst[q] = make_uint4(wv[0], wv[1], wv[2], wv[3]);
                        }
                    }
                    if (p.C32) {
                        float4* dst = reinterpret_cast<float4*>(p.C32 + (int64_t)ks * p.slice_stride + (int64_t)m * p.ldc + n);
                        if (p.accumulate) {
#pragma unroll
                            for (int q = 0; q < 8; ++q) {
                                const float4 o = dst[q];
                                dst[q] = make_float4(o.x + v[4 * q], o.y + v[4 * q + 1], o.z + v[4 * q + 2], o.w + v[4 * q + 3]);
                            }
                        } else {
#pragma unroll
                            for (int q = 0; q < 8; ++q) dst[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                        }
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (n + j < p.N) {
                            if (p.C16) p.C16[(int64_t)m * p.ldc + n + j] = __float2bfloat16_rn(v[j]);
                            if (p.C32) {
                                float* d = p.C32 + (int64_t)ks * p.slice_stride + (int64_t)m * p.ldc + n + j;
                                *d = p.accumulate ? *d + v[j] : v[j];
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
        tmem_dealloc(tmem_base, 512);
    }
}

// tensor map with a [64 x box_rows] box (64 bf16 = 128 B inner, SWIZZLE_128B); defined in linear_umma.cu
int make_tmap_bf16(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows);

}  // namespace dexgg

using namespace dexgg;

extern "C" int dexgg_gemm_bf16(const dexgg_gemm* g, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(g && g->M >= 0 && g->N >= 1 && g->n_seg >= 1 && g->n_seg <= GM_MAX_SEG && (g->C16 || g->C32), "gemm_bf16: bad arguments");
    if (g->M == 0) return DEXGG_OK;
    const int split = g->split_k > 1 ? g->split_k : 1;
    DEXGG_REQUIRE(split == 1 || (g->C32 && !g->C16 && !g->bias && g->act_from >= g->N && !g->accumulate),
                  "gemm_bf16: a split contraction produces raw fp32 partial sums only");
    DEXGG_REQUIRE(g->split_slice_stride >= 0 && g->split_slice_stride % 4 == 0 && (split > 1 || g->split_slice_stride == 0),
                  "gemm_bf16: bad split_slice_stride");
    int BN = 256;
    if (g->N <= 32) BN = 32;
    else if (g->N <= 64) BN = 64;
    else if (g->N <= 128) BN = 128;
    else {  // fewest waves of (128 + BN)-wide operand traffic: the big-K layers are bound by L2 -> SM bytes per tile, not MMA issue
        const int64_t mt = (g->M + GM_BM - 1) / GM_BM;
        int64_t best = -1;
        for (int bn : {256, 192, 128, 64}) {
            const int64_t tiles = mt * ((g->N + bn - 1) / bn) * split;
            const int64_t cost = ((tiles + sm_count() - 1) / sm_count()) * (128 + bn);
            if (best < 0 || cost < best) { best = cost; BN = bn; }
        }
    }
    if (g->tile_n == 64 || g->tile_n == 128 || g->tile_n == 192 || g->tile_n == 256) BN = g->tile_n;
    GemmMaps maps;
    GemmParams p;
    p.M = g->M; p.N = g->N; p.BN = BN; p.n_seg = g->n_seg; p.split = split;
    p.bias = g->bias; p.act_from = g->act_from < 0 ? 0 : g->act_from;
    p.C16 = reinterpret_cast<__nv_bfloat16*>(g->C16); p.C32 = g->C32; p.ldc = g->ldc; p.accumulate = g->accumulate;
    p.slice_stride = g->split_slice_stride;
    int nkb = 0;
    for (int s = 0; s < GM_MAX_SEG; ++s) { p.kb[s] = 0; p.a_mn[s] = p.b_mn[s] = 0; }
    for (int s = 0; s < g->n_seg; ++s) {
        const dexgg_gemm_seg& q = g->seg[s];
        DEXGG_REQUIRE(q.A && q.B && q.K >= 1, "gemm_bf16: segment %d is empty", s);
        DEXGG_REQUIRE(!q.b_mn_major || BN % 64 == 0, "gemm_bf16: an MN-major B operand needs N tiles of 64 columns (N = %d)", g->N);
        p.kb[s] = (q.K + GM_BK - 1) / GM_BK;
        p.a_mn[s] = q.a_mn_major ? 1 : 0;
        p.b_mn[s] = q.b_mn_major ? 1 : 0;
        nkb += p.kb[s];
        int rc;
        // K-major: matrix [rows = M or N, cols = K], box 64 k x (BM | BN) rows.  MN-major: matrix [rows = K, cols = M or N], box 64 x 64.
        if (q.a_mn_major) rc = make_tmap_bf16(&maps.a[s], q.A, q.K, g->M, q.lda, 64);
        else rc = make_tmap_bf16(&maps.a[s], q.A, g->M, q.K, q.lda, GM_BM);
        if (rc) return rc;
        if (q.b_mn_major) rc = make_tmap_bf16(&maps.b[s], q.B, q.K, g->N, q.ldb, 64);
        else rc = make_tmap_bf16(&maps.b[s], q.B, g->N, q.K, q.ldb, BN);
        if (rc) return rc;
    }
    for (int s = g->n_seg; s < GM_MAX_SEG; ++s) { maps.a[s] = maps.a[0]; maps.b[s] = maps.b[0]; }
    DEXGG_REQUIRE(split <= nkb, "gemm_bf16: split_k (%d) exceeds the number of k-blocks (%d)", split, nkb);
    static std::atomic<unsigned long long> attr_mask{0};
    DEXGG_SMEM_OPT_IN(gemm_umma_kernel, GM_SMEM_BYTES, attr_mask);
    const int64_t nwork = (int64_t)((g->M + GM_BM - 1) / GM_BM) * ((g->N + BN - 1) / BN) * split;
    const int grid = (int)(nwork < sm_count() ? nwork : sm_count());
    gemm_umma_kernel<<<grid, GM_THREADS, GM_SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(maps, p);
    count_launches(1);
    DEXGG_LAUNCH_OK("gemm_umma_kernel");
    return DEXGG_OK;
}
