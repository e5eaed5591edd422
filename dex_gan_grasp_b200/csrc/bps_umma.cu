// Stage 1 — Basis Point Set encoding, filter scan on the tensor cores (sm_100a tcgen05).
//
// Same contract as bps_encode.cu (reference call sites DexGanGrasp/data/bps_encoder.py:28-55; oracle/bps_oracle.c): per basis
// point the fp32 direct-difference arg-min over the cloud, lowest index on an exact tie, bit-exact.  Same two-phase scheme as
// bps_nn_filter_kernel: a cheap scan of a(k) = |x_k|^2 - 2 b.x_k only FILTERS 32-point chunks, the reference arithmetic is then
// evaluated on the candidate chunks.  What changes is where a(k) comes from: instead of three FFMA per (basis, point) pair on
// the FP32 pipe, the [128 basis x 256 points] tile of a() is ONE tcgen05.mma pair (K = 32 bf16):
//
//   every coordinate is split into three bf16 parts (8 + 8 + 8 significant bits: the fp32 value exactly, up to 2^-24 relative),
//   t = -2x = t0 + t1 + t2,  b = b0 + b1 + b2, and the K axis carries all nine part products of the three coordinates
//   (27 slots), |x|^2 = n0 + n1 + n2 against 1, 1, 1 (3 slots), 2 slots of zeros.  bf16 x bf16 products are exact in the fp32
//   accumulator, so the only error is the tensor core's fp32 accumulation of 30 terms and the <= 2^-24 relative residue of
//   the three-way split.  MEASURED (tools/bps_tc_error.py: the same operands through the same tcgen05 instruction against
//   float64, 5e8 pairs - surface clouds, Gaussian clouds, clouds 9 units off-centre, sigma = 30):
//       max |a_tc - a_exact| = 2.04 u (3 C^2 + 2 |b|_1 C),   u = 2^-24, C = max |coordinate| of the CTA's point range;
//   the kernel assumes E = 8 u (3 C^2 + 2 |b|_1 C) - four times that, the same constant as the FP32-pipe filter - and, like
//   that kernel, uses 1.5 x the resulting margin.
//   The candidate rule and its margin are those of bps_nn_filter_kernel with this E.
//
// The epilogue reads a() from TMEM (thread = TMEM lane = basis point, 32 columns = one chunk per tcgen05.ld) and needs ONE
// FMNMX3 per two pairs + ~8 ALU operations per chunk: the scan is bound by the ALU pipe (min / select, 64 lanes / clk / SM)
// against 128 / 3 = 43 for the FFMA form.
//
// CTA (two per SM) = 11 warps over 512 basis points: two scan groups of four warps (group g owns one 128-column accumulator
// and the M blocks 2j + g; the four scan warps of an SMSP belong to four different groups of the two resident CTAs, so one
// group's wait for its MMA is covered by the scans of the others), two converter warps, which turn the raw AoS tile (TMA bulk
// copy, as it lies in HBM) into the B operand (64-byte swizzle, K-major) of the NEXT tile while the current one is scanned, and
// the MMA issuer.  The basis points stay resident as the A operand (packed once per call by bps_basis_pack_kernel, one bulk
// copy per CTA).  BU_CFG 0 is the one-CTA-per-SM layout (1024 basis points, four groups): measured 10 % slower.
#include <atomic>
#include <cstdio>
#include <cstdlib>

#include "../../include/dexgg.h"
#include "common.cuh"
#include "umma.cuh"

namespace dexgg {

#ifndef BU_CFG
#define BU_CFG 1
#endif
#if BU_CFG == 0   // one CTA per SM: 1024 basis points, four scan groups
constexpr int BU_GROUPS = 4;                  // scan groups of four warps; group g owns the accumulator g and the M blocks GROUPS * j + g
constexpr int BU_CONV_WARPS = 4;              // the warps after the scan warps
constexpr int BU_MB = 8;                      // 128-row M blocks (basis points) per CTA
constexpr int BU_TILE = 512;                  // points per B buffer
constexpr int BU_CTAS_PER_SM = 1;
#else             // two CTAs per SM: 512 basis points, two scan groups each
constexpr int BU_GROUPS = 2;
constexpr int BU_CONV_WARPS = 2;              // (one converter warp per CTA measured slower: it falls behind the scan)
constexpr int BU_MB = 4;
constexpr int BU_TILE = 256;
constexpr int BU_CTAS_PER_SM = 2;
#endif
constexpr int BU_SCAN_WARPS = 4 * BU_GROUPS;
constexpr int BU_MMA_WARP = BU_SCAN_WARPS + BU_CONV_WARPS;
constexpr int BU_THREADS = 32 * (BU_MMA_WARP + 1);
constexpr int BU_NBLK = 128;                  // points per accumulator (N of one MMA)
constexpr int BU_SLOTS = 1;                   // accumulators per scan group (2 x 64 columns measured slower: more hand-offs per point)
constexpr int BU_RJ = BU_MB / BU_GROUPS;      // basis points per scan thread
constexpr int BU_BASIS = BU_MB * 128;
constexpr int BU_RING = 2048 / BU_TILE;       // raw tiles kept in shared memory (a range of <= 2048 points stays resident)
constexpr int BU_NACC = BU_GROUPS * BU_SLOTS;
constexpr int BU_TMEM_COLS = BU_NACC * BU_NBLK;
static_assert(BU_TMEM_COLS * BU_CTAS_PER_SM <= 512 && (BU_TMEM_COLS & (BU_TMEM_COLS - 1)) == 0, "TMEM budget");
constexpr int BU_CHUNK = 32;                  // points per candidate chunk
constexpr int BU_CPT = BU_TILE / BU_CHUNK;    // chunks per tile
constexpr int BU_A_BYTES = BU_BASIS * 64;     // [1024 x 32] bf16
constexpr int BU_B_BYTES = BU_TILE * 64;      // [512 x 32] bf16 per buffer
constexpr int BU_RAW_FLOATS = 3 * BU_TILE;
constexpr size_t BU_SMEM = 1024 + (size_t)BU_A_BYTES + 2 * (size_t)BU_B_BYTES + (size_t)BU_RING * BU_RAW_FLOATS * 4 + (size_t)BU_BASIS * 5;

__device__ __forceinline__ float bu_d2(float bx, float by, float bz, float x, float y, float z) {  // the reference arithmetic
    const float dx = __fsub_rn(x, bx), dy = __fsub_rn(y, by), dz = __fsub_rn(z, bz);
    return __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
}
__device__ __forceinline__ unsigned long long bu_key(float d2, uint32_t idx) {
    return ((unsigned long long)__float_as_uint(d2) << 32) | idx;
}
__device__ __forceinline__ uint32_t bf16_bits_rn(float v) {
    uint16_t h;
    asm("cvt.rn.bf16.f32 %0, %1;" : "=h"(h) : "f"(v));
    return h;
}
// v = p[0] + p[1] + p[2] (bf16 each) up to 2^-24 |v|
__device__ __forceinline__ void split3(float v, uint32_t (&p)[3]) {
    p[0] = bf16_bits_rn(v);
    const float r1 = v - __uint_as_float(p[0] << 16);
    p[1] = bf16_bits_rn(r1);
    const float r2 = r1 - __uint_as_float(p[1] << 16);
    p[2] = bf16_bits_rn(r2);
}
// minimum of the 32 values of one chunk: 16 FMNMX3 / FMNMX
__device__ __forceinline__ float min32(const uint32_t (&r)[32]) {
    float m10[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) m10[i] = min3(__uint_as_float(r[3 * i]), __uint_as_float(r[3 * i + 1]), __uint_as_float(r[3 * i + 2]));
    const float a = min3(m10[0], m10[1], m10[2]), b = min3(m10[3], m10[4], m10[5]), c = min3(m10[6], m10[7], m10[8]);
    const float d = min3(m10[9], __uint_as_float(r[30]), __uint_as_float(r[31]));
    return fminf(min3(a, b, c), d);
}
// one operand row of 32 bf16 (64 B) under the 64-byte swizzle: 16-byte chunk q of row r sits at chunk q ^ ((r >> 1) & 3)
__device__ __forceinline__ void store_row_sw64(unsigned char* base, int r, const uint32_t (&e)[32]) {
    uint4* row = reinterpret_cast<uint4*>(base + (size_t)r * 64);
    const int sw = (r >> 1) & 3;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint4 v;
        v.x = e[8 * q + 0] | (e[8 * q + 1] << 16);
        v.y = e[8 * q + 2] | (e[8 * q + 3] << 16);
        v.z = e[8 * q + 4] | (e[8 * q + 5] << 16);
        v.w = e[8 * q + 6] | (e[8 * q + 7] << 16);
        row[q ^ sw] = v;
    }
}

// The A operand of every CTA that shares blockIdx.x is the same: packed once per call (row p of apack = the 32 bf16 of basis
// point p, already in its swizzled position inside its 8-row group), with the two norms the margin needs per 1024-point block.
__global__ void __launch_bounds__(256) bps_basis_pack_kernel(const float* __restrict__ basis, int P, int rows, unsigned char* __restrict__ apack,
                                                             float* __restrict__ bmax /* [blocks][2], zeroed */) {
    const int p = blockIdx.x * 256 + threadIdx.x;
    float b1 = 0.f, b2 = 0.f;
    if (p < rows) {
        float q[3] = {0.f, 0.f, 0.f};
        if (p < P) {
            q[0] = basis[3 * p]; q[1] = basis[3 * p + 1]; q[2] = basis[3 * p + 2];
        }
        b1 = fabsf(q[0]) + fabsf(q[1]) + fabsf(q[2]);
        b2 = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
        uint32_t e[32];
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) {
            uint32_t parts[3];
            split3(q[cc], parts);
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) e[9 * cc + 3 * i + j] = parts[j];
        }
        e[27] = e[28] = e[29] = 0x3f80u;  // 1.0
        e[30] = e[31] = 0u;
        store_row_sw64(apack, p, e);
    }
    for (int o = 16; o; o >>= 1) {
        b1 = fmaxf(b1, __shfl_xor_sync(0xffffffffu, b1, o));
        b2 = fmaxf(b2, __shfl_xor_sync(0xffffffffu, b2, o));
    }
    if ((threadIdx.x & 31) == 0) {  // non-negative (or NaN / inf, which order above): the bits order like the values
        atomicMax(reinterpret_cast<unsigned int*>(bmax) + 2 * (p / BU_BASIS), __float_as_uint(b1));
        atomicMax(reinterpret_cast<unsigned int*>(bmax) + 2 * (p / BU_BASIS) + 1, __float_as_uint(b2));
    }
}

__global__ void __launch_bounds__(BU_THREADS, BU_CTAS_PER_SM)
bps_nn_umma_kernel(const float* __restrict__ basis, int P, const unsigned char* __restrict__ apack, const float* __restrict__ bmax,
                   const float* __restrict__ pts, const int64_t* __restrict__ offsets, int64_t uniform_n, int64_t total_points, int nsplit,
                   unsigned long long* __restrict__ keys, long long* dbg) {
    extern __shared__ unsigned char bu_smem_raw[];
#ifdef DEXGG_DEBUG
#define BU_CLK(i) do { if (dbg && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) dbg[i] = clock64(); } while (0)
#else
#define BU_CLK(i) do { } while (0)
#endif
    if (threadIdx.x == 0) BU_CLK(0);
    unsigned char* smem = bu_smem_raw + ((1024u - (smem_u32(bu_smem_raw) & 1023u)) & 1023u);  // keeps the shared address space
    unsigned char* sA = smem;
    unsigned char* sB = sA + BU_A_BYTES;
    float* raw = reinterpret_cast<float*>(sB + 2 * BU_B_BYTES);
    uint32_t* s_pack = reinterpret_cast<uint32_t*>(raw + BU_RING * BU_RAW_FLOATS);  // [BU_BASIS] last two candidate chunks
    unsigned char* s_cnt = reinterpret_cast<unsigned char*>(s_pack + BU_BASIS);      // [BU_BASIS] candidate count (saturating)
    __shared__ uint64_t raw_full[BU_RING], b_full[2], b_empty[2], acc_full[BU_NACC], acc_empty[BU_NACC], a_full;
    __shared__ uint32_t s_tmem, s_cmax;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int b = blockIdx.y;
    const int64_t o0 = offsets ? offsets[b] : (int64_t)b * uniform_n;
    const int64_t o1 = offsets ? offsets[b + 1] : o0 + uniform_n;
    const int64_t n = o1 - o0;
    const int64_t len = (((n + nsplit - 1) / nsplit) + 3) & ~(int64_t)3;
    const int64_t r0 = o0 + (int64_t)blockIdx.z * len;
    const int64_t r1 = (r0 + len < o1) ? r0 + len : o1;
    if (r0 >= r1) return;  // uniform for the CTA
    const int64_t g0 = r0 & ~(int64_t)3;  // local point 0 (48-byte aligned in HBM)
    const int64_t span = r1 - g0;
    const int ntiles = (int)((span + BU_TILE - 1) / BU_TILE);
    const int64_t bulk_end = total_points & ~(int64_t)3;
    const int pblock = blockIdx.x * BU_BASIS;

    if (tid == 0) {
        for (int i = 0; i < BU_RING; ++i) mbar_init(&raw_full[i], 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&b_full[i], BU_CONV_WARPS);
            mbar_init(&b_empty[i], 1);
        }
        for (int i = 0; i < BU_NACC; ++i) {
            mbar_init(&acc_full[i], 1);
            mbar_init(&acc_empty[i], 128);
        }
        mbar_init(&a_full, 1);
        mbar_fence_init();
        s_cmax = 0u;
    }
    if (warp == BU_MMA_WARP) tmem_alloc(&s_tmem, BU_TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    auto tile_cnt = [&](int t) {
        const int64_t c = span - (int64_t)t * BU_TILE;
        return (int)(c < BU_TILE ? c : BU_TILE);
    };
    auto issue_raw = [&](int t) {  // one thread
        const int slot = t % BU_RING;
        const int64_t gs = g0 + (int64_t)t * BU_TILE;
        int64_t ge4 = gs + ((tile_cnt(t) + 3) & ~3);
        if (ge4 > bulk_end) ge4 = bulk_end;
        if (ge4 > gs) {
            const uint32_t bytes = (uint32_t)((ge4 - gs) * 12);
            fence_proxy_async();
            mbar_expect_tx(&raw_full[slot], bytes);
            bulk_g2s(raw + slot * BU_RAW_FLOATS, pts + 3 * gs, bytes, &raw_full[slot]);
        } else {
            mbar_arrive(&raw_full[slot]);
        }
    };
    if (warp == BU_SCAN_WARPS && lane == 0) {
        for (int t = 0; t < ntiles && t < BU_RING; ++t) issue_raw(t);
        mbar_expect_tx(&a_full, BU_A_BYTES);  // this CTA's packed basis rows, one bulk copy
        bulk_g2s(sA, apack + (size_t)pblock * 64, BU_A_BYTES, &a_full);
    }

    // ---- C = max |coordinate| over the range: by the scan warps among themselves (the margin is theirs), while the converters
    // and the MMA issuer already work on the first tile ----
    if (warp < BU_SCAN_WARPS) {
        float c = 0.f;
        for (int64_t i = 3 * r0 + tid; i < 3 * r1; i += 32 * BU_SCAN_WARPS) c = fmaxf(c, fabsf(pts[i]));
        for (int o = 16; o; o >>= 1) c = fmaxf(c, __shfl_xor_sync(0xffffffffu, c, o));
        if (lane == 0) atomicMax(&s_cmax, __float_as_uint(c));  // c >= 0 (or NaN / inf, which order above)
        asm volatile("bar.sync 2, %0;" ::"n"(32 * BU_SCAN_WARPS) : "memory");
    }
    const uint32_t tmem = s_tmem;
    if (threadIdx.x == 0) BU_CLK(1);

    if (warp >= BU_SCAN_WARPS && warp < BU_MMA_WARP) {
        // ================================ converters: raw AoS tile -> B operand ================================
        const int cw = warp - BU_SCAN_WARPS;
        for (int t = 0; t < ntiles; ++t) {
            const int slot = t % BU_RING, buf = t & 1;
            float* st = raw + slot * BU_RAW_FLOATS;
            unsigned char* dst = sB + buf * BU_B_BYTES;
            mbar_wait_backoff(&raw_full[slot], (t / BU_RING) & 1, 100);
            mbar_wait_backoff(&b_empty[buf], ((t >> 1) & 1) ^ 1, 200);  // a tile ahead of the scan: no hurry
            const int64_t gs = g0 + (int64_t)t * BU_TILE;
            const int cnt = tile_cnt(t);
            const int rows = ((cnt + BU_NBLK - 1) / BU_NBLK) * BU_NBLK;  // whole N blocks in use
            const int lo_valid = (t == 0) ? (int)(r0 - g0) : 0;
            const int bulk_local_end = (int)((bulk_end - gs) < 0 ? 0 : ((bulk_end - gs) > cnt ? cnt : (bulk_end - gs)));
            for (int j = cw * 32 + lane; j < rows; j += 32 * BU_CONV_WARPS) {
                uint32_t e[32];
                const bool valid = j >= lo_valid && j < cnt;
                float x = 0.f, y = 0.f, z = 0.f;
                if (valid) {
                    if (j < bulk_local_end) {
                        x = st[3 * j]; y = st[3 * j + 1]; z = st[3 * j + 2];
                    } else {  // the <= 3 points past the last 48-byte group of the whole buffer
                        const float* gp = pts + 3 * (gs + j);
                        x = gp[0]; y = gp[1]; z = gp[2];
                        st[3 * j] = x; st[3 * j + 1] = y; st[3 * j + 2] = z;
                    }
                }
                const float c3[3] = {-2.f * x, -2.f * y, -2.f * z};
#pragma unroll
                for (int cc = 0; cc < 3; ++cc) {
                    uint32_t parts[3];
                    split3(c3[cc], parts);
#pragma unroll
                    for (int i = 0; i < 3; ++i)
#pragma unroll
                        for (int jj = 0; jj < 3; ++jj) e[9 * cc + 3 * i + jj] = parts[i];
                }
                if (valid) {
                    const float w = __fmaf_rn(z, z, __fmaf_rn(y, y, __fmul_rn(x, x)));
                    uint32_t parts[3];
                    split3(w, parts);
                    const bool fin = w < __int_as_float(0x7f800000);  // false for inf / NaN: the column is then inf / NaN as a whole
                    e[27] = parts[0]; e[28] = fin ? parts[1] : 0u; e[29] = fin ? parts[2] : 0u;
                } else {  // masked: a() = +inf for every basis point
                    e[27] = 0x7f80u; e[28] = 0u; e[29] = 0u;
                }
                e[30] = e[31] = 0u;
                store_row_sw64(dst, j, e);
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                if (cw == 0) BU_CLK(8 + t);
                mbar_arrive(&b_full[buf]);
            }
            if (t + BU_RING < ntiles) {  // the slot is re-filled with the tile BU_RING ahead once every converter is done with it
                asm volatile("bar.sync 1, %0;" ::"n"(32 * BU_CONV_WARPS) : "memory");
                if (cw == 0 && lane == 0) issue_raw(t + BU_RING);
            }
        }
    } else if (warp == BU_MMA_WARP) {
        // ================================ MMA issuer ================================
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(128, BU_NBLK);
            const uint32_t sA_u = smem_u32(sA), sB_u = smem_u32(sB);
            uint32_t use = 0u;  // units issued per group so far (the groups advance in step)
            mbar_wait_backoff(&a_full, 0, 100);
            for (int t = 0; t < ntiles; ++t) {
                const int buf = t & 1;
                mbar_wait_backoff(&b_full[buf], (t >> 1) & 1, 100);
                tc_fence_after();
                const int nbk = (tile_cnt(t) + BU_NBLK - 1) / BU_NBLK;
                for (int nb = 0; nb < nbk; ++nb) {
                    const uint64_t bd = umma_desc_sw64_kmajor(sB_u + buf * BU_B_BYTES + nb * (BU_NBLK * 64));
                    for (int j = 0; j < BU_RJ; ++j, ++use) {
                        for (int gg = 0; gg < BU_GROUPS; ++gg) {  // group gg's next accumulator <- M block GROUPS * j + gg
                            const uint32_t a = (uint32_t)gg * BU_SLOTS + use % BU_SLOTS;
                            mbar_wait(&acc_empty[a], ((use / BU_SLOTS) & 1u) ^ 1u);
                            tc_fence_after();
                            const uint64_t ad = umma_desc_sw64_kmajor(sA_u + (j * BU_GROUPS + gg) * (128 * 64));
                            umma_bf16(tmem + a * BU_NBLK, ad, bd, idesc, 0u);
                            umma_bf16(tmem + a * BU_NBLK, ad + 2, bd + 2, idesc, 1u);
                            umma_commit(&acc_full[a]);
                        }
                    }
                }
                umma_commit(&b_empty[buf]);
            }
        }
        __syncwarp();
    }

    // ================================ epilogue groups: scan a() out of TMEM ================================
    const int g = warp >> 2;              // scan group (warps 0..15)
    const int l = (warp & 3) * 32 + lane;  // TMEM lane = row of the M block
    float best[BU_RJ];
    uint32_t last2[BU_RJ], ncand[BU_RJ];  // [15:0] last candidate chunk, [31:16] the one before | candidates since the last clear
#pragma unroll
    for (int j = 0; j < BU_RJ; ++j) {
        best[j] = __int_as_float(0x7f800000);
        last2[j] = 0u;
        ncand[j] = 0u;
    }
    if (warp < BU_SCAN_WARPS) {
        float margin;
        {
            const float C = __uint_as_float(s_cmax), u = 5.9604645e-8f;
            const float b1max = bmax[2 * blockIdx.x], b2max = bmax[2 * blockIdx.x + 1];
            const float E = 8.f * u * (3.f * C * C + 2.f * b1max * C);
            const float dm = b2max + 1.7320508f * C;
            margin = 1.5f * (2.f * E + 9.f * u * dm * dm);  // NaN / inf coordinates give a NaN / inf margin: every chunk then
        }                                                    // overflows into the exact fallback scan
        const uint32_t tlane = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        uint32_t nunit = 0;
        for (int t = 0; t < ntiles; ++t) {
            const int cnt = tile_cnt(t);
            const int nbk = (cnt + BU_NBLK - 1) / BU_NBLK;
            for (int nb = 0; nb < nbk; ++nb) {
                const int left = cnt - nb * BU_NBLK;
                const int nch = left >= BU_NBLK ? BU_NBLK / BU_CHUNK : (left + BU_CHUNK - 1) / BU_CHUNK;
                const uint32_t gc0 = (uint32_t)(t * BU_CPT + nb * (BU_NBLK / BU_CHUNK));
#pragma unroll
                for (int j = 0; j < BU_RJ; ++j) {
                    const uint32_t a = (uint32_t)g * BU_SLOTS + nunit % BU_SLOTS;
                    mbar_wait(&acc_full[a], (nunit / BU_SLOTS) & 1u);
                    ++nunit;
                    tc_fence_after();
                    const uint32_t taddr = tlane + a * BU_NBLK;
                    auto release = [&]() {  // every value of this accumulator is in registers
                        tc_fence_before();
                        mbar_arrive(&acc_empty[a]);
                    };
                    auto rule = [&](float v, uint32_t gc) {
                        // v = +inf (masked columns only) may be recorded while m is still +inf: a finite chunk clears it later, and a
                        // range without a finite distance finds no valid point in the exact pass
                        const float m = best[j];
                        const bool cand = !(v > m + margin);  // a NaN margin keeps everything
                        const bool clear = v < m - margin;    // a clear improvement drops the older candidates
                        if (cand) last2[j] = (last2[j] << 16) + gc;
                        const uint32_t kept = clear ? 0u : ncand[j];
                        ncand[j] = cand ? kept + 1u : ncand[j];
                        best[j] = fminf(m, v);
                    };
                    // one chunk per tcgen05.ld; the other warps of the SMSP (other groups / the other CTA of the SM) cover its latency.
                    // (Measured and dropped: two 16-column half-chunk loads in flight alternately - the extra loads and waits cost more
                    // than the latency they hide - and two 64-column accumulators per group - more hand-offs per point.)
                    uint32_t ra[32];
                    if (nch == BU_NBLK / BU_CHUNK) {  // whole accumulator: straight-line
#pragma unroll
                        for (int c = 0; c < BU_NBLK / BU_CHUNK; ++c) {
                            tmem_ld_32x32(taddr + (uint32_t)c * 32u, ra);
                            tmem_ld_wait();
                            if (c + 1 == BU_NBLK / BU_CHUNK) release();
                            rule(min32(ra), gc0 + c);
                        }
                    } else {
                        for (int c = 0; c < nch; ++c) {
                            tmem_ld_32x32(taddr + (uint32_t)c * 32u, ra);
                            tmem_ld_wait();
                            if (c + 1 == nch) release();
                            rule(min32(ra), gc0 + c);
                        }
                    }
                }
            }
            if (tid == 0) BU_CLK(16 + t);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (tid == 0) BU_CLK(2);
    if (warp == BU_MMA_WARP) {
        tc_fence_after();
        tmem_dealloc(tmem, BU_TMEM_COLS);
    }

    // ---- exact arg-min on the candidates (reference arithmetic, lowest index on ties), as in bps_nn_filter_kernel ----
    if (warp < BU_SCAN_WARPS) {
#pragma unroll
        for (int j = 0; j < BU_RJ; ++j) {
            s_pack[(BU_GROUPS * j + g) * 128 + l] = last2[j];
            s_cnt[(BU_GROUPS * j + g) * 128 + l] = (unsigned char)min(ncand[j], 255u);
        }
    }
    __syncthreads();
    const int total_chunks = (int)((span + BU_CHUNK - 1) / BU_CHUNK);
    // per-lane best (distance bits, local point index) over chunk `ch` read from HBM / L2; strict < keeps the lower index on a tie
    auto lane_chunk = [&](int ch, float qx, float qy, float qz, uint32_t& bd, uint32_t& bi) {
        const int64_t gp = g0 + (int64_t)ch * BU_CHUNK + lane;
        const bool ok = gp >= r0 && gp < r1;
        const float* p3 = pts + 3 * (ok ? gp : r0);
        const float df = bu_d2(qx, qy, qz, p3[0], p3[1], p3[2]);
        const uint32_t d = (ok && !(df != df)) ? __float_as_uint(df) : 0xffffffffu;  // d2 >= 0: bits order like values
        if (d < bd) {
            bd = d;
            bi = (uint32_t)(gp - g0);
        }
    };
    auto commit = [&](int p, uint32_t bd, uint32_t bi) {  // warp reduction of the per-lane bests, then the atomic merge
        const uint32_t mn = __reduce_min_sync(0xffffffffu, bd);
        const uint32_t im = __reduce_min_sync(0xffffffffu, bd == mn ? bi : 0xffffffffu);
        if (lane == 0 && mn != 0xffffffffu) atomicMin(&keys[(size_t)b * P + p], bu_key(__uint_as_float(mn), (uint32_t)(g0 + im - o0)));
    };
    auto one = [&](int e) {  // the whole warp finishes basis point pblock + e
        const int p = pblock + e;
        if (e >= BU_BASIS || p >= P) return;
        const uint32_t pk = s_pack[e];
        const uint32_t cn = s_cnt[e];
        if (cn == 0u) return;  // no finite distance in this range
        const float qx = basis[3 * p], qy = basis[3 * p + 1], qz = basis[3 * p + 2];
        uint32_t bd = 0xffffffffu, bi = 0xffffffffu;
        if (cn <= 2u) {
            if (cn == 2u) lane_chunk((int)(pk >> 16), qx, qy, qz, bd, bi);  // the earlier chunk first
            lane_chunk((int)(pk & 0xffffu), qx, qy, qz, bd, bi);
        } else {
            for (int ch = 0; ch < total_chunks; ++ch) lane_chunk(ch, qx, qy, qz, bd, bi);
        }
        commit(p, bd, bi);
    };
    if (ntiles <= BU_RING) {
        // The range is still intact in the raw ring: every epilogue THREAD finishes its own basis points from shared memory.
        // Lane L visits a chunk's points in the order L, L+1, ... (mod 32): chunks are 96 floats apart, so unrotated lanes would
        // all hit the same bank.  The running minimum is the 64-bit (d2 bits, local index) key: lowest index on a tie whatever
        // the visiting order.
        if (warp >= BU_SCAN_WARPS) return;
        const uint32_t lo = (uint32_t)(r0 - g0), nvalid = (uint32_t)(r1 - r0);
#pragma unroll 1
        for (int j = 0; j < BU_RJ; ++j) {
            const int e = (BU_GROUPS * j + g) * 128 + l;
            const int p = pblock + e;
            const uint32_t pk = s_pack[e];
            const uint32_t cn = p < P ? (uint32_t)s_cnt[e] : 0u;
            if (cn == 1u || cn == 2u) {
                const float qx = basis[3 * p], qy = basis[3 * p + 1], qz = basis[3 * p + 2];
                unsigned long long bk = ~0ull;
                auto thread_chunk = [&](uint32_t ch) {
                    const uint32_t base = ch * BU_CHUNK;  // a chunk never straddles two ring slots
                    const float* sp = raw + (base / BU_TILE) * BU_RAW_FLOATS + 3 * (base % BU_TILE);
#pragma unroll 8
                    for (int i = 0; i < BU_CHUNK; ++i) {
                        const uint32_t jj = (uint32_t)(lane + i) & (BU_CHUNK - 1);
                        const float df = bu_d2(qx, qy, qz, sp[3 * jj], sp[3 * jj + 1], sp[3 * jj + 2]);
                        const uint32_t li = base + jj;
                        const uint32_t d = (li - lo < nvalid) ? __float_as_uint(df) : 0xffffffffu;
                        const unsigned long long k = ((unsigned long long)d << 32) | li;
                        bk = k < bk ? k : bk;
                    }
                };
                if (cn == 2u) thread_chunk(pk >> 16);
                thread_chunk(pk & 0xffffu);
                // d2 >= 0, so its bits order like its value; bits above +inf are a NaN distance (or "no valid point"): no result
                if ((uint32_t)(bk >> 32) <= 0x7f800000u)
                    atomicMin(&keys[(size_t)b * P + p], (bk & 0xffffffff00000000ull) | (uint32_t)(g0 + (bk & 0xffffffffu) - o0));
            }
            if (tid == 0) BU_CLK(3 + j);
            // more than two candidate chunks (near-ties): the whole-range scan of that basis point is done by the warp together
            for (unsigned ov = __ballot_sync(0xffffffffu, cn > 2u); ov; ov &= ov - 1) one((BU_GROUPS * j + g) * 128 + (l & ~31) + (__ffs(ov) - 1));
        }
        return;
    }
    for (int e = warp; e < BU_BASIS; e += BU_THREADS / 32) one(e);
}

int bps_umma_basis_per_cta() { return BU_BASIS; }
int bps_umma_ctas_per_sm() { return BU_CTAS_PER_SM; }

size_t bps_umma_extra_bytes(int P) {  // packed A operand + per-block norms, behind the keys in the caller's workspace
    const size_t blocks = ((size_t)P + BU_BASIS - 1) / BU_BASIS;
    return align_up(blocks * BU_BASIS * 64, 256) + align_up(blocks * 2 * sizeof(float), 256);
}

int bps_umma_launch(const float* basis, int P, const float* pts, const int64_t* offsets, int B, int64_t uniform_n, int64_t total,
                    int nsplit, unsigned long long* keys, void* extra, cudaStream_t st) {
    static std::atomic<unsigned long long> attr_mask{0};
    DEXGG_SMEM_OPT_IN(bps_nn_umma_kernel, BU_SMEM, attr_mask);
    const int blocks = (P + BU_BASIS - 1) / BU_BASIS, rows = blocks * BU_BASIS;
    unsigned char* apack = static_cast<unsigned char*>(extra);
    float* bmax = reinterpret_cast<float*>(apack + align_up((size_t)rows * 64, 256));
    DEXGG_CUDA_OK(cudaMemsetAsync(bmax, 0, (size_t)blocks * 2 * sizeof(float), st));
    bps_basis_pack_kernel<<<rows / 256, 256, 0, st>>>(basis, P, rows, apack, bmax);
    dim3 grid(blocks, B, nsplit);
    long long* dbg = nullptr;
#ifdef DEXGG_DEBUG
    if (getenv("DEXGG_BPS_CLOCKS")) {
        cudaMalloc(&dbg, 128 * 8);
        cudaMemset(dbg, 0, 128 * 8);
    }
#endif
    bps_nn_umma_kernel<<<grid, BU_THREADS, BU_SMEM, st>>>(basis, P, apack, bmax, pts, offsets, uniform_n, total, nsplit, keys, dbg);
    ::dexgg::count_launches(2);
#ifdef DEXGG_DEBUG
    if (dbg) {
        long long h[128];
        cudaMemcpy(h, dbg, sizeof(h), cudaMemcpyDeviceToHost);
        cudaFree(dbg);
        fprintf(stderr, "bps_umma clocks (CTA 0): prologue %lld | converter tile ends", h[1] - h[0]);
        for (int i = 0; i < 8; ++i) if (h[8 + i]) fprintf(stderr, " %lld", h[8 + i] - h[0]);
        fprintf(stderr, " | scan tile ends");
        for (int i = 0; i < 8; ++i) if (h[16 + i]) fprintf(stderr, " %lld", h[16 + i] - h[0]);
        fprintf(stderr, " | main loop end %lld | recheck per j", h[2] - h[0]);
        for (int i = 0; i < 4; ++i) fprintf(stderr, " %lld", h[3 + i] - h[0]);
        fprintf(stderr, "\n");
    }
#endif
    DEXGG_LAUNCH_OK("bps_nn_umma_kernel");
    return DEXGG_OK;
}

}  // namespace dexgg
