// Fused whole-network tcgen05 kernel for the DexGenerator / DexEvaluator forwards (eval-mode BatchNorm, hoisted BPS
// projection).  See chain.cuh for the interface and DESIGN.md §4 for the roofline.
//
// One persistent CTA per SM, 64 + 128 x CH_PARTS threads, one 128-grasp tile at a time:
//   warp 0      weight producer: walks the pre-swizzled weight tile stream (built once by chain_pack) with 1-D bulk TMA
//               copies into a 4-stage x 16 KB shared-memory ring (mbarrier expect_tx / complete_tx)
//   warp 1      MMA issuer: one thread issues tcgen05.mma.kind::f16 (M=128, N=256, K=16); A = activation k-blocks resident
//               in shared memory (9 slots x [128 rows x 64 bf16], 128B swizzle), B = ring stage, D = TMEM
//   warps 2..   epilogue (4 x CH_PARTS warps): tcgen05.ld -> per-object base + BatchNorm affine + LeakyReLU -> bf16 -> next layer's A operand in
//               shared memory (or packed back into TMEM with tcgen05.st for the residual branch)
// Per ResBlock (ResBlock.forward, networks.py:43-56), with X the 512-wide input (slots 0..7) and t the tail k-block (slot 8):
//   P1  D[0,512)   = [X|t] . W3^T                         fc3 (skip branch), 2 x N=256 per k-block
//   E1  R = bf16(lrelu(D + base3)) packed two per column into TMEM [0,128) and [384,512)
//   P2  D[128,384) = [X|t] . W1^T                         fc1, N=256
//   E2  H = bf16(lrelu(D + base1)) -> slots 0..3 (X is dead now; bn1 is folded into W1 and base1)
//   P3  D[128,384) = H . W2[0:256]^T ;  E3  Xout[:, 0:256]  = bf16(lrelu(R + D + sh2)) -> slots 4..7 (bn2 scale folded into W2)
//   P4  same for W2[256:512]                        ;  E4  Xout[:, 256:512] -> slots 0..3 (H is dead now)
// so every block leaves X columns 0..255 in slots 4..7 and columns 256..511 in slots 0..3; the stream is packed in that
// k-block order.  Heads: one N=32 (generator) / N=16 (evaluator, weight as hi+lo rows) MMA chain over the 8 X slots, then
// rot6d -> SO(3) + palm offset (utils.py:380-394, 19-48) or the sigmoid, straight from TMEM.
#include "chain.cuh"
#include "umma.cuh"

namespace dexgg {

#define CH_PARTS 4  // epilogue warps per TMEM lane quarter (four per SMSP); 2 software-pipelined ones (round 1) measured the same
constexpr int CH_SLOT_BYTES = 16384, CH_NSLOT = 9, CH_STAGES = 4, CH_EPI_THREADS = 128 * CH_PARTS, CH_THREADS = 64 + CH_EPI_THREADS, CH_HEAD_TILES = 8;
constexpr int CH_PARAM_FLOATS = 512;             // per block: sh2
constexpr int CH_BASE_ROWS = 512, CH_BASE_BYTES = CH_BASE_ROWS * 32;  // hoisted per-object base of one phase as a B operand: [rows x 16 bf16]
constexpr int CH_A_BYTES = CH_NSLOT * CH_SLOT_BYTES;
// 227 KB per CTA: the alignment slack is what is left (the dynamic window starts 1024-aligned when the kernel has no static shared
// memory; the kernel traps if the slack does not cover the offset it finds)
constexpr int CH_BAR_BYTES = 128;
constexpr int CH_SMEM_USED = CH_A_BYTES + CH_STAGES * CH_SLOT_BYTES + CH_BASE_BYTES + CH_BAR_BYTES + CH_PARAM_FLOATS * 4;
constexpr int CH_SMEM_BYTES = 227 * 1024;
static_assert(CH_SMEM_USED <= CH_SMEM_BYTES, "shared memory budget");
constexpr int CH_TAIL_SLOT = 8;
#ifdef DEXGG_DEBUG
constexpr bool CH_DEBUG = true;   // the DEXGG_CHAIN_NOSTREAM / DEXGG_CHAIN_CLOCKS experiments of DESIGN.md §4.2
#else
constexpr bool CH_DEBUG = false;  // release: the debug branches fold away
#endif

__host__ __device__ inline int ch_block_tiles(int has_x) { return has_x ? 70 : 22; }  // fc3 4/kb + fc1 2/kb + fc2 16
__host__ __device__ inline int ch_head_rows(int kind) { return kind == DEXGG_NET_GENERATOR ? 32 : 16; }

// ---------------------------------------------------------------- weight stream packing
struct ChTile {
    const float* src;   // source matrix (fp32, row-major)
    const float* rowscale;  // optional per-source-row scale folded in before the bf16 rounding (BatchNorm affine)
    int ld, row0, nrows;  // rows [row0, row0 + nrows) are real, the rest of the tile is zero
    int col0;           // first source column
    int mode;           // 0: 64 plain columns; 1: tail [hi16|hi16|lo16|0]; 2: score head (row 0 = hi, row 8 = lo);
                        // 3: discriminator tail [w nt | w nt | 0] against inputs [hi nt | lo nt | one-hot | 0], nt <= 24
    int nt;             // mode 3: tail inputs
    int tile_rows;      // 128, 32 or 16
    uint32_t pad;
    uint64_t dst_off;   // byte offset of the tile in the stream
};

__global__ void chain_pack_kernel(const ChTile* __restrict__ tiles, unsigned char* __restrict__ stream) {
    const ChTile t = tiles[blockIdx.x];
    // pad = 0: [tile_rows x 64] tile, 128-byte swizzle.  pad = 1 + h: K half h of a [256 x 64] tile as a [256 x 32] tile under the
    // 64-byte swizzle (main layers: a ring stage is then one K half of an N = 256 operand and is consumed - and released - on
    // its own, instead of waiting for the partner stage that holds the other 128 rows)
    const int kw = t.pad ? 32 : 64, c_off = t.pad ? 32 * ((int)t.pad - 1) : 0;
    for (int e = threadIdx.x; e < t.tile_rows * kw; e += blockDim.x) {
        const int r = e / kw, cl = e % kw, c = cl + c_off;
        float v = 0.f;
        bool lo = false;
        if (t.mode == 0) {
            if (r < t.nrows) v = t.src[(int64_t)(t.row0 + r) * t.ld + t.col0 + c];
        } else if (t.mode == 1) {
            if (r < t.nrows && c < 48) {
                v = t.src[(int64_t)(t.row0 + r) * t.ld + t.col0 + (c & 15)];
                lo = c >= 32;
            }
        } else if (t.mode == 3) {
            if (r < t.nrows && c < 2 * t.nt) v = t.src[(int64_t)(t.row0 + r) * t.ld + t.col0 + (c % t.nt)];
        } else {
            if (r == 0 || r == 8) {
                v = t.src[t.col0 + c];
                lo = r == 8;
            }
        }
        if (t.rowscale && r < t.nrows) v *= t.rowscale[t.row0 + r];
        __nv_bfloat16 hi = __float2bfloat16_rn(v);
        if (lo) hi = __float2bfloat16_rn(v - __bfloat162float(hi));
        const uint64_t off = t.pad ? t.dst_off + (uint64_t)r * 64 + (uint64_t)((((cl >> 3) ^ ((r >> 1) & 3)) << 4) + ((cl & 7) << 1))
                                   : t.dst_off + (uint64_t)r * 128 + (uint64_t)((((c >> 3) ^ (r & 7)) << 4) + ((c & 7) << 1));
        *reinterpret_cast<__nv_bfloat16*>(stream + off) = hi;
    }
}

// ---------------------------------------------------------------- TMEM helpers (16-column variants, store)
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}
// LeakyReLU(0.2) on a packed bf16 pair, max(h, 0.2 h) as HMUL2.BF16 + HMNMX2.BF16: two instructions per pair after the fp32 sum
// is rounded to bf16, against FMUL2 + 2 FMNMX before it (for h < 0 the product is rounded to bf16 a second time; measured
// generator 2.05 -> 1.99 ms, evaluator 3.07 -> 2.98 ms, same filtered grasp set)
__device__ __forceinline__ uint32_t lrelu_bf16x2(uint32_t h) {
    uint32_t t, o;
    asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(t) : "r"(h), "r"(0x3e4d3e4du));  // bf16(0.2) = 0x3e4d
    asm("max.bf16x2 %0, %1, %2;" : "=r"(o) : "r"(h), "r"(t));
    return o;
}
// bf16x2( leaky_relu(bf16(acc + add)) ) for two adjacent columns: packed fp32x2 add (FADD2), one F2FP, then the bf16x2 LeakyReLU
__device__ __forceinline__ uint32_t lrelu_pack2(uint32_t a0, uint32_t a1, float b0, float b1) {
    const f32x2 x = add2(pack2(__uint_as_float(a0), __uint_as_float(a1)), pack2(b0, b1));
    float x0, x1;
    unpack2(x, x0, x1);
    return lrelu_bf16x2(pack_bf16x2(x0, x1));
}
// the same without an addend (the per-object base is accumulated by the tensor core)
__device__ __forceinline__ uint32_t lrelu_cvt2(uint32_t a0, uint32_t a1) {
    return lrelu_bf16x2(pack_bf16x2(__uint_as_float(a0), __uint_as_float(a1)));
}
// accumulator + bf16x2-packed residual only (the shift is accumulated by the tensor core)
__device__ __forceinline__ uint32_t lrelu_res2(uint32_t a0, uint32_t a1, uint32_t r2) {
    const f32x2 x = add2(pack2(__uint_as_float(a0), __uint_as_float(a1)), pack2(__uint_as_float(r2 << 16), __uint_as_float(r2 & 0xffff0000u)));
    float x0, x1;
    unpack2(x, x0, x1);
    return lrelu_bf16x2(pack_bf16x2(x0, x1));
}
// v = p0 + p1 + p2 (bf16 each) up to 2^-24 |v|
__device__ __forceinline__ void ch_split3(float v, uint32_t (&p)[3]) {
    const __nv_bfloat16 h0 = __float2bfloat16_rn(v);
    const float r1 = v - __bfloat162float(h0);
    const __nv_bfloat16 h1 = __float2bfloat16_rn(r1);
    const __nv_bfloat16 h2 = __float2bfloat16_rn(r1 - __bfloat162float(h1));
    p[0] = __bfloat16_as_ushort(h0); p[1] = __bfloat16_as_ushort(h1); p[2] = __bfloat16_as_ushort(h2);
}
// same with a third, bf16x2-packed addend (the residual branch)
__device__ __forceinline__ uint32_t lrelu_pack3(uint32_t a0, uint32_t a1, float b0, float b1, uint32_t r2) {
    f32x2 x = add2(pack2(__uint_as_float(a0), __uint_as_float(a1)), pack2(b0, b1));
    x = add2(x, pack2(__uint_as_float(r2 << 16), __uint_as_float(r2 & 0xffff0000u)));
    float x0, x1;
    unpack2(x, x0, x1);
    return lrelu_bf16x2(pack_bf16x2(x0, x1));
}
// 32 consecutive bf16 columns, already packed in pairs, into an A-operand slot (4 swizzle chunks from chunk `ci0`)
__device__ __forceinline__ void store_row32p(unsigned char* slot, int row, int ci0, const uint32_t (&o)[16]) {
#pragma unroll
    for (int q = 0; q < 4; ++q)
        *reinterpret_cast<uint4*>(slot + row * 128 + (((ci0 + q) ^ (row & 7)) << 4)) = make_uint4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
}
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
// 32 consecutive bf16 columns (= 4 swizzle chunks starting at chunk `ci0`) of row `row` into an A-operand slot
__device__ __forceinline__ void store_row32(unsigned char* slot, int row, int ci0, const float (&v)[32]) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint4 u = make_uint4(pack_bf16x2(v[8 * q], v[8 * q + 1]), pack_bf16x2(v[8 * q + 2], v[8 * q + 3]),
                                   pack_bf16x2(v[8 * q + 4], v[8 * q + 5]), pack_bf16x2(v[8 * q + 6], v[8 * q + 7]));
        *reinterpret_cast<uint4*>(slot + row * 128 + (((ci0 + q) ^ (row & 7)) << 4)) = u;
    }
}

}  // namespace dexgg

namespace dexgg {

// ---------------------------------------------------------------- the kernel
template <int KIND>
__global__ void __launch_bounds__(CH_THREADS, 1) chain_umma_kernel(const ChainArgs a) {
    extern __shared__ unsigned char ch_smem_raw[];
    // 1024-byte alignment by an OFFSET into the __shared__ array (a pointer rebuilt from an integer loses its address space and
    // every access to it becomes a generic LD / ST with 64-bit address arithmetic instead of LDS / STS)
    unsigned char* smem = ch_smem_raw + ((1024u - (smem_u32(ch_smem_raw) & 1023u)) & 1023u);
    unsigned char* sA = smem;                 // 9 activation k-block slots
    unsigned char* sR = smem + CH_A_BYTES;    // weight ring
    unsigned char* sBase = sR + CH_STAGES * CH_SLOT_BYTES;  // base B-operand tile (32-byte swizzle)
    uint64_t* bars = reinterpret_cast<uint64_t*>(sBase + CH_BASE_BYTES);
    uint64_t* full = bars;                    // [STAGES]
    uint64_t* empty = bars + CH_STAGES;       // [STAGES]
    uint64_t* mma_bar = bars + 2 * CH_STAGES; // MMA phase complete -> epilogue
    uint64_t* epi_bar = mma_bar + 1;          // epilogue phase complete (8 warps) -> MMA issuer
    uint64_t* base_bar = epi_bar + 1;         // base tile written (all epilogue warps) -> MMA issuer
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(base_bar + 1);
    float* s_par = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(bars) + CH_BAR_BYTES);  // [CH_PARAM_FLOATS]
    if ((int)(smem - ch_smem_raw) + CH_SMEM_USED > CH_SMEM_BYTES) __trap();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntiles = (int)((a.M + 127) / 128);
    constexpr int HEAD_ROWS = KIND == DEXGG_NET_GENERATOR ? 32 : 16;
    constexpr uint32_t HEAD_BYTES = HEAD_ROWS * 128;
    int n_main = 0;
    for (int b = 0; b < a.n_blocks; ++b) n_main += ch_block_tiles(a.has_x[b]);

    if (threadIdx.x == 0) {
        for (int s = 0; s < CH_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_init(mma_bar, 1);
        mbar_init(epi_bar, 4 * CH_PARTS);
        mbar_init(base_bar, 4 * CH_PARTS);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== weight producer =====================
        if (lane == 0) {
            int stage = 0, phase = 0;
            const unsigned char* src = reinterpret_cast<const unsigned char*>(a.stream);
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
                size_t off = 0;
                for (int i = 0; i < n_main; ++i) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    if (CH_DEBUG && a.debug_no_stream) { mbar_arrive(&full[stage]); } else {
                    mbar_expect_tx(&full[stage], CH_SLOT_BYTES);
                    bulk_g2s(sR + stage * CH_SLOT_BYTES, src + off, CH_SLOT_BYTES, &full[stage]); }
                    off += CH_SLOT_BYTES;
                    if (++stage == CH_STAGES) { stage = 0; phase ^= 1; }
                }
                // All eight head tiles in ONE copy (a copy costs ~140 cycles + bytes / 74 whatever its size), into a whole even /
                // odd stage pair (n_main is even; the evaluator's 16 KB fill only the first stage).  (Issuing it one pair earlier,
                // ahead of the last fc2 pair, was measured slower: it delays that pair and the head MMAs are not data-bound.)
                for (int j = 0; j < 2; ++j) mbar_wait(&empty[stage + j], phase ^ 1);
                if (CH_DEBUG && a.debug_no_stream) { mbar_arrive(&full[stage]); } else {
                mbar_expect_tx(&full[stage], CH_HEAD_TILES * HEAD_BYTES);
                bulk_g2s(sR + stage * CH_SLOT_BYTES, src + off, CH_HEAD_TILES * HEAD_BYTES, &full[stage]); }
                mbar_arrive(&full[stage + 1]);  // the pair's second barrier completes its phase too (parities stay in step)
                stage += 2;
                if (stage == CH_STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc256 = umma_idesc_bf16(128, 256), idesc_h = umma_idesc_bf16(128, HEAD_ROWS);
            const uint32_t sA_u = smem_u32(sA), sR_u = smem_u32(sR);
            int stage = 0, phase = 0, ephase = 0, head_stage = 0, head_phase = 0;
            int dbg_n = 0;
            auto dbg = [&]() {
                if (CH_DEBUG && a.debug_clock && blockIdx.x == 0 && dbg_n < 256) a.debug_clock[dbg_n++] = clock64();
            };
            auto wait_epi = [&]() {
                mbar_wait(epi_bar, ephase);
                ephase ^= 1;
                tc_fence_after();
                dbg();
            };
            auto mma_tile = [&](uint32_t d_col, int slot, uint32_t id, bool first) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                const uint64_t ad = umma_desc_sw128_kmajor(sA_u + slot * CH_SLOT_BYTES);
                const uint64_t bd = umma_desc_sw128_kmajor(sR_u + stage * CH_SLOT_BYTES);
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_bf16(tmem_base + d_col, ad + 2 * k, bd + 2 * k, id, !(first && k == 0));
                umma_commit(&empty[stage]);
                if (++stage == CH_STAGES) { stage = 0; phase ^= 1; }
            };
            // one [256 x 64] B tile = two consecutive ring stages, each ONE K HALF ([256 x 32], 64-byte swizzle) of it: N = 256
            // halves the shared-memory operand traffic per MAC against N = 128, and every stage is consumed and released on its
            // own (two K = 16 MMAs) as soon as it lands
            auto mma_tile2 = [&](uint32_t d_col, int slot, bool first) {
#pragma unroll
                for (int hk = 0; hk < 2; ++hk) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint64_t ad = umma_desc_sw128_kmajor(sA_u + slot * CH_SLOT_BYTES) + 4 * hk;
                    const uint64_t bd = umma_desc_sw64_kmajor(sR_u + stage * CH_SLOT_BYTES);
#pragma unroll
                    for (int k = 0; k < 2; ++k) umma_bf16(tmem_base + d_col, ad + 2 * k, bd + 2 * k, idesc256, !(first && hk == 0 && k == 0));
                    umma_commit(&empty[stage]);
                    if (++stage == CH_STAGES) { stage = 0; phase ^= 1; }
                }
            };
            // The per-object base of fc3 / fc1 (the hoisted BPS projection, fp32) is accumulated by the tensor core itself: the tail
            // k-block carries one-hot columns 48..53 "row belongs to the tile's first / second object" (three each), and the epilogue
            // warps write base[obj0 | obj1] of the phase's output columns as three bf16 parts each into a [rows x 16] B tile: one
            // K = 16 MMA per N = 256 adds it exactly to 2^-24 - no base loads, no fp32 adds in the epilogue.  The bn2 shift of fc2
            // goes the same way (its three parts in the slots of BOTH objects: the two one-hot groups sum to one).
            const uint64_t ad_ind = umma_desc_sw128_kmajor(sA_u + CH_TAIL_SLOT * CH_SLOT_BYTES) + 6;
            const uint32_t sB_u = smem_u32(sBase);
            int bphase = 0;
            auto base_mma = [&](bool staged, bool fresh, int ngroups, int row0, uint32_t d_col) {
                if (fresh) {  // a new tile was written for this phase
                    mbar_wait(base_bar, bphase);
                    bphase ^= 1;
                    tc_fence_after();
                }
                if (staged)
                    for (int q = 0; q < ngroups; ++q)
                        umma_bf16(tmem_base + d_col + q * 256, ad_ind, umma_desc_sw32_kmajor(sB_u + (row0 + q * 256) * 32), idesc256, 1u);
            };
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
                const int64_t m_last = ((int64_t)t * 128 + 127 < a.M) ? (int64_t)t * 128 + 127 : a.M - 1;
                const bool staged = (m_last / a.rows_per_obj) - ((int64_t)t * 128) / a.rows_per_obj <= 1;  // at most two objects
                for (int b = 0; b < a.n_blocks; ++b) {
                    const int nkb = a.has_x[b] ? 9 : 1, s0 = a.has_x[b] ? 0 : CH_TAIL_SLOT;
                    wait_epi();  // P1: fc3 -> D[0,512)
                    for (int i = 0; i < nkb; ++i)
                        for (int q = 0; q < 2; ++q) mma_tile2(q * 256, s0 + i, i == 0);
                    base_mma(staged, true, 2, 0, 0);
                    umma_commit(mma_bar); dbg();
                    wait_epi();  // P2: fc1 -> D[128,384)
                    for (int i = 0; i < nkb; ++i) mma_tile2(128, s0 + i, i == 0);
                    base_mma(staged, true, 1, 0, 128);
                    umma_commit(mma_bar); dbg();
                    for (int h = 0; h < 2; ++h) {  // P3 / P4: fc2 halves -> D[128,384), A = H in slots 0..3
                        wait_epi();
                        for (int kb = 0; kb < 4; ++kb) mma_tile2(128, kb, kb == 0);
                        base_mma(staged, h == 0, 1, h * 256, 128);  // + sh2 (the bn2 shift): same tile form, both one-hot slots
                        umma_commit(mma_bar); dbg();
                    }
                }
                wait_epi();  // heads over the 8 X slots: the 8 weight tiles arrive as one copy in the next stage pair
                head_stage = stage;
                head_phase = phase;
                stage += 2;
                if (stage == CH_STAGES) { stage = 0; phase ^= 1; }
                mbar_wait(&full[head_stage], head_phase);
                tc_fence_after();
                for (int s = 0; s < CH_HEAD_TILES; ++s) {
                    const uint64_t ad = umma_desc_sw128_kmajor(sA_u + s * CH_SLOT_BYTES);
                    const uint64_t bd = umma_desc_sw128_kmajor(sR_u + head_stage * CH_SLOT_BYTES + s * HEAD_BYTES);
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(tmem_base + 128, ad + 2 * k, bd + 2 * k, idesc_h, !(s == 0 && k == 0));
                }
                umma_commit(&empty[head_stage]);
                umma_commit(&empty[head_stage + 1]);
                umma_commit(mma_bar); dbg();
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int quarter = warp & 3, part = (warp - 2) >> 2;  // TMEM lanes 32 * (warp % 4) ..; `part`: which share of a phase's columns
        const int half = part >> 1;  // which 256-column half of the fc3 accumulator
        const int row = quarter * 32 + lane;
        const int etid = threadIdx.x - 64;  // 0..255 among the epilogue threads
        const uint32_t tlane = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const int64_t ldbase = 768 * (int64_t)a.n_blocks;
        float* s_sh2 = s_par;                // [512]
        int mphase = 0;
        int dbg_n = 0;
        auto dbg = [&]() {
            if (CH_DEBUG && a.debug_clock && blockIdx.x == 0 && warp == 2 && lane == 0 && dbg_n < 256) a.debug_clock[256 + dbg_n++] = clock64();
        };
        auto wait_mma = [&]() {
            mbar_wait(mma_bar, mphase);
            mphase ^= 1;
            tc_fence_after();
            dbg();
        };
        auto arrive_epi = [&]() {
            dbg();
            tc_fence_before();
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(epi_bar);
        };
        auto epi_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(CH_EPI_THREADS) : "memory"); };
        // ---- tail k-block (slot 8) of tile `tt`: per-row inputs as [hi16 | lo16 | hi16 | 0].  Done by the half-1 warps: for the first
        // tile before the loop, afterwards during the head phase of the previous tile (slot 8 is free then, the half-0 warps are
        // busy with the head epilogue), so the global loads of the next tile's inputs are off the critical path ----
        auto pack_tail = [&](int tt) {
            {
                const int64_t m = (int64_t)tt * 128 + row;
                const int64_t mm = m < a.M ? m : a.M - 1;
                const int64_t rel = mm / a.rows_per_obj - ((int64_t)tt * 128) / a.rows_per_obj;
                const float ind0 = rel == 0 ? 1.f : 0.f, ind1 = rel == 1 ? 1.f : 0.f;
                unsigned char* slot = sA + CH_TAIL_SLOT * CH_SLOT_BYTES;
                if (KIND == DEXGG_NET_DISCRIMINATOR) {
                    // 24 inputs (rot 9 | transl 3 | joints) as [hi 24 | lo 24 | one-hot 6 | 0]: the weights of these columns are plain bf16
                    float v[24];
#pragma unroll
                    for (int j = 0; j < 24; ++j) v[j] = 0.f;
#pragma unroll
                    for (int j = 0; j < 9; ++j) v[j] = fmaf(__ldg(a.tail_a + j), a.in_rot[mm * 9 + j], __ldg(a.tail_c + j));
#pragma unroll
                    for (int j = 0; j < 3; ++j) v[9 + j] = fmaf(__ldg(a.tail_a + 9 + j), a.in_transl[mm * 3 + j], __ldg(a.tail_c + 9 + j));
#pragma unroll
                    for (int j = 0; j < 12; ++j)
                        if (j < a.nJ) v[12 + j] = fmaf(__ldg(a.tail_a + 12 + j), a.in_joint[mm * a.nJ + j], __ldg(a.tail_c + 12 + j));
                    float w[32];
#pragma unroll
                    for (int j = 0; j < 24; ++j) w[j] = __bfloat162float(__float2bfloat16_rn(v[j]));
#pragma unroll
                    for (int j = 0; j < 8; ++j) w[24 + j] = v[j] - w[j];
                    store_row32(slot, row, 0, w);  // chunks 0..3: hi 24 | lo 0..7
                    float w2[32];
#pragma unroll
                    for (int j = 0; j < 16; ++j) w2[j] = v[8 + j] - w[8 + j];
                    w2[16] = w2[17] = w2[18] = ind0;
                    w2[19] = w2[20] = w2[21] = ind1;
#pragma unroll
                    for (int j = 22; j < 32; ++j) w2[j] = 0.f;
                    store_row32(slot, row, 4, w2);  // chunks 4..7: lo 8..23 | one-hot | 0
                    return;
                }
                float v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = 0.f;
                if (KIND == DEXGG_NET_GENERATOR) {
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        if (j < a.zD) v[j] = a.z[mm * a.zD + j];
                } else {
#pragma unroll
                    for (int j = 0; j < 9; ++j) v[j] = fmaf(__ldg(a.tail_a + j), a.in_rot[mm * 9 + j], __ldg(a.tail_c + j));
#pragma unroll
                    for (int j = 0; j < 3; ++j) v[9 + j] = fmaf(__ldg(a.tail_a + 9 + j), a.in_transl[mm * 3 + j], __ldg(a.tail_c + 9 + j));
                }
                float w[32];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float hi = __bfloat162float(__float2bfloat16_rn(v[j]));
                    w[j] = hi;
                    w[16 + j] = v[j] - hi;
                }
                store_row32(slot, row, 0, w);  // chunks 0..3: hi | lo
#pragma unroll
                for (int j = 0; j < 16; ++j) { w[16 + j] = 0.f; }
                // columns 48..53: one-hot "this row belongs to the tile's first / second object" (three each: the base MMA's A operand)
                w[16] = w[17] = w[18] = ind0;
                w[19] = w[20] = w[21] = ind1;
                store_row32(slot, row, 4, w);  // chunks 4..7: hi | one-hot | 0
            }
        };
        if (part == CH_PARTS - 1 && (int)blockIdx.x < ntiles) pack_tail((int)blockIdx.x);
        arrive_epi();
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
            const int64_t m = (int64_t)t * 128 + row;
            const int64_t mm = m < a.M ? m : a.M - 1;
            const int64_t obj = mm / a.rows_per_obj;
            const int64_t obj0 = ((int64_t)t * 128) / a.rows_per_obj;
            const int64_t m_last = ((int64_t)t * 128 + 127 < a.M) ? (int64_t)t * 128 + 127 : a.M - 1;
            const bool staged = (m_last / a.rows_per_obj) - obj0 <= 1;  // the tile touches at most two objects
            const int64_t obj1 = obj0 + 1 < a.n_obj ? obj0 + 1 : obj0;
            const float* base_obj = a.base + obj * ldbase;
            for (int b = 0; b < a.n_blocks; ++b) {
                // ---- stage sh2 and the fc3 base tile of this block (while the fc3 MMAs run; the base MMAs are the last of the phase).
                // Row n of a base tile = base[obj0][col0 + n] and base[obj1][col0 + n] as three bf16 parts each, K = 16 (32 bytes),
                // 32-byte swizzle.  The previous user of the buffer (the fc1 base MMA of the previous block) completed before this
                // warp left that block's E2. ----
                auto write_base_tile = [&](int n, int col0) {
                    const float v0 = __ldg(col0 < 0 ? a.sh2[b] + n : a.base + obj0 * ldbase + b * 768 + col0 + n);
                    const float v1 = __ldg(col0 < 0 ? a.sh2[b] + n : a.base + obj1 * ldbase + b * 768 + col0 + n);
                    uint32_t p0[3], p1[3];
                    ch_split3(v0, p0);
                    ch_split3(v1, p1);
                    const int sw = (n >> 2) & 1;
                    *reinterpret_cast<uint4*>(sBase + n * 32 + (sw << 4)) = make_uint4(p0[0] | (p0[1] << 16), p0[2] | (p1[0] << 16), p1[1] | (p1[2] << 16), 0u);
                    *reinterpret_cast<uint4*>(sBase + n * 32 + ((sw ^ 1) << 4)) = make_uint4(0u, 0u, 0u, 0u);
                };
                auto arrive_base = [&]() {
                    fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(base_bar);
                };
                epi_sync();  // every epilogue warp is done with the previous block's parameters
                if (etid < 128) reinterpret_cast<float4*>(s_sh2)[etid] = __ldg(reinterpret_cast<const float4*>(a.sh2[b]) + etid);
                if (staged) write_base_tile(etid, 256);  // fc3 columns: one of the 512 rows per epilogue thread
                arrive_base();
                epi_sync();
                const float* base_g = base_obj + b * 768;  // tiles over more than two objects: the base is added here, per row, from global
                // Four warps per TMEM lane quarter, i.e. four epilogue warps per SMSP: a warp's tcgen05.ld latency is covered by the
                // other three (plain load - wait - convert loops, 1/4 of a phase's columns each).
                // ---- E1: skip branch R = bf16(lrelu(fc3 + base)), packed in place into TMEM: [0,256) upwards into [0,128) by parts
                // 0 / 1, [256,512) DOWNWARDS into [384,512) by parts 2 / 3, [128,384) left free as one 256-column block.  The two warps
                // of a pair alternate 32-column chunks (step s: chunks 2s and 2s + 1 in walking order); the packed chunk c lands in
                // the columns of chunk c / 2, which one of the two has read in a step <= s: a 64-thread barrier between the loads and
                // the stores of a step keeps every store behind every load it could overwrite. ----
                wait_mma();
                if (staged && etid < 256) write_base_tile(etid, 0);  // fc1 columns (the fc3 base MMAs are complete: the buffer is free)
                arrive_base();
                {
                    const int sub = part & 1;
                    uint32_t ra[32];
#pragma unroll
                    for (int st = 0; st < 4; ++st) {
                        const int cc = 2 * st + sub, c = half ? 7 - cc : cc;
                        tmem_ld_32x32(tlane + half * 256 + c * 32, ra);
                        tmem_ld_wait();
                        // the fc1 MMAs only need D[128,384) to have been READ: arrive as soon as this warp's last load has landed
                        if (st == 3) arrive_epi();
                        asm volatile("bar.sync %0, 64;" ::"r"(2 + quarter * 2 + half) : "memory");
                        uint32_t o[16];
                        if (staged) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) o[j] = lrelu_cvt2(ra[2 * j], ra[2 * j + 1]);
                        } else {
                            const float* bp = base_g + 256 + half * 256 + c * 32;
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const float4 bv = *reinterpret_cast<const float4*>(bp + 4 * j);
                                o[2 * j] = lrelu_pack2(ra[4 * j], ra[4 * j + 1], bv.x, bv.y);
                                o[2 * j + 1] = lrelu_pack2(ra[4 * j + 2], ra[4 * j + 3], bv.z, bv.w);
                            }
                        }
                        tmem_st_32x16(tlane + (half ? 384 : 0) + c * 16, o);
                    }
                }
                tmem_st_wait();
                // ---- E2: H = bf16(lrelu(fc1 + base)) -> slots 0..3 (part p owns H columns [64p, 64p + 64) = slot p) ----
                wait_mma();
                if (staged) write_base_tile(etid, -1);  // sh2 of both fc2 halves (the fc1 base MMA is complete: the buffer is free)
                arrive_base();
                {
                    uint32_t ra[32];
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        tmem_ld_32x32(tlane + 128 + part * 64 + c * 32, ra);
                        tmem_ld_wait();
                        uint32_t o[16];
                        if (staged) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) o[j] = lrelu_cvt2(ra[2 * j], ra[2 * j + 1]);
                        } else {
                            const float* bp = base_g + part * 64 + c * 32;
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const float4 bv = *reinterpret_cast<const float4*>(bp + 4 * j);
                                o[2 * j] = lrelu_pack2(ra[4 * j], ra[4 * j + 1], bv.x, bv.y);
                                o[2 * j + 1] = lrelu_pack2(ra[4 * j + 2], ra[4 * j + 3], bv.z, bv.w);
                            }
                        }
                        store_row32p(sA + part * CH_SLOT_BYTES, row, c * 4, o);
                    }
                }
                arrive_epi();
                // ---- E3 / E4: Xout = bf16(lrelu(R + fc2 + sh2)) -> slots 4..7 (h = 0) / 0..3 (h = 1) ----
                for (int h = 0; h < 2; ++h) {
                    wait_mma();
                    uint32_t ra[32], qa[16];
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        tmem_ld_32x32(tlane + 128 + part * 64 + c * 32, ra);
                        tmem_ld_32x16(tlane + (h ? 384 : 0) + part * 32 + c * 16, qa);
                        tmem_ld_wait();
                        // the second fc2 half reads H (slots 0..3) and only needs D[128,384) read: start it under this warp's last chunk
                        if (h == 0 && c == 1) arrive_epi();
                        uint32_t o[16];
                        if (staged) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) o[j] = lrelu_res2(ra[2 * j], ra[2 * j + 1], qa[j]);
                        } else {
                            const float* hp = s_sh2 + h * 256 + part * 64 + c * 32;
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const float4 hv = *reinterpret_cast<const float4*>(hp + 4 * j);
                                o[2 * j] = lrelu_pack3(ra[4 * j], ra[4 * j + 1], hv.x, hv.y, qa[2 * j]);
                                o[2 * j + 1] = lrelu_pack3(ra[4 * j + 2], ra[4 * j + 3], hv.z, hv.w, qa[2 * j + 1]);
                            }
                        }
                        store_row32p(sA + ((h ? 0 : 4) + part) * CH_SLOT_BYTES, row, c * 4, o);
                    }
                    if (h == 1) arrive_epi();
                }
            }
            // ---- heads (part-0 warps) | next tile's tail pack (last part's warps) ----
            if (part == CH_PARTS - 1 && t + (int)gridDim.x < ntiles) pack_tail(t + (int)gridDim.x);
            wait_mma();
            if (part == 0) {
                if (KIND == DEXGG_NET_GENERATOR) {
                    uint32_t r[32];
                    tmem_ld_32x32(tlane + 128, r);
                    tmem_ld_wait();
                    if (m < a.M) {
                        const int nJ = a.nJ;
                        float hv[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) hv[j] = __uint_as_float(r[j]) + (j < a.nh ? __ldg(a.bh + j) : 0.f);
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (j < nJ) a.joint[m * nJ + j] = hv[j];
                        float g[9];
#pragma unroll
                        for (int j = 0; j < 32; ++j)  // static indexing keeps hv in registers
                            if (j >= nJ && j < nJ + 9) g[j - nJ] = hv[j];
                        const float a0 = g[0], a1 = g[1], a2 = g[2], b0 = g[3], b1 = g[4], b2 = g[5];
                        if (a.rot6d) {
                            float* r6 = a.rot6d + m * 6;
                            r6[0] = a0; r6[1] = a1; r6[2] = a2; r6[3] = b0; r6[4] = b1; r6[5] = b2;
                        }
                        const float na = fmaxf(sqrtf(a0 * a0 + a1 * a1 + a2 * a2), 1e-8f);
                        const float x0 = a0 / na, x1 = a1 / na, x2 = a2 / na;
                        float z0 = x1 * b2 - x2 * b1, z1 = x2 * b0 - x0 * b2, z2 = x0 * b1 - x1 * b0;
                        const float nz = fmaxf(sqrtf(z0 * z0 + z1 * z1 + z2 * z2), 1e-8f);
                        z0 /= nz; z1 /= nz; z2 /= nz;
                        const float y0 = z1 * x2 - z2 * x1, y1 = z2 * x0 - z0 * x2, y2 = z0 * x1 - z1 * x0;
                        float* ro = a.rot + m * 9;
                        ro[0] = x0; ro[1] = y0; ro[2] = z0;
                        ro[3] = x1; ro[4] = y1; ro[5] = z1;
                        ro[6] = x2; ro[7] = y2; ro[8] = z2;
                        const double off = (double)a.z_offset;
                        a.transl[m * 3 + 0] = (float)((double)g[6] - (double)x0 * off);
                        a.transl[m * 3 + 1] = (float)((double)g[7] - (double)x1 * off);
                        a.transl[m * 3 + 2] = (float)((double)g[8] - (double)x2 * off);
                    }
                } else {
                    uint32_t r[16];
                    tmem_ld_32x16(tlane + 128, r);
                    tmem_ld_wait();
                    if (m < a.M) {
                        const float s = __uint_as_float(r[0]) + __uint_as_float(r[8]) + __ldg(a.bh);
                        a.p[m] = 1.f / (1.f + expf(-s));
                    }
                }
            }
            arrive_epi();  // head accumulators read, next tail packed: the MMA issuer may start the next tile
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace dexgg

// ------------------------------------------------------------------------------------------------ host side
#include <stdlib.h>
#include <vector>

namespace dexgg {

static size_t ch_data_bytes(int kind, int n_blocks, const int* has_x) {
    size_t n = 0;
    for (int b = 0; b < n_blocks; ++b) n += (size_t)ch_block_tiles(has_x[b]) * CH_SLOT_BYTES;
    return n + (size_t)CH_HEAD_TILES * ch_head_rows(kind) * 128;
}
static int ch_total_tiles(int n_blocks, const int* has_x) {
    int n = CH_HEAD_TILES;
    for (int b = 0; b < n_blocks; ++b) n += ch_block_tiles(has_x[b]);
    return n;
}

size_t chain_stream_bytes(int kind, int n_blocks, const int* has_x) {
    return align_up(ch_data_bytes(kind, n_blocks, has_x), 256) + align_up((size_t)ch_total_tiles(n_blocks, has_x) * sizeof(ChTile), 256);
}

// X k-block i of a block's first layer -> source columns (slots 0..3 hold X[256:512], slots 4..7 hold X[0:256])
static inline int ch_xcol(int i) { return i < 4 ? 256 + 64 * i : 64 * (i - 4); }

int chain_pack(const ChainPackSrc& s, __nv_bfloat16* stream, cudaStream_t st) {
    const bool disc = s.kind == DEXGG_NET_DISCRIMINATOR;
    DEXGG_REQUIRE((disc ? (s.tailw == 32 && s.n_tail <= 24) : (s.tailw == 16 && (s.kind == DEXGG_NET_GENERATOR || s.kind == DEXGG_NET_EVALUATOR))) &&
                      s.n_blocks <= CH_MAX_BLOCKS, "chain_pack: unsupported network for the fused kernel");
    std::vector<ChTile> tiles;
    uint64_t off = 0;
    auto push = [&](const float* src, int ld, int row0, int nrows, int col0, int mode, int tile_rows, const float* rowscale = nullptr,
                    uint32_t khalf = 0) {
        ChTile t;
        t.rowscale = rowscale;
        t.src = src; t.ld = ld; t.row0 = row0; t.nrows = nrows; t.col0 = col0; t.mode = mode; t.tile_rows = tile_rows; t.pad = khalf;
        t.nt = 24;
        t.dst_off = off;
        off += (uint64_t)tile_rows * (khalf ? 64 : 128);
        tiles.push_back(t);
    };
    for (int b = 0; b < s.n_blocks; ++b) {
        const int nkb = s.has_x[b] ? 9 : 1, KT = s.KT[b];
        auto first_layer = [&](int row_base, int nq) {
            for (int i = 0; i < nkb; ++i) {
                const bool tail = (i == nkb - 1);
                const int col0 = tail ? (s.has_x[b] ? 512 : 0) : ch_xcol(i);
                for (int g = 0; g < nq / 2; ++g)
                    for (int hk = 0; hk < 2; ++hk) push(s.Wt[b], KT, row_base + g * 256, 256, col0, tail ? (disc ? 3 : 1) : 0, 256, s.sc[b], 1 + hk);
            }
        };
        first_layer(256, 4);  // fc3 rows 256..767
        first_layer(0, 2);    // fc1 rows 0..255
        for (int h = 0; h < 2; ++h)
            for (int kb = 0; kb < 4; ++kb)
                for (int hk = 0; hk < 2; ++hk) push(s.W2[b], 256, h * 256, 256, kb * 64, 0, 256, s.sc2[b], 1 + hk);
    }
    const int hr = ch_head_rows(s.kind);
    for (int i = 0; i < CH_HEAD_TILES; ++i) {
        if (s.kind == DEXGG_NET_GENERATOR) push(s.Wh, 512, 0, s.nh, ch_xcol(i), 0, hr);
        else push(s.Wh, 512, 0, 1, ch_xcol(i), 2, hr);  // evaluator / discriminator: one score row as hi + lo
    }
    const int has_x[CH_MAX_BLOCKS] = {s.has_x[0], s.has_x[1], s.has_x[2]};
    DEXGG_REQUIRE(off == ch_data_bytes(s.kind, s.n_blocks, has_x) && (int)tiles.size() == ch_total_tiles(s.n_blocks, has_x),
                  "chain_pack: internal size mismatch");
    unsigned char* base = reinterpret_cast<unsigned char*>(stream);
    ChTile* table = reinterpret_cast<ChTile*>(base + align_up((size_t)off, 256));
    DEXGG_CUDA_OK(cudaMemcpyAsync(table, tiles.data(), tiles.size() * sizeof(ChTile), cudaMemcpyHostToDevice, st));
    chain_pack_kernel<<<(unsigned)tiles.size(), 256, 0, st>>>(table, base); count_launches(1);
    DEXGG_LAUNCH_OK("chain_pack_kernel");
    return DEXGG_OK;
}

int chain_forward(const ChainArgs& a, cudaStream_t st) {
    DEXGG_REQUIRE(a.M >= 1 && a.stream && a.base && a.rows_per_obj >= 1, "chain_forward: bad arguments");
    static std::atomic<unsigned long long> attr_g{0}, attr_e{0}, attr_d{0};
    DEXGG_SMEM_OPT_IN(chain_umma_kernel<DEXGG_NET_GENERATOR>, CH_SMEM_BYTES, attr_g);
    DEXGG_SMEM_OPT_IN(chain_umma_kernel<DEXGG_NET_EVALUATOR>, CH_SMEM_BYTES, attr_e);
    DEXGG_SMEM_OPT_IN(chain_umma_kernel<DEXGG_NET_DISCRIMINATOR>, CH_SMEM_BYTES, attr_d);
    const int64_t ntiles = (a.M + 127) / 128;
#ifdef DEXGG_DEBUG  // timing experiments only (never in a release build): weight stream off / phase timeline of CTA 0
    { const char* e = getenv("DEXGG_CHAIN_NOSTREAM"); const_cast<ChainArgs&>(a).debug_no_stream = (e && e[0] == '1') ? 1 : 0; }
    static long long* dbg_buf = nullptr;
    if (getenv("DEXGG_CHAIN_CLOCKS")) {
        if (!dbg_buf) DEXGG_CUDA_OK(cudaMalloc(&dbg_buf, 512 * sizeof(long long)));
        DEXGG_CUDA_OK(cudaMemsetAsync(dbg_buf, 0, 512 * sizeof(long long), st));
        const_cast<ChainArgs&>(a).debug_clock = dbg_buf;
    }
#endif
    const int grid = (int)(ntiles < sm_count() ? ntiles : sm_count());
    if (a.kind == DEXGG_NET_GENERATOR) {
        DEXGG_REQUIRE(a.z && a.rot && a.transl && a.joint && a.zD >= 1 && a.zD <= 16 && a.nh <= 32 && a.nJ + 9 == a.nh, "chain_forward: bad generator arguments");
        chain_umma_kernel<DEXGG_NET_GENERATOR><<<grid, CH_THREADS, CH_SMEM_BYTES, st>>>(a);
    } else if (a.kind == DEXGG_NET_DISCRIMINATOR) {
        DEXGG_REQUIRE(a.in_rot && a.in_transl && a.in_joint && a.nJ >= 0 && a.nJ <= 12 && a.tail_a && a.tail_c && a.p, "chain_forward: bad discriminator arguments");
        chain_umma_kernel<DEXGG_NET_DISCRIMINATOR><<<grid, CH_THREADS, CH_SMEM_BYTES, st>>>(a);
    } else {
        DEXGG_REQUIRE(a.kind == DEXGG_NET_EVALUATOR && a.in_rot && a.in_transl && a.tail_a && a.tail_c && a.p, "chain_forward: bad evaluator arguments");
        chain_umma_kernel<DEXGG_NET_EVALUATOR><<<grid, CH_THREADS, CH_SMEM_BYTES, st>>>(a);
    }
    count_launches(1);
    DEXGG_LAUNCH_OK("chain_umma_kernel");
#ifdef DEXGG_DEBUG
    if (a.debug_clock) {
        std::vector<long long> h(512);
        DEXGG_CUDA_OK(cudaMemcpy(h.data(), dbg_buf, 512 * sizeof(long long), cudaMemcpyDeviceToHost));
        FILE* f = fopen(getenv("DEXGG_CHAIN_CLOCKS"), "a");
        if (f) {
            fprintf(f, "kind %d\n", a.kind);
            for (int r = 0; r < 2; ++r) {
                fprintf(f, r ? "epi:" : "mma:");
                for (int i = 0; i < 256 && h[r * 256 + i]; ++i) fprintf(f, " %lld", h[r * 256 + i] - h[0]);
                fprintf(f, "\n");
            }
            fclose(f);
        }
    }
#endif
    return DEXGG_OK;
}

}  // namespace dexgg
