// Stage 1 — Basis Point Set encoding on sm_100a.
//
// For every basis point b (P = 4096 in the reference) and every cloud: idx = argmin_k d2(x_k, b) with
//   d2 = fma(dz,dz, fma(dy,dy, dx*dx))  (fp32, direct differences, lowest index wins an exact tie)
// and dist = sqrt((dx^2 + dz^2) + dy^2) of the winner, the summation order that reproduces the reference's stored
// encodings bit-exactly (oracle/bps_oracle.c; reference call sites DexGanGrasp/data/bps_encoder.py:28-55).
//
// Kernel design (bps_nn_packed_kernel):
//   * one CTA = 128 threads x R basis points per thread, held in registers as packed f32x2 pairs;
//   * the cloud is streamed through shared memory in tiles of 1024 points by 1-D TMA bulk copies
//     (cp.async.bulk + mbarrier, double buffered) exactly as it lies in HBM (AoS, 12 B / point); three LDS.128
//     broadcast four points to the whole CTA;
//   * per (basis pair x point): FADD2 x3, FMUL2, FFMA2 x2 — the same six IEEE roundings per lane as the scalar
//     form, at half the issue slots — and the running minimum is kept with one FMNMX3 per two points (ALU pipe);
//   * the arg-min is recovered exactly: per 32-point chunk the CTA records which chunk first lowered the minimum
//     (strict <), and at the end one warp re-evaluates that chunk (lane = point), `ballot(d == best)` gives the
//     lowest matching index;
//   * clouds can be split over blockIdx.z; partial results merge through atomicMin on the 64-bit key
//     (d2 bits << 32 | index), which orders by distance first and by index second (d2 >= 0, so the float bit
//     pattern is monotonic) — the same tie rule.
//   * bps_finish_kernel gathers the winner and writes dist / idx.
// Roofline: FP32 pipe (6 lane-ops per pair); see DESIGN.md.
#include "common.cuh"

namespace dexgg {

constexpr int BPS_THREADS = 128;
constexpr int BPS_TILE = 1024;  // points per shared-memory stage
constexpr int BPS_CHUNK = 32;   // points per arg-min chunk (= one warp in the rescan)
constexpr int BPS_STAGE_FLOATS = 3 * (BPS_TILE + BPS_CHUNK);
#define BPS_SENTINEL 3.0e38f  // masked point: every distance to it is +inf

__device__ __forceinline__ float search_d2(float bx, float by, float bz, float x, float y, float z) {
    const float dx = __fsub_rn(x, bx), dy = __fsub_rn(y, by), dz = __fsub_rn(z, bz);
    return __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
}

__device__ __forceinline__ f32x2 search_d2x2(f32x2 bx, f32x2 by, f32x2 bz, float x, float y, float z) {
    const f32x2 dx = sub2(pack2(x, x), bx), dy = sub2(pack2(y, y), by), dz = sub2(pack2(z, z), bz);
    return fma2(dz, dz, fma2(dy, dy, mul2(dx, dx)));
}

__device__ __forceinline__ unsigned long long make_key(float d2, uint32_t idx) {
    return ((unsigned long long)__float_as_uint(d2) << 32) | idx;
}

template <int R>
__global__ void __launch_bounds__(BPS_THREADS)
bps_nn_packed_kernel(const float* __restrict__ basis, int P, const float* __restrict__ pts,
                     const int64_t* __restrict__ offsets, int64_t uniform_n, int64_t total_points, int nsplit,
                     unsigned long long* __restrict__ keys) {
    static_assert(R % 2 == 0, "R must be even");
    extern __shared__ __align__(128) unsigned char bps_smem[];
    float* stage0 = reinterpret_cast<float*>(bps_smem);
    float* stage1 = stage0 + BPS_STAGE_FLOATS;
    float* s_best = stage1 + BPS_STAGE_FLOATS;                 // [R][BPS_THREADS]
    int* s_chunk = reinterpret_cast<int*>(s_best + R * BPS_THREADS);  // [R][BPS_THREADS]
    __shared__ uint64_t bar[2];

    const int tid = threadIdx.x;
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
    const int ntiles = (int)((span + BPS_TILE - 1) / BPS_TILE);
    const int64_t bulk_end = total_points & ~(int64_t)3;

    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();

    auto issue = [&](int t) {  // tid 0 only
        float* dst = (t & 1) ? stage1 : stage0;
        const int64_t gs = g0 + (int64_t)t * BPS_TILE;
        int64_t cnt = span - (int64_t)t * BPS_TILE;
        if (cnt > BPS_TILE) cnt = BPS_TILE;
        int64_t ge4 = gs + ((cnt + 3) & ~(int64_t)3);
        if (ge4 > bulk_end) ge4 = bulk_end;
        if (ge4 > gs) {
            const uint32_t bytes = (uint32_t)((ge4 - gs) * 12);
            fence_proxy_async();
            mbar_expect_tx(&bar[t & 1], bytes);
            bulk_g2s(dst, pts + 3 * gs, bytes, &bar[t & 1]);
        } else {
            mbar_arrive(&bar[t & 1]);
        }
    };
    if (tid == 0) issue(0);

    // basis points of this thread: p(r) = blockIdx.x * (THREADS*R) + r*THREADS + tid, pairs (2j, 2j+1)
    const int pbase = blockIdx.x * (BPS_THREADS * R) + tid;
    f32x2 bx[R / 2], by[R / 2], bz[R / 2];
#pragma unroll
    for (int j = 0; j < R / 2; ++j) {
        float c[2][3];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int p = pbase + (2 * j + h) * BPS_THREADS;
            const bool ok = p < P;
            c[h][0] = ok ? basis[3 * p] : 0.f;
            c[h][1] = ok ? basis[3 * p + 1] : 0.f;
            c[h][2] = ok ? basis[3 * p + 2] : 0.f;
        }
        bx[j] = pack2(c[0][0], c[1][0]);
        by[j] = pack2(c[0][1], c[1][1]);
        bz[j] = pack2(c[0][2], c[1][2]);
    }
    float best[R];
    int bchunk[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        best[r] = __int_as_float(0x7f800000);
        bchunk[r] = -1;
    }

    for (int t = 0; t < ntiles; ++t) {
        if (tid == 0 && t + 1 < ntiles) issue(t + 1);
        float* st = (t & 1) ? stage1 : stage0;
        mbar_wait(&bar[t & 1], (t >> 1) & 1);
        // ---- fix-up: mask points outside [r0, r1), plain-load the <= 3 points past the last 48-byte group ----
        const int64_t gs = g0 + (int64_t)t * BPS_TILE;
        int cnt = (int)((span - (int64_t)t * BPS_TILE) < BPS_TILE ? (span - (int64_t)t * BPS_TILE) : BPS_TILE);
        const int npad = (cnt + BPS_CHUNK - 1) & ~(BPS_CHUNK - 1);
        const int lo_valid = (t == 0) ? (int)(r0 - g0) : 0;
        int bulk_local_end = (int)((bulk_end - gs) < 0 ? 0 : ((bulk_end - gs) > cnt ? cnt : (bulk_end - gs)));
        for (int j = tid; j < lo_valid; j += BPS_THREADS) {
            st[3 * j] = BPS_SENTINEL;
            st[3 * j + 1] = BPS_SENTINEL;
            st[3 * j + 2] = BPS_SENTINEL;
        }
        for (int j = bulk_local_end + tid; j < npad; j += BPS_THREADS) {
            float x = BPS_SENTINEL, y = BPS_SENTINEL, z = BPS_SENTINEL;
            if (j < cnt) {
                const float* g = pts + 3 * (gs + j);
                x = g[0];
                y = g[1];
                z = g[2];
            }
            st[3 * j] = x;
            st[3 * j + 1] = y;
            st[3 * j + 2] = z;
        }
        __syncthreads();

        const float4* s4 = reinterpret_cast<const float4*>(st);
        const int nchunks = npad / BPS_CHUNK;
        for (int c = 0; c < nchunks; ++c) {
            float cur[R];
#pragma unroll
            for (int r = 0; r < R; ++r) cur[r] = best[r];
#pragma unroll 2
            for (int q = 0; q < BPS_CHUNK / 4; ++q) {
                const float4 v0 = s4[(c * (BPS_CHUNK / 4) + q) * 3 + 0];
                const float4 v1 = s4[(c * (BPS_CHUNK / 4) + q) * 3 + 1];
                const float4 v2 = s4[(c * (BPS_CHUNK / 4) + q) * 3 + 2];
#pragma unroll
                for (int j = 0; j < R / 2; ++j) {
                    const f32x2 da = search_d2x2(bx[j], by[j], bz[j], v0.x, v0.y, v0.z);
                    const f32x2 db = search_d2x2(bx[j], by[j], bz[j], v0.w, v1.x, v1.y);
                    const f32x2 dc = search_d2x2(bx[j], by[j], bz[j], v1.z, v1.w, v2.x);
                    const f32x2 dd = search_d2x2(bx[j], by[j], bz[j], v2.y, v2.z, v2.w);
                    float a0, a1, b0, b1, c0, c1, d0, d1;
                    unpack2(da, a0, a1);
                    unpack2(db, b0, b1);
                    unpack2(dc, c0, c1);
                    unpack2(dd, d0, d1);
                    cur[2 * j] = min3(min3(cur[2 * j], a0, b0), c0, d0);
                    cur[2 * j + 1] = min3(min3(cur[2 * j + 1], a1, b1), c1, d1);
                }
            }
            const int gc = t * (BPS_TILE / BPS_CHUNK) + c;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (cur[r] < best[r]) {
                    best[r] = cur[r];
                    bchunk[r] = gc;
                }
            }
        }
        __syncthreads();
    }

    // ---- exact arg-min: one warp per basis point re-evaluates the winning 32-point chunk ----
#pragma unroll
    for (int r = 0; r < R; ++r) {
        s_best[r * BPS_THREADS + tid] = best[r];
        s_chunk[r * BPS_THREADS + tid] = bchunk[r];
    }
    __syncthreads();
    const int lane = tid & 31, warp = tid >> 5;
    const int pblock = blockIdx.x * (BPS_THREADS * R);
    for (int e = warp; e < BPS_THREADS * R; e += BPS_THREADS / 32) {
        const int p = pblock + e;  // e = r*THREADS + t  <->  p(r) of thread t
        if (p >= P) break;
        const int ch = s_chunk[e];
        if (ch < 0) continue;  // no finite distance in this range
        const float bd = s_best[e];
        const float qx = basis[3 * p], qy = basis[3 * p + 1], qz = basis[3 * p + 2];
        const int64_t g = g0 + (int64_t)ch * BPS_CHUNK + lane;
        float d = __int_as_float(0x7f800000);
        if (g >= r0 && g < r1) {
            const float* gp = pts + 3 * g;
            d = search_d2(qx, qy, qz, gp[0], gp[1], gp[2]);
        }
        const unsigned m = __ballot_sync(0xffffffffu, d == bd);
        if (lane == 0 && m != 0) {
            const int64_t gi = g0 + (int64_t)ch * BPS_CHUNK + (__ffs(m) - 1);
            atomicMin(&keys[(size_t)b * P + p], make_key(bd, (uint32_t)(gi - o0)));
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Filtered search (default): the scan runs on the EXPANDED form  a(k) = |x_k|^2 - 2 b.x_k  (= d2 - |b|^2, three FFMA per pair
// instead of the six FP32-pipe operations of the direct form) and only FILTERS; the reference arithmetic (direct differences,
// search_d2) is then evaluated on the few candidates, so indices and distances stay bit-exact.
//   * |a_fp32(k) - a_exact(k)| <= E = 8u (3C^2 + 2 |b|_1 C),  u = 2^-24, C = max |coordinate| of the CTA's point range
//     (three fma roundings on intermediates bounded by 3C^2 + 2|b|_1 C, plus the three roundings of |x|^2);
//   * the fp32 direct-form winner k* has exact d2 within (1 + 9u) of the exact minimum, so
//     a_fp32(k*) <= min_k a_fp32(k) + 2E + 9u (|b|_2 + sqrt(3) C)^2 =: M  (the kernel uses 1.5 M, one value per thread: the
//     maximum over its R basis points);
//   * online rule per chunk with value v = min a over the chunk and running minimum m: v < m - M clears the
//     candidate list (every earlier chunk is > final minimum + M), v <= m + M appends the chunk (chunks of BPSF_CHUNK points here).  The list keeps the last
//     two chunks and a count in one register; a count above two (near-ties in more than two chunks: practically never)
//     falls back to an exact scan of the whole range for that basis point;
//   * masked points carry (0,0,0) with |x|^2 = +inf, so they never win and never produce a NaN.
constexpr int BPSF_W_FLOATS = BPS_TILE + BPS_CHUNK;
constexpr int BPSF_CHUNK = 32;   // points per candidate chunk of the filter
constexpr int BPSF_SUBS = BPSF_CHUNK / 32;  // points per lane in the exact evaluation of a chunk

__device__ __forceinline__ f32x2 filter_a2(f32x2 nbx, f32x2 nby, f32x2 nbz, float x, float y, float z, float w) {
    return fma2(nbx, pack2(x, x), fma2(nby, pack2(y, y), fma2(nbz, pack2(z, z), pack2(w, w))));
}

template <int R>
__global__ void __launch_bounds__(BPS_THREADS, R >= 16 ? 3 : (R >= 8 ? 5 : 7))
bps_nn_filter_kernel(const float* __restrict__ basis, int P, const float* __restrict__ pts,
                     const int64_t* __restrict__ offsets, int64_t uniform_n, int64_t total_points, int nsplit,
                     unsigned long long* __restrict__ keys) {
    static_assert(R % 2 == 0, "R must be even");
    extern __shared__ __align__(128) unsigned char bps_smem[];
    float* stage0 = reinterpret_cast<float*>(bps_smem);
    float* stage1 = stage0 + BPS_STAGE_FLOATS;
    float* s_w = stage1 + BPS_STAGE_FLOATS;                              // [BPSF_W_FLOATS] |x|^2 of the current tile
    uint32_t* s_pack = reinterpret_cast<uint32_t*>(s_w + BPSF_W_FLOATS);  // [R][BPS_THREADS] candidate lists
    __shared__ uint64_t bar[2];
    __shared__ uint32_t s_cmax;

    const int tid = threadIdx.x;
    const int b = blockIdx.y;
    const int64_t o0 = offsets ? offsets[b] : (int64_t)b * uniform_n;
    const int64_t o1 = offsets ? offsets[b + 1] : o0 + uniform_n;
    const int64_t n = o1 - o0;
    const int64_t len = (((n + nsplit - 1) / nsplit) + 3) & ~(int64_t)3;
    const int64_t r0 = o0 + (int64_t)blockIdx.z * len;
    const int64_t r1 = (r0 + len < o1) ? r0 + len : o1;
    if (r0 >= r1) return;  // uniform for the CTA
    const int64_t g0 = r0 & ~(int64_t)3;
    const int64_t span = r1 - g0;
    const int ntiles = (int)((span + BPS_TILE - 1) / BPS_TILE);
    const int64_t bulk_end = total_points & ~(int64_t)3;

    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        mbar_fence_init();
        s_cmax = 0u;
    }
    __syncthreads();

    auto issue = [&](int t) {  // tid 0 only
        float* dst = (t & 1) ? stage1 : stage0;
        const int64_t gs = g0 + (int64_t)t * BPS_TILE;
        int64_t cnt = span - (int64_t)t * BPS_TILE;
        if (cnt > BPS_TILE) cnt = BPS_TILE;
        int64_t ge4 = gs + ((cnt + 3) & ~(int64_t)3);
        if (ge4 > bulk_end) ge4 = bulk_end;
        if (ge4 > gs) {
            const uint32_t bytes = (uint32_t)((ge4 - gs) * 12);
            fence_proxy_async();
            mbar_expect_tx(&bar[t & 1], bytes);
            bulk_g2s(dst, pts + 3 * gs, bytes, &bar[t & 1]);
        } else {
            mbar_arrive(&bar[t & 1]);
        }
    };
    if (tid == 0) issue(0);

    // ---- C = max |coordinate| over the range (the first tile's copy is in flight meanwhile) ----
    {
        float c = 0.f;
        for (int64_t i = 3 * r0 + tid; i < 3 * r1; i += BPS_THREADS) c = fmaxf(c, fabsf(pts[i]));
        for (int o = 16; o; o >>= 1) c = fmaxf(c, __shfl_xor_sync(0xffffffffu, c, o));
        if ((tid & 31) == 0) atomicMax(&s_cmax, __float_as_uint(c));  // c >= 0 (or NaN / inf, which order above)
    }

    const int pbase = blockIdx.x * (BPS_THREADS * R) + tid;
    f32x2 nbx[R / 2], nby[R / 2], nbz[R / 2];
    float b1max = 0.f, b2max = 0.f;
#pragma unroll
    for (int j = 0; j < R / 2; ++j) {
        float c[2][3];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int p = pbase + (2 * j + h) * BPS_THREADS;
            const bool ok = p < P;
            c[h][0] = ok ? basis[3 * p] : 0.f;
            c[h][1] = ok ? basis[3 * p + 1] : 0.f;
            c[h][2] = ok ? basis[3 * p + 2] : 0.f;
            b1max = fmaxf(b1max, fabsf(c[h][0]) + fabsf(c[h][1]) + fabsf(c[h][2]));
            b2max = fmaxf(b2max, sqrtf(c[h][0] * c[h][0] + c[h][1] * c[h][1] + c[h][2] * c[h][2]));
        }
        nbx[j] = pack2(-2.f * c[0][0], -2.f * c[1][0]);
        nby[j] = pack2(-2.f * c[0][1], -2.f * c[1][1]);
        nbz[j] = pack2(-2.f * c[0][2], -2.f * c[1][2]);
    }
    __syncthreads();
    float margin;
    {
        const float C = __uint_as_float(s_cmax), u = 5.9604645e-8f;
        const float E = 8.f * u * (3.f * C * C + 2.f * b1max * C);
        const float dm = b2max + 1.7320508f * C;
        margin = 1.5f * (2.f * E + 9.f * u * dm * dm);  // NaN / inf coordinates give a NaN / inf margin: every chunk then
    }                                                    // overflows into the exact fallback scan
    float best[R];
    uint32_t pack[R];  // [31:18] last candidate chunk, [17:4] the one before, [3:0] count (saturating)
#pragma unroll
    for (int r = 0; r < R; ++r) {
        best[r] = __int_as_float(0x7f800000);
        pack[r] = 0u;
    }

    for (int t = 0; t < ntiles; ++t) {
        if (tid == 0 && t + 1 < ntiles) issue(t + 1);
        float* st = (t & 1) ? stage1 : stage0;
        mbar_wait(&bar[t & 1], (t >> 1) & 1);
        const int64_t gs = g0 + (int64_t)t * BPS_TILE;
        int cnt = (int)((span - (int64_t)t * BPS_TILE) < BPS_TILE ? (span - (int64_t)t * BPS_TILE) : BPS_TILE);
        const int npad = (cnt + BPS_CHUNK - 1) & ~(BPS_CHUNK - 1);
        const int lo_valid = (t == 0) ? (int)(r0 - g0) : 0;
        int bulk_local_end = (int)((bulk_end - gs) < 0 ? 0 : ((bulk_end - gs) > cnt ? cnt : (bulk_end - gs)));
        // ---- fix-up + |x|^2: masked points become (0,0,0) with w = +inf; the <= 3 points past the last 48-byte group of the
        // whole buffer are plain-loaded ----
        for (int j = tid; j < npad; j += BPS_THREADS) {
            float x = 0.f, y = 0.f, z = 0.f, w = __int_as_float(0x7f800000);
            if (j >= lo_valid && j < cnt) {
                if (j < bulk_local_end) {
                    x = st[3 * j]; y = st[3 * j + 1]; z = st[3 * j + 2];
                } else {
                    const float* g = pts + 3 * (gs + j);
                    x = g[0]; y = g[1]; z = g[2];
                }
                w = __fmaf_rn(z, z, __fmaf_rn(y, y, __fmul_rn(x, x)));
            }
            if (j < lo_valid || j >= bulk_local_end) {
                st[3 * j] = x; st[3 * j + 1] = y; st[3 * j + 2] = z;
            }
            s_w[j] = w;
        }
        __syncthreads();

        const float4* s4 = reinterpret_cast<const float4*>(st);
        const float4* w4 = reinterpret_cast<const float4*>(s_w);
        const int nquads = npad / 4;
        const int nchunks = (npad + BPSF_CHUNK - 1) / BPSF_CHUNK;
        for (int c = 0; c < nchunks; ++c) {
            float cur[R];
#pragma unroll
            for (int r = 0; r < R; ++r) cur[r] = __int_as_float(0x7f800000);
            const int q1 = min((c + 1) * (BPSF_CHUNK / 4), nquads);
#pragma unroll 2
            for (int q = c * (BPSF_CHUNK / 4); q < q1; ++q) {
                const float4 v0 = s4[q * 3 + 0];
                const float4 v1 = s4[q * 3 + 1];
                const float4 v2 = s4[q * 3 + 2];
                const float4 ww = w4[q];
#pragma unroll
                for (int j = 0; j < R / 2; ++j) {
                    const f32x2 da = filter_a2(nbx[j], nby[j], nbz[j], v0.x, v0.y, v0.z, ww.x);
                    const f32x2 db = filter_a2(nbx[j], nby[j], nbz[j], v0.w, v1.x, v1.y, ww.y);
                    const f32x2 dc = filter_a2(nbx[j], nby[j], nbz[j], v1.z, v1.w, v2.x, ww.z);
                    const f32x2 dd = filter_a2(nbx[j], nby[j], nbz[j], v2.y, v2.z, v2.w, ww.w);
                    float a0, a1, b0, b1, c0, c1, d0, d1;
                    unpack2(da, a0, a1);
                    unpack2(db, b0, b1);
                    unpack2(dc, c0, c1);
                    unpack2(dd, d0, d1);
                    cur[2 * j] = min3(min3(cur[2 * j], a0, b0), c0, d0);
                    cur[2 * j + 1] = min3(min3(cur[2 * j + 1], a1, b1), c1, d1);
                }
            }
            const uint32_t gc = (uint32_t)(t * (BPS_TILE / BPSF_CHUNK) + c);
#pragma unroll
            for (int r = 0; r < R; ++r) {  // branch-free: selects only
                const float v = cur[r], m = best[r];
                const bool cand = (v < __int_as_float(0x7f800000)) && !(v > m + margin);  // a NaN margin keeps everything
                const uint32_t pk = pack[r];
                const uint32_t cn = (v < m - margin) ? 0u : (pk & 15u);  // a clear improvement drops the older candidates
                const uint32_t npk = (gc << 18) | ((pk >> 14) & 0x3fff0u) | min(cn + 1u, 15u);
                pack[r] = cand ? npk : pk;
                best[r] = fminf(m, v);
            }
        }
        __syncthreads();
    }

    // ---- exact arg-min on the candidates: one warp per basis point, reference arithmetic (search_d2), lowest index on ties.
    // A lane evaluates its points of a chunk with all loads in flight together, and two basis points are
    // handled per iteration, so the L2 latency of the candidate points is paid once per eight sub-chunks ----
#pragma unroll
    for (int r = 0; r < R; ++r) s_pack[r * BPS_THREADS + tid] = pack[r];
    __syncthreads();
    const int lane = tid & 31, warp = tid >> 5;
    const int pblock = blockIdx.x * (BPS_THREADS * R);
    const int total_chunks = (int)((span + BPSF_CHUNK - 1) / BPSF_CHUNK);
    const bool resident = ntiles <= 2;  // tile 0 is still in stage0 and tile 1 in stage1
    // per-lane best (distance bits, local point index) over chunk `ch`; strict < keeps the lower index on a tie
    auto lane_chunk = [&](int ch, float qx, float qy, float qz, uint32_t& bd, uint32_t& bi) {
        const int64_t gbase = g0 + (int64_t)ch * BPSF_CHUNK + lane;
        float px[BPSF_SUBS], py[BPSF_SUBS], pz[BPSF_SUBS];
#pragma unroll
        for (int sub = 0; sub < BPSF_SUBS; ++sub) {
            const int64_t g = gbase + sub * 32;
            const bool ok = g >= r0 && g < r1;
            const float* gp = pts + 3 * (ok ? g : r0);
            px[sub] = gp[0]; py[sub] = gp[1]; pz[sub] = gp[2];
        }
#pragma unroll
        for (int sub = 0; sub < BPSF_SUBS; ++sub) {
            const int64_t g = gbase + sub * 32;
            const float df = search_d2(qx, qy, qz, px[sub], py[sub], pz[sub]);
            const uint32_t d = (g >= r0 && g < r1 && !(df != df)) ? __float_as_uint(df) : 0xffffffffu;  // d2 >= 0: bits order like values
            if (d < bd) {
                bd = d;
                bi = (uint32_t)(g - g0);
            }
        }
    };
    auto commit = [&](int p, uint32_t bd, uint32_t bi) {  // warp reduction of the per-lane bests, then the atomic merge
        const uint32_t mn = __reduce_min_sync(0xffffffffu, bd);
        const uint32_t im = __reduce_min_sync(0xffffffffu, bd == mn ? bi : 0xffffffffu);
        if (lane == 0 && mn != 0xffffffffu)
            atomicMin(&keys[(size_t)b * P + p], make_key(__uint_as_float(mn), (uint32_t)(g0 + im - o0)));
    };
    auto one = [&](int e) {
        const int p = pblock + e;
        if (e >= BPS_THREADS * R || p >= P) return;
        const uint32_t pk = s_pack[e];
        const uint32_t cn = pk & 15u;
        if (cn == 0u) return;  // no finite distance in this range
        const float qx = basis[3 * p], qy = basis[3 * p + 1], qz = basis[3 * p + 2];
        uint32_t bd = 0xffffffffu, bi = 0xffffffffu;
        if (cn <= 2u) {
            if (cn == 2u) lane_chunk((int)((pk >> 4) & 0x3fffu), qx, qy, qz, bd, bi);  // the earlier chunk first
            lane_chunk((int)(pk >> 18), qx, qy, qz, bd, bi);
        } else {
            for (int ch = 0; ch < total_chunks; ++ch) lane_chunk(ch, qx, qy, qz, bd, bi);
        }
        commit(p, bd, bi);
    };
    if (resident) {
        // The range's (at most two) tiles are still intact in the stages (valid points are never rewritten), so every THREAD
        // finishes its own R basis points from shared memory: 32 points per candidate chunk, ~19 instructions per point and
        // 32 basis points per warp in flight, against ~50 warp instructions per basis point (L2 loads, two REDUX, an atomic
        // by one lane) for the warp-per-basis-point form below.  Lane L visits the chunk's points in the order L, L+1, ...
        // (mod 32): chunks are 96 floats apart, so unrotated lanes would all hit the same bank.  The running minimum is the
        // 64-bit (d2 bits, local index) key, which keeps the lowest index on a tie whatever the visiting order.
        const uint32_t lo = (uint32_t)(r0 - g0), nvalid = (uint32_t)(r1 - r0);
#pragma unroll 1
        for (int r = 0; r < R; ++r) {
            const int p = pbase + r * BPS_THREADS;
            const uint32_t pk = s_pack[r * BPS_THREADS + tid];
            const uint32_t cn = p < P ? (pk & 15u) : 0u;
            if (cn == 1u || cn == 2u) {
                const float qx = basis[3 * p], qy = basis[3 * p + 1], qz = basis[3 * p + 2];
                unsigned long long bk = ~0ull;
                auto thread_chunk = [&](uint32_t ch) {
                    const uint32_t base = ch * BPSF_CHUNK;  // a chunk never straddles the two stages (BPS_TILE % BPSF_CHUNK == 0)
                    const float* sp = ((base >= BPS_TILE) ? stage1 : stage0) + 3 * (base & (BPS_TILE - 1));
#pragma unroll 8
                    for (int i = 0; i < BPSF_CHUNK; ++i) {
                        const uint32_t j = (uint32_t)(lane + i) & (BPSF_CHUNK - 1);
                        const float df = search_d2(qx, qy, qz, sp[3 * j], sp[3 * j + 1], sp[3 * j + 2]);
                        const uint32_t li = base + j;
                        const uint32_t d = (li - lo < nvalid) ? __float_as_uint(df) : 0xffffffffu;
                        const unsigned long long k = ((unsigned long long)d << 32) | li;
                        bk = k < bk ? k : bk;
                    }
                };
                if (cn == 2u) thread_chunk((pk >> 4) & 0x3fffu);
                thread_chunk(pk >> 18);
                // d2 >= 0, so its bits order like its value; bits above +inf are a NaN distance (or "no valid point"): no result
                if ((uint32_t)(bk >> 32) <= 0x7f800000u)
                    atomicMin(&keys[(size_t)b * P + p], (bk & 0xffffffff00000000ull) | (uint32_t)(g0 + (bk & 0xffffffffu) - o0));
            }
            // more than two candidate chunks (near-ties): the whole-range scan of that basis point is done by the warp together
            for (unsigned ov = __ballot_sync(0xffffffffu, cn > 2u); ov; ov &= ov - 1)
                one(r * BPS_THREADS + (tid & ~31) + (__ffs(ov) - 1));
        }
        return;
    }
    for (int e = warp; e < BPS_THREADS * R; e += 2 * (BPS_THREADS / 32)) {
        const int e1 = e + BPS_THREADS / 32;
        const int p0 = pblock + e, p1 = pblock + e1;
        const uint32_t pk0 = p0 < P ? s_pack[e] : 0u, pk1 = (e1 < BPS_THREADS * R && p1 < P) ? s_pack[e1] : 0u;
        if ((pk0 & 15u) == 1u && (pk1 & 15u) == 1u) {  // common case: one candidate chunk each, eight loads in flight per lane
            const float ax = basis[3 * p0], ay = basis[3 * p0 + 1], az = basis[3 * p0 + 2];
            const float cx = basis[3 * p1], cy = basis[3 * p1 + 1], cz = basis[3 * p1 + 2];
            uint32_t d0 = 0xffffffffu, i0 = 0xffffffffu, d1 = 0xffffffffu, i1 = 0xffffffffu;
            lane_chunk((int)(pk0 >> 18), ax, ay, az, d0, i0);
            lane_chunk((int)(pk1 >> 18), cx, cy, cz, d1, i1);
            commit(p0, d0, i0);
            commit(p1, d1, i1);
        } else {
            one(e);
            one(e1);
        }
    }
}

// Simple scalar compare-select form: one thread per (cloud, basis point); cross-check of the packed kernel.
__global__ void __launch_bounds__(128)
bps_nn_scalar_kernel(const float* __restrict__ basis, int P, const float* __restrict__ pts,
                     const int64_t* __restrict__ offsets, int64_t uniform_n, unsigned long long* __restrict__ keys) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (p >= P) return;
    const int64_t o0 = offsets ? offsets[b] : (int64_t)b * uniform_n;
    const int64_t o1 = offsets ? offsets[b + 1] : o0 + uniform_n;
    const float qx = basis[3 * p], qy = basis[3 * p + 1], qz = basis[3 * p + 2];
    float best = __int_as_float(0x7f800000);
    int64_t bi = -1;
    for (int64_t k = o0; k < o1; ++k) {
        const float d = search_d2(qx, qy, qz, pts[3 * k], pts[3 * k + 1], pts[3 * k + 2]);
        if (bi < 0 || d < best) {
            best = d;
            bi = k;
        }
    }
    if (bi >= 0) keys[(size_t)b * P + p] = make_key(best, (uint32_t)(bi - o0));
}

__global__ void bps_finish_kernel(const float* __restrict__ basis, int P, const float* __restrict__ pts,
                                  const int64_t* __restrict__ offsets, int64_t uniform_n, int B,
                                  const unsigned long long* __restrict__ keys, float* __restrict__ dists,
                                  int32_t* __restrict__ idx_out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)B * P) return;
    const int b = (int)(i / P), p = (int)(i % P);
    const unsigned long long key = keys[i];
    const uint32_t k = (uint32_t)(key & 0xffffffffu);
    float dist = __int_as_float(0x7f800000);
    int32_t id = -1;
    if (key != ~0ull) {
        const int64_t o0 = offsets ? offsets[b] : (int64_t)b * uniform_n;
        const float* g = pts + 3 * (o0 + k);
        const float dx = __fsub_rn(g[0], basis[3 * p]), dy = __fsub_rn(g[1], basis[3 * p + 1]),
                    dz = __fsub_rn(g[2], basis[3 * p + 2]);
        const float sx = __fmul_rn(dx, dx), sy = __fmul_rn(dy, dy), sz = __fmul_rn(dz, dz);
        dist = __fsqrt_rn(__fadd_rn(__fadd_rn(sx, sz), sy));
        id = (int32_t)k;
    }
    dists[i] = dist;
    if (idx_out) idx_out[i] = id;
}

// ---- FP32-pipe probe (roofline denominator for the BPS kernel) ----
template <int KIND>
__global__ void __launch_bounds__(256) fp32_probe_kernel(int iters, float* out) {
    float a[8], s = 1.0000001f + threadIdx.x * 1e-9f, t = 1e-7f;
    f32x2 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        a[i] = i * 0.5f + threadIdx.x;
        v[i] = pack2(a[i], a[i] + 1.f);
    }
    const f32x2 s2 = pack2(s, s), t2 = pack2(t, t);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (KIND == 0) a[i] = __fmaf_rn(a[i], s, t);
                if (KIND == 1) v[i] = fma2(v[i], s2, t2);
                if (KIND == 2) v[i] = add2(v[i], t2);
                if (KIND == 3) a[i] = min3(a[i], a[(i + 1) & 7] + 0.f, s);
            }
        }
    }
    float r = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float lo, hi;
        unpack2(v[i], lo, hi);
        r += a[i] + lo + hi;
    }
    if (r == 123.456f) out[0] = r;
}

}  // namespace dexgg

namespace dexgg {
int bps_umma_launch(const float* basis, int P, const float* pts, const int64_t* offsets, int B, int64_t uniform_n, int64_t total,
                    int nsplit, unsigned long long* keys, void* extra, cudaStream_t st);  // bps_umma.cu
size_t bps_umma_extra_bytes(int P);
int bps_umma_basis_per_cta();
int bps_umma_ctas_per_sm();
}

using namespace dexgg;

extern "C" size_t dexgg_bps_workspace_bytes(int B, int P) {
    if (B < 0 || P < 0) return 0;
    return align_up((size_t)B * (size_t)P * sizeof(unsigned long long), 256) + bps_umma_extra_bytes(P);
}

template <int R>
static int launch_packed(const float* basis, int P, const float* pts, const int64_t* offsets, int B, int64_t uniform_n,
                         int64_t total, int nsplit, unsigned long long* keys, cudaStream_t st) {
    const size_t smem = 2 * BPS_STAGE_FLOATS * sizeof(float) + (size_t)R * BPS_THREADS * 8;
    static std::atomic<unsigned long long> attr_mask{0};
    if (smem > 48 * 1024) DEXGG_SMEM_OPT_IN(bps_nn_packed_kernel<R>, smem, attr_mask);
    dim3 grid((P + BPS_THREADS * R - 1) / (BPS_THREADS * R), B, nsplit);
    bps_nn_packed_kernel<R><<<grid, BPS_THREADS, smem, st>>>(basis, P, pts, offsets, uniform_n, total, nsplit, keys); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("bps_nn_packed_kernel");
    return DEXGG_OK;
}

template <int R>
static int launch_filter(const float* basis, int P, const float* pts, const int64_t* offsets, int B, int64_t uniform_n,
                         int64_t total, int nsplit, unsigned long long* keys, cudaStream_t st) {
    const size_t smem = (2 * BPS_STAGE_FLOATS + BPSF_W_FLOATS) * sizeof(float) + (size_t)R * BPS_THREADS * 4;
    static std::atomic<unsigned long long> attr_mask{0};
    if (smem > 48 * 1024) DEXGG_SMEM_OPT_IN(bps_nn_filter_kernel<R>, smem, attr_mask);
    dim3 grid((P + BPS_THREADS * R - 1) / (BPS_THREADS * R), B, nsplit);
    bps_nn_filter_kernel<R><<<grid, BPS_THREADS, smem, st>>>(basis, P, pts, offsets, uniform_n, total, nsplit, keys); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("bps_nn_filter_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_bps_encode(const float* basis, int P, const float* pts, const int64_t* offsets, int B,
                                int64_t max_cloud_points, int64_t total_points, float* dists, int32_t* idx, int flags,
                                void* workspace, size_t workspace_bytes, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(P >= 1 && B >= 0 && max_cloud_points >= 0 && total_points >= 0, "bps_encode: bad sizes P=%d B=%d", P, B);
    if (B == 0) return DEXGG_OK;
    DEXGG_REQUIRE(basis && dists && (pts || total_points == 0), "bps_encode: null pointer");
    DEXGG_REQUIRE(B <= 65535, "bps_encode: at most 65535 clouds per call (got %d)", B);
    DEXGG_REQUIRE((reinterpret_cast<uintptr_t>(pts) & 15) == 0, "bps_encode: pts must be 16-byte aligned");
    DEXGG_REQUIRE(max_cloud_points < (1ll << 32), "bps_encode: cloud too large");
    if (!offsets) DEXGG_REQUIRE(total_points == (int64_t)B * max_cloud_points, "bps_encode: uniform clouds need total_points == B*max_cloud_points");
    if (!workspace || workspace_bytes < dexgg_bps_workspace_bytes(B, P)) {
        set_error("bps_encode: workspace of %zu bytes needed, %zu given", dexgg_bps_workspace_bytes(B, P), workspace_bytes);
        return DEXGG_E_WORKSPACE;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned long long* keys = static_cast<unsigned long long*>(workspace);
    DEXGG_CUDA_OK(cudaMemsetAsync(keys, 0xff, (size_t)B * P * sizeof(unsigned long long), st));
    if (total_points > 0 && max_cloud_points > 0) {
        if (flags & DEXGG_BPS_SCALAR) {
            dim3 grid((P + 127) / 128, B);
            bps_nn_scalar_kernel<<<grid, 128, 0, st>>>(basis, P, pts, offsets, max_cloud_points, keys); ::dexgg::count_launches(1);
            DEXGG_LAUNCH_OK("bps_nn_scalar_kernel");
        } else if (!(flags & (DEXGG_BPS_DIRECT | DEXGG_BPS_FMA))) {
            // default: the filter scan on the tensor cores (bps_umma.cu), 512 basis points x one point range per CTA, two CTAs per
            // SM.  Splits of the point range (merged by atomicMin): enough CTAs to fill the GPU when there are few clouds (the
            // online one-object shape, dexgangrasp_real_robot.py:139), and ranges of at most 2048 points, which stay resident in
            // shared memory for the exact pass; never below 512 points per split.
            const int gx = (P + bps_umma_basis_per_cta() - 1) / bps_umma_basis_per_cta();
            int nsplit = (flags >> 8) & 0xff;
            if (nsplit == 0) {
                const int64_t have = (int64_t)gx * B, slots = (int64_t)bps_umma_ctas_per_sm() * sm_count();
                int64_t s = (max_cloud_points + 2047) / 2048;
                if (have * s < slots) s = (slots + have - 1) / have;
                const int64_t smax = (max_cloud_points + 511) / 512;
                if (s > smax) s = smax;
                if (s < 1) s = 1;
                if (s > 4096) s = 4096;
                nsplit = (int)s;
            }
            if ((max_cloud_points + nsplit - 1) / nsplit + 8 > (int64_t)65536 * 32) {  // 16-bit chunk ids
                int rc = launch_packed<16>(basis, P, pts, offsets, B, max_cloud_points, total_points, nsplit, keys, st);
                if (rc != DEXGG_OK) return rc;
            } else {
                int rc = bps_umma_launch(basis, P, pts, offsets, B, max_cloud_points, total_points, nsplit, keys,
                                         static_cast<unsigned char*>(workspace) + align_up((size_t)B * P * sizeof(unsigned long long), 256), st);
                if (rc != DEXGG_OK) return rc;
            }
        } else {
            // basis points per thread: 8 for the filter kernel (94 registers -> 5 CTAs / 20 warps per SM; measured 1.52 ms at C3
            // against 1.65 ms for 16 and 1.55 ms for 4), 16 for the direct kernel as tuned in round 1
            constexpr int RF = 8, RD = 16;
            const int R = (flags & DEXGG_BPS_DIRECT) ? RD : RF;
            const int gx = (P + BPS_THREADS * R - 1) / (BPS_THREADS * R);
            int nsplit = (flags >> 8) & 0xff;
            if (nsplit == 0) {
                // Splits of the point range (merged by atomicMin) so that the grid fills whole waves of 5 resident CTAs per SM:
                // the smallest split count whose last wave is >= 90 % full, else the fullest; never below one 1024-point tile per
                // split (a handful of clouds - the online one-object shape, dexgangrasp_real_robot.py:139 - gets one tile per split)
                const int64_t have = (int64_t)gx * B, slots = 5ll * sm_count();
                int64_t smax = (max_cloud_points + 1023) / 1024;
                if (smax > 64) smax = 64;
                if (smax < 1) smax = 1;
                int64_t s = 1;
                double best_eff = 0.0;
                for (int64_t c = 1; c <= smax; ++c) {
                    if (have * c >= 16 * slots && c > 1) break;  // many waves already: quantisation is below 6 %
                    const double waves = (double)(have * c) / (double)slots;
                    const double eff = waves / (double)((have * c + slots - 1) / slots);
                    if (eff > best_eff + 1e-9) { best_eff = eff; s = c; }
                    if (eff >= 0.9) break;
                }
                nsplit = (int)s;
            }
            // the filtered kernel addresses chunks with 14 bits: at most 2^14 * 32 points per split
            const bool direct = (flags & DEXGG_BPS_DIRECT) || (max_cloud_points + nsplit - 1) / nsplit + 8 > (int64_t)16384 * BPSF_CHUNK;
            int rc = direct ? launch_packed<RD>(basis, P, pts, offsets, B, max_cloud_points, total_points, nsplit, keys, st)
                            : launch_filter<RF>(basis, P, pts, offsets, B, max_cloud_points, total_points, nsplit, keys, st);
            if (rc != DEXGG_OK) return rc;
        }
    }
    const size_t tot = (size_t)B * P;
    bps_finish_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(basis, P, pts, offsets, max_cloud_points, B, keys, dists, idx); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("bps_finish_kernel");
    return DEXGG_OK;
}

// ---- allocation-free variant: the workspace is a file-scope device array (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).  Every GPU that loads the library holds the array until exit.  In HOST code the array's name is only a stand-in for
// the symbol, not a device address: the address is asked for once per device, with that device current (cudaGetSymbolAddress), and
// that pointer is what the kernels get.  All calls on a device share the array and it keeps the previous call's contents: the keys
// and the basis norms are re-initialised in stream order by every call (dexgg_bps_encode does that for any workspace); two calls
// that overlap on the GPU (different streams without an ordering between them) race on it.
constexpr size_t BPS_STATIC_MAX_KEYS = 64 * 4096;  // B * P of one call: e.g. 64 clouds x the reference's 4096 basis points
constexpr size_t BPS_STATIC_BYTES = BPS_STATIC_MAX_KEYS * sizeof(unsigned long long) + (512u << 10);  // + packed basis operand, norms
static __device__ __align__(256) unsigned char bps_static_scratch[BPS_STATIC_BYTES];

extern "C" size_t dexgg_bps_encode_static_max_keys(void) { return BPS_STATIC_MAX_KEYS; }

extern "C" int dexgg_bps_encode_static(const float* basis, int P, const float* pts, const int64_t* offsets, int B,
                                       int64_t max_cloud_points, int64_t total_points, float* dists, int32_t* idx, int flags,
                                       void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(P >= 1 && B >= 0, "bps_encode_static: bad sizes P=%d B=%d", P, B);
    const size_t need = dexgg_bps_workspace_bytes(B, P);
    DEXGG_REQUIRE((size_t)B * (size_t)P <= BPS_STATIC_MAX_KEYS && need <= BPS_STATIC_BYTES,
                  "bps_encode_static: B * P = %zu exceeds the built-in scratch (at most %zu keys, %zu bytes; this call needs %zu bytes): "
                  "use dexgg_bps_encode with a caller workspace", (size_t)B * (size_t)P, BPS_STATIC_MAX_KEYS, BPS_STATIC_BYTES, need);
    static std::atomic<void*> dev_ptr[64];  // per device: the array's DEVICE address
    int dev = 0;
    DEXGG_CUDA_OK(cudaGetDevice(&dev));
    DEXGG_REQUIRE(dev >= 0 && dev < 64, "bps_encode_static: device ordinal %d not supported", dev);
    void* ws = dev_ptr[dev].load(std::memory_order_acquire);
    if (!ws) {
        DEXGG_CUDA_OK(cudaGetSymbolAddress(&ws, bps_static_scratch));
        DEXGG_REQUIRE(ws != nullptr, "bps_encode_static: cudaGetSymbolAddress returned a null address");
        dev_ptr[dev].store(ws, std::memory_order_release);
    }
    return dexgg_bps_encode(basis, P, pts, offsets, B, max_cloud_points, total_points, dists, idx, flags, ws, BPS_STATIC_BYTES, stream);
}

// ---- the step in front of the encoder: obj_pcd.translate(-obj_pcd.get_center()) (dexgangrasp_real_robot.py:133-134,
// bps_encoder.py:69).  Open3D holds the points as doubles: the centre is the float64 mean, the shift is done in float64 and the
// cloud goes back to float32 when bps_torch takes it.  One CTA per cloud sums in float64 in a fixed order (deterministic). ----
__global__ void __launch_bounds__(256) cloud_center_kernel(const float* __restrict__ pts, const int64_t* __restrict__ offsets,
                                                            int64_t uniform_n, double* __restrict__ centers) {
    __shared__ double s[3][256];
    const int b = blockIdx.x;
    const int64_t o0 = offsets ? offsets[b] : (int64_t)b * uniform_n, o1 = offsets ? offsets[b + 1] : o0 + uniform_n;
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (int64_t i = o0 + threadIdx.x; i < o1; i += 256) {
        ax += (double)pts[3 * i]; ay += (double)pts[3 * i + 1]; az += (double)pts[3 * i + 2];
    }
    s[0][threadIdx.x] = ax; s[1][threadIdx.x] = ay; s[2][threadIdx.x] = az;
    __syncthreads();
    for (int w = 128; w; w >>= 1) {
        if ((int)threadIdx.x < w)
            for (int c = 0; c < 3; ++c) s[c][threadIdx.x] += s[c][threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x < 3) centers[3 * b + threadIdx.x] = o1 > o0 ? s[threadIdx.x][0] / (double)(o1 - o0) : 0.0;
}
__global__ void __launch_bounds__(256) cloud_shift_kernel(const float* __restrict__ pts, const int64_t* __restrict__ offsets,
                                                           int64_t uniform_n, const double* __restrict__ centers, float* __restrict__ out) {
    const int b = blockIdx.y;
    const int64_t o0 = offsets ? offsets[b] : (int64_t)b * uniform_n, o1 = offsets ? offsets[b + 1] : o0 + uniform_n;
    const double cx = centers[3 * b], cy = centers[3 * b + 1], cz = centers[3 * b + 2];
    for (int64_t i = o0 + (int64_t)blockIdx.x * 256 + threadIdx.x; i < o1; i += (int64_t)gridDim.x * 256) {
        out[3 * i] = (float)((double)pts[3 * i] - cx);
        out[3 * i + 1] = (float)((double)pts[3 * i + 1] - cy);
        out[3 * i + 2] = (float)((double)pts[3 * i + 2] - cz);
    }
}

extern "C" int dexgg_center_clouds(const float* pts, const int64_t* offsets, int B, int64_t max_cloud_points, int64_t total_points,
                                   float* pts_out, double* centers, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(B >= 0 && max_cloud_points >= 0 && total_points >= 0, "center_clouds: bad sizes");
    if (B == 0) return DEXGG_OK;
    DEXGG_REQUIRE(centers && (total_points == 0 || (pts && pts_out)) && B <= 65535, "center_clouds: null pointer or more than 65535 clouds");
    if (!offsets) DEXGG_REQUIRE(total_points == (int64_t)B * max_cloud_points, "center_clouds: uniform clouds need total_points == B*max_cloud_points");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cloud_center_kernel<<<B, 256, 0, st>>>(pts, offsets, max_cloud_points, centers);
    int gx = (int)((max_cloud_points + 2047) / 2048);
    if (gx < 1) gx = 1;
    if (gx > 1024) gx = 1024;
    cloud_shift_kernel<<<dim3(gx, B), 256, 0, st>>>(pts, offsets, max_cloud_points, centers, pts_out);
    ::dexgg::count_launches(2);
    DEXGG_LAUNCH_OK("center_clouds kernels");
    return DEXGG_OK;
}

extern "C" int dexgg_bps_encode_host(const float* basis, int P, const float* pts, const int64_t* offsets, int B,
                                     float* dists, int32_t* idx, int flags) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(P >= 1 && B >= 0 && basis && offsets && dists, "bps_encode_host: bad arguments");
    if (B == 0) return DEXGG_OK;
    const int64_t total = offsets[B];
    int64_t maxn = 0;
    for (int b = 0; b < B; ++b) {
        DEXGG_REQUIRE(offsets[b + 1] >= offsets[b], "bps_encode_host: offsets must be non-decreasing");
        if (offsets[b + 1] - offsets[b] > maxn) maxn = offsets[b + 1] - offsets[b];
    }
    float *d_basis = nullptr, *d_pts = nullptr, *d_dists = nullptr;
    int64_t* d_off = nullptr;
    int32_t* d_idx = nullptr;
    void* ws = nullptr;
    const size_t wsb = dexgg_bps_workspace_bytes(B, P);
    int rc = DEXGG_OK;
    cudaError_t e;
#define HOST_TRY(x)                                                          \
    if (rc == DEXGG_OK && (e = (x)) != cudaSuccess) {                        \
        set_error("%s failed: %s", #x, cudaGetErrorString(e));               \
        rc = DEXGG_E_CUDA;                                                   \
    }
    HOST_TRY(cudaMalloc(&d_basis, (size_t)P * 12));
    HOST_TRY(cudaMalloc(&d_pts, (size_t)(total > 0 ? total : 1) * 12 + 16));
    HOST_TRY(cudaMalloc(&d_off, (size_t)(B + 1) * 8));
    HOST_TRY(cudaMalloc(&d_dists, (size_t)B * P * 4));
    HOST_TRY(cudaMalloc(&d_idx, (size_t)B * P * 4));
    HOST_TRY(cudaMalloc(&ws, wsb));
    HOST_TRY(cudaMemcpy(d_basis, basis, (size_t)P * 12, cudaMemcpyHostToDevice));
    if (total > 0) HOST_TRY(cudaMemcpy(d_pts, pts, (size_t)total * 12, cudaMemcpyHostToDevice));
    HOST_TRY(cudaMemcpy(d_off, offsets, (size_t)(B + 1) * 8, cudaMemcpyHostToDevice));
    if (rc == DEXGG_OK)
        rc = dexgg_bps_encode(d_basis, P, d_pts, d_off, B, maxn, total, d_dists, d_idx, flags, ws, wsb, nullptr);
    HOST_TRY(cudaMemcpy(dists, d_dists, (size_t)B * P * 4, cudaMemcpyDeviceToHost));
    if (idx) HOST_TRY(cudaMemcpy(idx, d_idx, (size_t)B * P * 4, cudaMemcpyDeviceToHost));
#undef HOST_TRY
    cudaFree(d_basis);
    cudaFree(d_pts);
    cudaFree(d_off);
    cudaFree(d_dists);
    cudaFree(d_idx);
    cudaFree(ws);
    return rc;
}

extern "C" int dexgg_fp32_pipe_probe(int kind, int iters, float* ms, double* lane_ops) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(kind >= 0 && kind <= 3 && iters > 0 && ms && lane_ops, "fp32_pipe_probe: bad arguments");
    float* out = nullptr;
    DEXGG_CUDA_OK(cudaMalloc(&out, 4));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = sm_count() * 8, block = 256;
    for (int rep = 0; rep < 2; ++rep) {  // first launch warms up
        cudaEventRecord(e0);
        if (kind == 0) { fp32_probe_kernel<0><<<grid, block>>>(iters, out); ::dexgg::count_launches(1); }
        if (kind == 1) { fp32_probe_kernel<1><<<grid, block>>>(iters, out); ::dexgg::count_launches(1); }
        if (kind == 2) { fp32_probe_kernel<2><<<grid, block>>>(iters, out); ::dexgg::count_launches(1); }
        if (kind == 3) { fp32_probe_kernel<3><<<grid, block>>>(iters, out); ::dexgg::count_launches(1); }
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
    }
    cudaEventElapsedTime(ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    DEXGG_LAUNCH_OK("fp32_probe_kernel");
    const double per_thread = (double)iters * 64.0 * ((kind == 1 || kind == 2) ? 2.0 : 1.0);
    *lane_ops = per_thread * (double)grid * block;
    return DEXGG_OK;
}
