// Fused non-GEMM kernels of the bf16 training step (train_engine.py drives them; the GEMMs are gemm_umma.cu).
//
// Reference: DexGanGrasp.update_dexgangrasp / update_dexevaluator (DexGanGrasp/models/dexgangrasp.py:739-823) over
// ResBlock / Generator / Discriminator / DexEvaluator in training mode (models/networks.py:43-56, 88-103, 148-159, 315-337),
// losses (dexgangrasp.py:142-147, 213-232, 272-307; models/losses.py:61-76), utils.rot_matrix_from_ortho6d (utils.py:380-394).
//
// Layout of the step (all activations bf16, all statistics / losses / gradients of parameters fp32):
//   * The input BatchNorm of every network normalises the SAME BPS batch: xhat = (bps - mean) / sqrt(var + eps) is computed
//     once per step (bn_input_stats_kernel, xhat_kernel) and the per-network gamma / beta are folded into the first-layer
//     weights when they are converted to bf16 (prep_weights_kernel); unfold_grads_kernel maps the gradients of the folded
//     weights back to fc*.weight and to the input BatchNorm's gamma / beta, so no GEMM ever produces a [rows, 4096] gradient.
//   * A train-mode BatchNorm between two GEMMs is ONE column-slab kernel: a CTA owns 16 columns over all rows of a batch group,
//     computes the batch statistics (shifted single pass), then normalises / adds the skip branch / applies LeakyReLU (and
//     dropout) and writes the bf16 operand of the next GEMM (bn_fwd_slab_kernel).  Its backward is the mirror image
//     (bn_bwd_slab_kernel: LeakyReLU', the two column sums of the BatchNorm backward, then dx and the skip-branch gradient).
//   * Heads, losses and the rot6d Gram-Schmidt are row kernels (score_head_kernel, gen_pose_kernel, gen_loss_kernel).
// All of them are HBM / L2-bandwidth-bound streaming kernels (2 - 8 B per element).
#include <cuda_bf16.h>

#include "common.cuh"
#include "pose.cuh"

#include <string.h>

namespace dexgg {

constexpr float TF_SLOPE = 0.2f;

__device__ __forceinline__ void unpack8(const uint4 u, float (&v)[8]) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        v[2 * i] = __uint_as_float(w[i] << 16);
        v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
}
__device__ __forceinline__ uint4 pack8(const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ uint4 ld8(const __nv_bfloat16* p) { return *reinterpret_cast<const uint4*>(p); }

// Dropout keep-scale of element `idx` of dropout layer `stream` at step seed `seed`: 1 / (1 - p) with probability 1 - p, else 0
// (nn.Dropout(0.3), networks.py:293).  Counter-based (splitmix64 finaliser), so backward regenerates the forward's mask.
__device__ __forceinline__ float drop_scale(unsigned long long seed, unsigned stream, unsigned long long idx, float p) {
    unsigned long long z = seed + 0x9e3779b97f4a7c15ull * (idx + 1) + ((unsigned long long)stream << 56);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    z ^= z >> 31;
    const float u = (float)(z >> 40) * (1.0f / 16777216.0f);
    return u >= p ? 1.0f / (1.0f - p) : 0.f;
}

// Sum NQ x 8 per-thread values over the 128 row lanes of a 256-thread slab CTA (thread t: column half t & 1, row lane t >> 1).
// Result for this thread's own 8 columns in `acc` on return (valid in every thread).
template <int NQ>
__device__ __forceinline__ void slab_reduce(float (&acc)[NQ][8], float* sm /* [8 warps][2][NQ][8] + [2][NQ][8] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = threadIdx.x & 1;
#pragma unroll
    for (int q = 0; q < NQ; ++q)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float v = acc[q][j];
#pragma unroll
            for (int o = 2; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            acc[q][j] = v;
        }
    __syncthreads();  // sm may still be read by the previous reduction
    if (lane < 2) {
#pragma unroll
        for (int q = 0; q < NQ; ++q)
#pragma unroll
            for (int j = 0; j < 8; ++j) sm[((warp * 2 + sub) * NQ + q) * 8 + j] = acc[q][j];
    }
    __syncthreads();
    float* out = sm + 8 * 2 * NQ * 8;
    for (int i = threadIdx.x; i < 2 * NQ * 8; i += blockDim.x) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += sm[w * 2 * NQ * 8 + i];
        out[i] = s;
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < NQ; ++q)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[q][j] = out[(sub * NQ + q) * 8 + j];
}

// ---- column sums across the CTAs of a slab kernel AND across the GPUs of a data-parallel job, inside the kernel ----
// Every rank owns a peer-mapped scratch (cudaIpc, NVLink P2P): two parity buffers of SLAB_SLOTS slots; a slab launch uses one
// slot: [local accumulator 2048 f | one 2048-float slice per rank | counters].  Protocol of a launch (every rank runs the same
// launch sequence, so slot numbers and grid sizes agree):
//   1. each CTA adds its partial column sums to the LOCAL accumulator (atomics) and bumps the local arrival counter;
//   2. the last CTA to arrive copies the local totals into slice[rank] of EVERY rank's slot (plain stores over NVLink, a
//      system-scope fence) and then bumps every rank's `published` counter;
//   3. every CTA spins until its own rank's `published` == world, then sums the slices in rank order: all ranks get bit-identical
//      totals, with ONE exchange per BatchNorm and no NCCL launch.  world == 1 is the same code (one slice).
// The grids hold at most one CTA per SM and run stream-ordered behind their producers, so every CTA is resident while it spins.
// The parity buffer alternates per step (dexgg_slab_step_begin flips it on the device and zeroes the other one for the next
// step), which keeps a fast rank's next-step traffic out of a slow rank's current-step slots.
constexpr int SLAB_ACC = 2048, SLAB_MAXW = 8, SLAB_SLOT_FLOATS = SLAB_ACC * (1 + SLAB_MAXW) + 64, SLAB_SLOTS = 64;
struct SlabComm {
    float* const* peers;       // device array [world]: scratch base of every rank (peers[rank] = own)
    int world, rank, slot;
    const unsigned* parity;    // device word, 0 / 1
};
__device__ __forceinline__ float* slab_slot(const SlabComm& c, int peer) {
    return c.peers[peer] + ((size_t)(*c.parity) * SLAB_SLOTS + c.slot) * SLAB_SLOT_FLOATS;
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_sys(unsigned* p, unsigned v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// n floats of the local accumulator are exchanged; returns (in every thread) a pointer to this rank's slot
__device__ __forceinline__ float* slab_exchange(const SlabComm& c, int n, unsigned nblocks) {
    __shared__ int s_last;
    float* mine = slab_slot(c, c.rank);
    unsigned* cnt = reinterpret_cast<unsigned*>(mine + SLAB_ACC * (1 + SLAB_MAXW));
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(cnt, 1u) == nblocks - 1;
    }
    __syncthreads();
    if (s_last) {
        if (c.world == 1) {  // single rank: the local accumulator IS the total
            if (threadIdx.x == 0) {
                __threadfence();
                red_release_sys(cnt + 1, 1u);
            }
        } else {
            __threadfence();
            for (int i = threadIdx.x; i < n; i += blockDim.x) {
                const float v = __ldcg(mine + i);
                for (int p = 0; p < c.world; ++p) __stcg(slab_slot(c, p) + SLAB_ACC * (1 + c.rank) + i, v);
            }
            __threadfence_system();
            __syncthreads();
            if ((int)threadIdx.x < c.world)
                red_release_sys(reinterpret_cast<unsigned*>(slab_slot(c, threadIdx.x) + SLAB_ACC * (1 + SLAB_MAXW)) + 1, 1u);
        }
    }
    if (threadIdx.x == 0) {
        while (ld_acquire_sys(cnt + 1) < (unsigned)c.world) __nanosleep(20);
    }
    __syncthreads();
    return mine;
}
__device__ __forceinline__ float slab_total(const SlabComm& c, const float* mine, int i) {
    if (c.world == 1) return __ldcv(mine + i);
    float t = 0.f;
    for (int r = 0; r < c.world; ++r) t += __ldcv(mine + SLAB_ACC * (1 + r) + i);
    return t;
}
// Small collective over the same peer scratch (the statistics of the input / tail BatchNorm and the tail backward sums: a few
// hundred floats that would otherwise each cost an NCCL launch): CTA b owns the 2048-float chunk b of `src` and slot c.slot + b.
// It stores its chunk into slice[rank] of every rank's slot, flags, waits for all ranks, then writes the rank-ordered sum to
// sum_out (may alias src) and / or every rank's chunk to gather_out [world, n].
__global__ void __launch_bounds__(256) p2p_small_kernel(SlabComm c, const float* src, int n, float* sum_out, float* gather_out) {
    c.slot += blockIdx.x;
    const int i0 = blockIdx.x * SLAB_ACC, m = (n - i0 < SLAB_ACC) ? n - i0 : SLAB_ACC;
    float* mine = slab_slot(c, c.rank);
    unsigned* cnt = reinterpret_cast<unsigned*>(mine + SLAB_ACC * (1 + SLAB_MAXW));
    for (int i = threadIdx.x; i < m; i += blockDim.x) {
        const float v = src[i0 + i];
        for (int p = 0; p < c.world; ++p) __stcg(slab_slot(c, p) + SLAB_ACC * (1 + c.rank) + i, v);
    }
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < c.world) red_release_sys(reinterpret_cast<unsigned*>(slab_slot(c, threadIdx.x) + SLAB_ACC * (1 + SLAB_MAXW)) + 1, 1u);
    if (threadIdx.x == 0) {
        while (ld_acquire_sys(cnt + 1) < (unsigned)c.world) __nanosleep(20);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < m; i += blockDim.x) {
        float t = 0.f;
        for (int r = 0; r < c.world; ++r) {
            const float v = __ldcv(mine + SLAB_ACC * (1 + r) + i);
            t += v;
            if (gather_out) gather_out[(int64_t)r * n + i0 + i] = v;
        }
        if (sum_out) sum_out[i0 + i] = t;
    }
}

// Gradient allreduce of a data-parallel step over the SAME peer mapping, one launch, no NCCL: every rank's flat gradient buffer
// lives in peer-mapped memory (bufs[p] = rank p's buffer as seen from here).  Two-shot: rank r owns the r-th n/world chunk; it
// reads that chunk from every rank over NVLink, sums in rank order, scales (the 1/world of the mean) and stores the result into
// the chunk of EVERY rank - so all ranks hold bit-identical gradients.  Two flag barriers through one exchange slot:
//   ready (cnt[1]): every rank's kernel has started, i.e. its gradients are complete in stream order;
//   done  (cnt[2]): every owner's stores have landed; only the last CTA of a rank waits for it, which keeps the kernel (and the
//                   Adam update behind it in the stream) from completing early.
// A small grid of narrow CTAs (<= 85 registers x 256 threads): it always fits beside the capped slab grid of the other chain of
// a step, so it may spin while that chain runs - unlike an NCCL kernel, whose CTAs do not fit an SM holding a slab CTA.
__global__ void __launch_bounds__(256, 3) grad_allreduce_kernel(SlabComm c, float4* const* bufs, long long n4, float scale) {
    __shared__ int s_last;
    float* mine = slab_slot(c, c.rank);
    unsigned* cnt = reinterpret_cast<unsigned*>(mine + SLAB_ACC * (1 + SLAB_MAXW));
    if (blockIdx.x == 0 && (int)threadIdx.x < c.world)
        red_release_sys(reinterpret_cast<unsigned*>(slab_slot(c, threadIdx.x) + SLAB_ACC * (1 + SLAB_MAXW)) + 1, 1u);
    if (threadIdx.x == 0) {
        while (ld_acquire_sys(cnt + 1) < (unsigned)c.world) __nanosleep(20);
    }
    __syncthreads();
    float4* b[SLAB_MAXW];
#pragma unroll
    for (int p = 0; p < SLAB_MAXW; ++p) b[p] = bufs[p < c.world ? p : 0];
    const long long per = (n4 + c.world - 1) / c.world, lo = per * c.rank, hi = (lo + per < n4) ? lo + per : n4;
    for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (long long)gridDim.x * blockDim.x) {
        float4 v[SLAB_MAXW];
#pragma unroll
        for (int p = 0; p < SLAB_MAXW; ++p)
            if (p < c.world) v[p] = __ldcg(b[p] + i);
        float4 t = v[0];
#pragma unroll
        for (int p = 1; p < SLAB_MAXW; ++p)
            if (p < c.world) { t.x += v[p].x; t.y += v[p].y; t.z += v[p].z; t.w += v[p].w; }
        t.x *= scale; t.y *= scale; t.z *= scale; t.w *= scale;
#pragma unroll
        for (int p = 0; p < SLAB_MAXW; ++p)
            if (p < c.world) __stcg(b[p] + i, t);
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        s_last = atomicAdd(cnt, 1u) == gridDim.x - 1;
        __threadfence();
    }
    __syncthreads();
    if (s_last) {
        if ((int)threadIdx.x < c.world)
            red_release_sys(reinterpret_cast<unsigned*>(slab_slot(c, threadIdx.x) + SLAB_ACC * (1 + SLAB_MAXW)) + 2, 1u);
        if (threadIdx.x == 0) {
            while (ld_acquire_sys(cnt + 2) < (unsigned)c.world) __nanosleep(20);
        }
    }
}

// flips the step parity and zeroes the OTHER parity buffer (used by the next step)
__global__ void slab_step_begin_kernel(float* base, unsigned* parity) {
    __shared__ unsigned s_p;
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) s_p = *parity ^ 1u;   // block 0 flips; the other blocks compute the same value without writing
        else s_p = *parity ^ 1u;
    }
    __syncthreads();
    float* other = base + (size_t)(s_p ^ 1u) * SLAB_SLOTS * SLAB_SLOT_FLOATS;
    const size_t n = (size_t)SLAB_SLOTS * SLAB_SLOT_FLOATS;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) other[i] = 0.f;
}
__global__ void slab_parity_flip_kernel(unsigned* parity) { *parity ^= 1u; }

// ------------------------------------------------------------------------------------------------ BatchNorm forward slab
// grid (C / 16 column slabs, row splits, n_groups): a CTA owns 16 columns x one row range of one batch group.  Pass 1 adds the
// CTA's shifted partial sums (shift = the layer's running mean: common to every CTA and every rank) to acc [n_groups, 2, C];
// after the grid barrier every CTA derives mean / invstd and normalises its rows (pass 2).
struct BnFwd {
    const __nv_bfloat16* x; int64_t ldx;
    int C; int64_t rpg; int n_groups;
    const float *gamma, *beta, *shift;
    float *mean, *invstd;          // [n_groups, C]
    float eps, count;              // count = rows per group over all ranks
    const __nv_bfloat16* res; int64_t ldr;
    int act;
    const float* mask; int64_t ldm; float drop_p; const unsigned long long* seed; unsigned stream;
    __nv_bfloat16* y; int64_t ldy;
    int mode;                      // 0 fused (in-kernel exchange), 1 partial sums only (NCCL fallback: allreduce acc, then mode 2), 2 apply from acc
    float* acc;                    // modes 1 / 2: [n_groups, 2, C] zeroed by the caller
    SlabComm comm;                 // mode 0
};

__global__ void __launch_bounds__(256, 3) bn_fwd_slab_kernel(const BnFwd a) {
    __shared__ float sm[8 * 2 * 2 * 8 + 2 * 2 * 8];
    const int sub = threadIdx.x & 1, rl = threadIdx.x >> 1, g = blockIdx.z;
    const int c0 = blockIdx.x * 16 + sub * 8;
    const int64_t chunk = ((a.rpg + gridDim.y - 1) / gridDim.y + 127) / 128 * 128;
    const int64_t g0 = (int64_t)g * a.rpg;
    const int64_t r0 = g0 + (int64_t)blockIdx.y * chunk, r1 = (r0 + chunk < g0 + a.rpg) ? r0 + chunk : g0 + a.rpg;
    float s[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] = a.shift ? a.shift[c0 + j] : 0.f;
    float* accl = a.mode == 0 ? slab_slot(a.comm, a.comm.rank) : a.acc;  // local accumulator [n_groups, 2, C]
    float* accg = accl + (int64_t)g * 2 * a.C;
    if (a.mode != 2) {
        float acc[2][8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[0][j] = acc[1][j] = 0.f;
#pragma unroll 4
        for (int64_t r = r0 + rl; r < r1; r += 128) {
            float v[8];
            unpack8(ld8(a.x + r * a.ldx + c0), v);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float d = v[j] - s[j];
                acc[0][j] += d;
                acc[1][j] = fmaf(d, d, acc[1][j]);
            }
        }
        slab_reduce<2>(acc, sm);
        if (rl == 0 && r0 < r1) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                atomicAdd(accg + c0 + j, acc[0][j]);
                atomicAdd(accg + a.C + c0 + j, acc[1][j]);
            }
        }
        if (a.mode == 1) return;
        slab_exchange(a.comm, a.n_groups * 2 * a.C, gridDim.x * gridDim.y * gridDim.z);
    }
    float sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int i0 = g * 2 * a.C + c0 + j;
        const float sd = a.mode == 0 ? slab_total(a.comm, accl, i0) : __ldcg(accg + c0 + j);
        const float sq = a.mode == 0 ? slab_total(a.comm, accl, i0 + a.C) : __ldcg(accg + a.C + c0 + j);
        const float mean = s[j] + sd / a.count;
        const float var = fmaxf(sq / a.count - (sd / a.count) * (sd / a.count), 0.f);
        const float istd = rsqrtf(var + a.eps);
        if (rl == 0 && blockIdx.y == 0) {
            a.mean[(int64_t)g * a.C + c0 + j] = mean;
            a.invstd[(int64_t)g * a.C + c0 + j] = istd;
        }
        sc[j] = a.gamma[c0 + j] * istd;
        sh[j] = a.beta[c0 + j] - mean * sc[j];
    }
    const unsigned long long seed = (a.drop_p > 0.f && !a.mask && a.seed) ? *a.seed : 0ull;
#pragma unroll 2
    for (int64_t r = r0 + rl; r < r1; r += 128) {
        float v[8];
        unpack8(ld8(a.x + r * a.ldx + c0), v);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = fmaf(v[j], sc[j], sh[j]);
        if (a.res) {
            float q[8];
            unpack8(ld8(a.res + r * a.ldr + c0), q);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] += q[j];
        }
        if (a.act) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = v[j] > 0.f ? v[j] : TF_SLOPE * v[j];
        }
        if (a.mask) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] *= a.mask[r * a.ldm + c0 + j];
        } else if (a.drop_p > 0.f) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] *= drop_scale(seed, a.stream, (unsigned long long)r * a.C + c0 + j, a.drop_p);
        }
        *reinterpret_cast<uint4*>(a.y + r * a.ldy + c0) = pack8(v);
    }
}

// ------------------------------------------------------------------------------------------------ BatchNorm backward slab
struct BnBwd {
    const __nv_bfloat16* g; int64_t ldg;        // incoming gradient wrt the forward output, or
    const float *dl, *w;                        // rank-1 form dl[m] * w[c] (score head: out_success, networks.py:158,336)
    float* dw_out;                              // [C] += sum_m dl[m] * out[m,c]   (rank-1 form only)
    const __nv_bfloat16* out; int64_t ldo;      // saved forward OUTPUT (LeakyReLU' sign; also the head's input)
    int act;
    const float* mask; int64_t ldm; float drop_p; const unsigned long long* seed; unsigned stream;
    const __nv_bfloat16* x; int64_t ldx;        // saved forward INPUT (pre-normalisation)
    const float *mean, *invstd, *gamma;
    int C; int64_t rpg; int n_groups; float count;
    float *dgamma, *dbeta;                      // [C] += (may be NULL)
    __nv_bfloat16* gx; int64_t ldgx;            // out: gradient wrt x
    const __nv_bfloat16* skip; int64_t lds;     // skip-branch activation (post-LeakyReLU) or NULL
    __nv_bfloat16* gskip; int64_t ldgs;         // out: g * LeakyReLU'(skip)
    float* dbias_skip;                          // [C] += sum_m gskip (may be NULL)
    int mode;                                   // 0 fused (in-kernel exchange), 1 column sums only (NCCL fallback), 2 apply from acc
    float* acc;                                 // modes 1 / 2: [n_groups, 2, C] zeroed by the caller: sum g, sum g * xhat
    SlabComm comm;                              // mode 0
};

__global__ void __launch_bounds__(256, 2) bn_bwd_slab_kernel(const BnBwd a) {
    __shared__ float sm[8 * 2 * 4 * 8 + 2 * 4 * 8];
    const int sub = threadIdx.x & 1, rl = threadIdx.x >> 1, gi = blockIdx.z;
    const int c0 = blockIdx.x * 16 + sub * 8;
    const int64_t chunk = ((a.rpg + gridDim.y - 1) / gridDim.y + 127) / 128 * 128;
    const int64_t g0 = (int64_t)gi * a.rpg;
    const int64_t r0 = g0 + (int64_t)blockIdx.y * chunk, r1 = (r0 + chunk < g0 + a.rpg) ? r0 + chunk : g0 + a.rpg;
    float mean[8], istd[8], wv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        mean[j] = a.mean[(int64_t)gi * a.C + c0 + j];
        istd[j] = a.invstd[(int64_t)gi * a.C + c0 + j];
        wv[j] = a.g ? 0.f : a.w[c0 + j];
    }
    const unsigned long long seed = (a.drop_p > 0.f && !a.mask && a.seed) ? *a.seed : 0ull;
    // gradient wrt the pre-activation sum (after dropout' and LeakyReLU'), 8 columns of row r
    auto grad_row = [&](int64_t r, float (&gv)[8], float (&ov)[8]) {
        if (a.g) unpack8(ld8(a.g + r * a.ldg + c0), gv);
        else {
            const float d = a.dl[r];
#pragma unroll
            for (int j = 0; j < 8; ++j) gv[j] = d * wv[j];
        }
        if (a.mask) {
#pragma unroll
            for (int j = 0; j < 8; ++j) gv[j] *= a.mask[r * a.ldm + c0 + j];
        } else if (a.drop_p > 0.f) {
#pragma unroll
            for (int j = 0; j < 8; ++j) gv[j] *= drop_scale(seed, a.stream, (unsigned long long)r * a.C + c0 + j, a.drop_p);
        }
        if (a.out) {
            unpack8(ld8(a.out + r * a.ldo + c0), ov);
            if (a.act) {
                // a dropped element is stored as 0: its gradient is already 0, the slope chosen for it does not matter
#pragma unroll
                for (int j = 0; j < 8; ++j) gv[j] = ov[j] > 0.f ? gv[j] : TF_SLOPE * gv[j];
            }
        }
    };
    float* accl = a.mode == 0 ? slab_slot(a.comm, a.comm.rank) : a.acc;
    float* accg = accl + (int64_t)gi * 2 * a.C;
    if (a.mode != 2) {
        float S[4][8];  // 0: sum g, 1: sum g * xhat, 2: sum dl * out, 3: sum gskip
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int j = 0; j < 8; ++j) S[q][j] = 0.f;
#pragma unroll 2
        for (int64_t r = r0 + rl; r < r1; r += 128) {
            float gv[8], ov[8], xv[8];
            grad_row(r, gv, ov);
            if (!a.g && a.out) {  // rank-1 form: the head's weight gradient needs the saved output itself
                const float d = a.dl[r];
#pragma unroll
                for (int j = 0; j < 8; ++j) S[2][j] = fmaf(d, ov[j], S[2][j]);
            }
            unpack8(ld8(a.x + r * a.ldx + c0), xv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                S[0][j] += gv[j];
                S[1][j] = fmaf(gv[j], (xv[j] - mean[j]) * istd[j], S[1][j]);
            }
            if (a.skip) {
                float sv[8];
                unpack8(ld8(a.skip + r * a.lds + c0), sv);
#pragma unroll
                for (int j = 0; j < 8; ++j) S[3][j] += sv[j] > 0.f ? gv[j] : TF_SLOPE * gv[j];
            }
        }
        slab_reduce<4>(S, sm);
        if (rl == 0 && r0 < r1) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                atomicAdd(accg + c0 + j, S[0][j]);
                atomicAdd(accg + a.C + c0 + j, S[1][j]);
                if (a.dgamma) atomicAdd(a.dgamma + c0 + j, S[1][j]);
                if (a.dbeta) atomicAdd(a.dbeta + c0 + j, S[0][j]);
                if (a.dw_out && !a.g) atomicAdd(a.dw_out + c0 + j, S[2][j]);
                if (a.dbias_skip && a.skip) atomicAdd(a.dbias_skip + c0 + j, S[3][j]);
            }
        }
        if (a.mode == 1) return;
        slab_exchange(a.comm, a.n_groups * 2 * a.C, gridDim.x * gridDim.y * gridDim.z);
    }
    float k0[8], k1[8], k2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int i0 = gi * 2 * a.C + c0 + j;
        k0[j] = a.gamma[c0 + j] * istd[j];
        k1[j] = (a.mode == 0 ? slab_total(a.comm, accl, i0) : __ldcg(accg + c0 + j)) / a.count;
        k2[j] = (a.mode == 0 ? slab_total(a.comm, accl, i0 + a.C) : __ldcg(accg + a.C + c0 + j)) / a.count;
    }
#pragma unroll 2
    for (int64_t r = r0 + rl; r < r1; r += 128) {
        float gv[8], ov[8], xv[8], o[8];
        grad_row(r, gv, ov);
        unpack8(ld8(a.x + r * a.ldx + c0), xv);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = k0[j] * (gv[j] - k1[j] - (xv[j] - mean[j]) * istd[j] * k2[j]);
        *reinterpret_cast<uint4*>(a.gx + r * a.ldgx + c0) = pack8(o);
        if (a.skip && a.gskip) {
            float sv[8];
            unpack8(ld8(a.skip + r * a.lds + c0), sv);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = sv[j] > 0.f ? gv[j] : TF_SLOPE * gv[j];
            *reinterpret_cast<uint4*>(a.gskip + r * a.ldgs + c0) = pack8(o);
        }
    }
}

// ------------------------------------------------------------------------------------------------ input BatchNorm (BPS columns)
// Column statistics of the fp32 BPS batch [B, C] (C % 4 == 0): grid (C / 128, row splits); a thread owns 4 adjacent columns
// (float4 loads, a warp reads 512 contiguous bytes of a row), warps stride over the rows.  Shifted sums (shift = row 0 of the batch,
// common to all CTAs) are added to acc [2, C] (zeroed by the caller); bn_input_finalize_kernel turns them into mean / var / invstd.
__global__ void __launch_bounds__(256) bn_input_stats_kernel(const float* __restrict__ x, int64_t ldx, int64_t B, int C, float* __restrict__ acc) {
    __shared__ float4 s0[8][32], s1[8][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = (blockIdx.x * 32 + lane) * 4;
    const int64_t chunk = (B + gridDim.y - 1) / gridDim.y;
    const int64_t r0 = (int64_t)blockIdx.y * chunk, r1 = (r0 + chunk < B) ? r0 + chunk : B;
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
    if (c < C) {
        const float4 sh = *reinterpret_cast<const float4*>(x + c);
#pragma unroll 4
        for (int64_t r = r0 + warp; r < r1; r += 8) {
            const float4 v = *reinterpret_cast<const float4*>(x + r * ldx + c);
            const float dx = v.x - sh.x, dy = v.y - sh.y, dz = v.z - sh.z, dw = v.w - sh.w;
            a0.x += dx; a0.y += dy; a0.z += dz; a0.w += dw;
            a1.x = fmaf(dx, dx, a1.x); a1.y = fmaf(dy, dy, a1.y); a1.z = fmaf(dz, dz, a1.z); a1.w = fmaf(dw, dw, a1.w);
        }
    }
    s0[warp][lane] = a0;
    s1[warp][lane] = a1;
    __syncthreads();
    if (warp == 0 && c < C) {
#pragma unroll
        for (int j = 1; j < 8; ++j) {
            const float4 p = s0[j][lane], q = s1[j][lane];
            a0.x += p.x; a0.y += p.y; a0.z += p.z; a0.w += p.w;
            a1.x += q.x; a1.y += q.y; a1.z += q.z; a1.w += q.w;
        }
        atomicAdd(acc + c, a0.x); atomicAdd(acc + c + 1, a0.y); atomicAdd(acc + c + 2, a0.z); atomicAdd(acc + c + 3, a0.w);
        atomicAdd(acc + C + c, a1.x); atomicAdd(acc + C + c + 1, a1.y); atomicAdd(acc + C + c + 2, a1.z); atomicAdd(acc + C + c + 3, a1.w);
    }
}
// local statistics from the shifted sums: mean, biased variance, invstd, and M2 = n * var (for the cross-rank Chan merge)
__global__ void bn_input_finalize_kernel(const float* __restrict__ x, const float* __restrict__ acc, int64_t B, int C, float eps,
                                         float* __restrict__ mean, float* __restrict__ var, float* __restrict__ invstd) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float n = (float)B, sd = acc[c], sq = acc[C + c];
    const float m2 = fmaxf(sq - sd * sd / n, 0.f);
    mean[c] = x[c] + sd / n;
    var[c] = m2 / n;
    invstd[c] = rsqrtf(m2 / n + eps);
}

// xhat = (x - mean) * invstd as bf16 [B, C] (C % 4 == 0), written `copies` times back to back (the discriminator's real | fake
// double batch reads the same rows twice)
__global__ void xhat_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ mean, const float* __restrict__ invstd,
                            int64_t B, int C, __nv_bfloat16* __restrict__ out, int copies) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int c4 = C / 4;
    if (i >= B * c4) return;
    const int64_t r = i / c4;
    const int c = (int)(i % c4) * 4;
    const float4 v = *reinterpret_cast<const float4*>(x + r * ldx + c);
    const float4 m = *reinterpret_cast<const float4*>(mean + c), s = *reinterpret_cast<const float4*>(invstd + c);
    const __nv_bfloat162 lo = __floats2bfloat162_rn((v.x - m.x) * s.x, (v.y - m.y) * s.y);
    const __nv_bfloat162 hi = __floats2bfloat162_rn((v.z - m.z) * s.z, (v.w - m.w) * s.w);
    const uint2 o = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
    for (int k = 0; k < copies; ++k) *reinterpret_cast<uint2*>(out + ((int64_t)k * B + r) * C + c) = o;
}

// ------------------------------------------------------------------------------------------------ per-row inputs ("tail")
struct TailSrc {          // up to three sources concatenated along columns: (rot 9 | transl 3 | joint nJ) or (z)
    const void* p[3];
    int n[3];
    int f64;              // sources are float64 (the reference DataLoader's dtype, train.py:211-215) else float32
};
__device__ __forceinline__ float tail_value(const TailSrc& s, int64_t r, int j) {
    int k = 0;
    while (k < 2 && j >= s.n[k]) { j -= s.n[k]; ++k; }
    const int64_t i = r * s.n[k] + j;
    return s.f64 ? (float)static_cast<const double*>(s.p[k])[i] : static_cast<const float*>(s.p[k])[i];
}
struct TailDst3 {
    __nv_bfloat16* p[3];  // destination matrices (row r of this call = row r of the destination)
    int64_t ld[3];
    int col0[3];
};
// One CTA per tail column: train-mode BatchNorm statistics of that column over the B rows (two passes over a tiny column), the
// normalised value WITHOUT gamma / beta (folded into the weights) written as bf16 into every destination, running statistics
// updated.  bn == 0: plain copy (the generator's z).  mode 1: statistics only -> stat [2, T] (local mean, M2);  mode 2: mean /
// invstd given.
__global__ void __launch_bounds__(256) tail_pack_kernel(const TailSrc src, int T, int64_t B, int bn, float eps, float count, float momentum,
                                                        float* __restrict__ mean, float* __restrict__ invstd, float* running_mean,
                                                        float* running_var, const TailDst3 dst, int mode, float* stat) {
    __shared__ float red[256];
    const int j = blockIdx.x;
    float mu = 0.f, is = 1.f;
    if (bn && mode != 2) {
        float s = 0.f;
        for (int64_t r = threadIdx.x; r < B; r += 256) s += tail_value(src, r, j);
        red[threadIdx.x] = s;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) {
            if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
            __syncthreads();
        }
        mu = red[0] / (float)B;
        __syncthreads();
        float q = 0.f;
        for (int64_t r = threadIdx.x; r < B; r += 256) {
            const float d = tail_value(src, r, j) - mu;
            q = fmaf(d, d, q);
        }
        red[threadIdx.x] = q;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) {
            if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
            __syncthreads();
        }
        const float m2 = red[0];
        if (mode == 1) {
            if (threadIdx.x == 0) {
                stat[j] = mu;
                stat[T + j] = m2;
            }
            return;
        }
        const float var = m2 / (float)B;
        is = rsqrtf(var + eps);
        if (threadIdx.x == 0) {
            mean[j] = mu;
            invstd[j] = is;
            if (running_mean) {
                running_mean[j] = (1.f - momentum) * running_mean[j] + momentum * mu;
                running_var[j] = (1.f - momentum) * running_var[j] + momentum * (count > 1.f ? var * (count / (count - 1.f)) : var);
            }
        }
    } else if (bn) {  // statistics merged over the ranks by the caller
        mu = mean[j];
        is = invstd[j];
        if (threadIdx.x == 0 && running_mean) {
            const float var = fmaxf(1.f / (is * is) - eps, 0.f);
            running_mean[j] = (1.f - momentum) * running_mean[j] + momentum * mu;
            running_var[j] = (1.f - momentum) * running_var[j] + momentum * (count > 1.f ? var * (count / (count - 1.f)) : var);
        }
    }
    for (int64_t r = threadIdx.x; r < B; r += 256) {
        const __nv_bfloat16 v = __float2bfloat16_rn((tail_value(src, r, j) - mu) * is);
#pragma unroll
        for (int d = 0; d < 3; ++d)
            if (dst.p[d]) dst.p[d][r * dst.ld[d] + dst.col0[d] + j] = v;
    }
}

// Backward of the tail's BatchNorm normalisation (generator step: the discriminator's pose inputs are the generator's outputs,
// dexgangrasp.py:800-816).  g_hat = gA[:, a0 + j] + gB[:, b0 + j] is the gradient wrt the normalised column (gamma is folded
// into the weights).  out [B, T] f32 = invstd * (g_hat - mean(g_hat) - xhat * mean(g_hat * xhat)).  One CTA per column.
__global__ void __launch_bounds__(256) tail_bn_bwd_kernel(const TailSrc src, int T, int64_t B, float count, const float* __restrict__ mean,
                                                          const float* __restrict__ invstd, const __nv_bfloat16* gA, int64_t lda, int a0,
                                                          const __nv_bfloat16* gB, int64_t ldb, int b0, float* __restrict__ out,
                                                          int mode, float* sums) {
    __shared__ float r0s[256], r1s[256];
    const int j = blockIdx.x;
    const float mu = mean[j], is = invstd[j];
    auto ghat = [&](int64_t r) {
        float g = __bfloat162float(gA[r * lda + a0 + j]);
        if (gB) g += __bfloat162float(gB[r * ldb + b0 + j]);
        return g;
    };
    float s0, s1;
    if (mode != 2) {
        float a = 0.f, b = 0.f;
        for (int64_t r = threadIdx.x; r < B; r += 256) {
            const float g = ghat(r);
            a += g;
            b = fmaf(g, (tail_value(src, r, j) - mu) * is, b);
        }
        r0s[threadIdx.x] = a;
        r1s[threadIdx.x] = b;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) {
            if (threadIdx.x < o) {
                r0s[threadIdx.x] += r0s[threadIdx.x + o];
                r1s[threadIdx.x] += r1s[threadIdx.x + o];
            }
            __syncthreads();
        }
        s0 = r0s[0];
        s1 = r1s[0];
        if (mode == 1) {
            if (threadIdx.x == 0) {
                sums[j] = s0;
                sums[T + j] = s1;
            }
            return;
        }
    } else {
        s0 = sums[j];
        s1 = sums[T + j];
    }
    for (int64_t r = threadIdx.x; r < B; r += 256) {
        const float xh = (tail_value(src, r, j) - mu) * is;
        out[r * T + j] = is * (ghat(r) - s0 / count - xh * s1 / count);
    }
}

// ------------------------------------------------------------------------------------------------ heads and losses
// Score head of the discriminator / evaluator: p = sigmoid(X . w + b), BCELoss (log clamped at -100, mean over the group's rows),
// dl = d(weighted loss)/d(logit) = (p - t) * dscale[g].  One warp per row, K = 512 bf16.
// target: labels[row] if labels else tconst[g].   loss[g] += -(t log p + (1-t) log(1-p)) * lscale[g];  db += sum dl.
struct ScoreHead {
    const __nv_bfloat16* X; int64_t ldx;
    const float *w, *b;
    int64_t rpg; int n_groups;
    const float* labels;
    float tconst[2], dscale[2], lscale[2];
    float *p, *dl, *loss, *db;
    int wass;  // WGAN variant (dexgangrasp.py:204-218, 234-270): loss = mean score, d loss / d logit = dscale * p (1 - p)
};
__global__ void __launch_bounds__(256) score_head_kernel(const ScoreHead a) {
    __shared__ float s_loss[8], s_db[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row = (int64_t)blockIdx.x * 8 + warp;
    const int64_t rows = a.rpg * a.n_groups;
    float loss = 0.f, d = 0.f;
    int g = 0;
    if (row < rows) {
        g = (int)(row / a.rpg);
        const __nv_bfloat16* x = a.X + row * a.ldx;
        float s = 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float v[8];
            unpack8(ld8(x + h * 256 + lane * 8), v);
            const float4 w0 = *reinterpret_cast<const float4*>(a.w + h * 256 + lane * 8), w1 = *reinterpret_cast<const float4*>(a.w + h * 256 + lane * 8 + 4);
            s = fmaf(v[0], w0.x, s); s = fmaf(v[1], w0.y, s); s = fmaf(v[2], w0.z, s); s = fmaf(v[3], w0.w, s);
            s = fmaf(v[4], w1.x, s); s = fmaf(v[5], w1.y, s); s = fmaf(v[6], w1.z, s); s = fmaf(v[7], w1.w, s);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const float pr = 1.f / (1.f + expf(-(s + a.b[0])));
        if (lane == 0) {
            const float t = a.labels ? a.labels[row] : a.tconst[g];
            if (a.wass) {
                loss = pr * a.lscale[g];
                d = a.dscale[g] * pr * (1.f - pr);
            } else {
                loss = -(t * fmaxf(logf(pr), -100.f) + (1.f - t) * fmaxf(logf(1.f - pr), -100.f)) * a.lscale[g];
                d = (pr - t) * a.dscale[g];
            }
            if (a.p) a.p[row] = pr;
            a.dl[row] = d;
        }
    }
    // rows of one CTA (8 consecutive rows) may straddle the two groups: reduce per group
    for (int gg = 0; gg < a.n_groups; ++gg) {
        if (lane == 0) {
            s_loss[warp] = (row < rows && g == gg) ? loss : 0.f;
            s_db[warp] = (row < rows && g == gg) ? d : 0.f;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            float l = 0.f, q = 0.f;
            for (int i = 0; i < 8; ++i) { l += s_loss[i]; q += s_db[i]; }
            if (l != 0.f) atomicAdd(a.loss + gg, l);
            if (a.db && q != 0.f) atomicAdd(a.db, q);
        }
        __syncthreads();
    }
}

// Generator heads -> fake grasp: out [B, ldo] f32 = (joint nJ | rot6d 6 | transl 3) -> rot [B,9] (columns x, y, z of the
// Gram-Schmidt), transl [B,3], joint [B,nJ]  (networks.py:98-103, dexgangrasp.py:776)
__global__ void gen_pose_kernel(const float* __restrict__ out, int64_t ldo, int nJ, int64_t B, float* __restrict__ rot,
                                float* __restrict__ transl, float* __restrict__ joint) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= B) return;
    const float* h = out + m * ldo;
    for (int j = 0; j < nJ; ++j) joint[m * nJ + j] = h[j];
    V3 x, y, z;
    rot6d_columns(V3{h[nJ], h[nJ + 1], h[nJ + 2]}, V3{h[nJ + 3], h[nJ + 4], h[nJ + 5]}, x, y, z);
    float* r = rot + m * 9;
    r[0] = x.x; r[1] = y.x; r[2] = z.x;
    r[3] = x.y; r[4] = y.y; r[5] = z.y;
    r[6] = x.z; r[7] = y.z; r[8] = z.z;
    for (int j = 0; j < 3; ++j) transl[m * 3 + j] = h[nJ + 6 + j];
}

// Generator loss terms and their gradient wrt the heads (compute_loss_dexgangrasp_generator, dexgangrasp.py:272-307;
// losses.transl_rot_6D_l2_loss, losses.py:61-76):  100 MSE(transl) + 1 MSE(vec R) + 10 MSE(joint), means over B_total rows, plus
// the gradient that came back through the discriminator (g_pose [B, 12 + nJ] wrt rot 9 | transl 3 | joint nJ; may be NULL).
//   g_heads [B, ldg] bf16 = (joint nJ | rot6d 6 | transl 3 | 0...);  loss[0..2] += transl, rot, conf terms;  dbias [ldg] += column sums
__global__ void __launch_bounds__(128) gen_loss_kernel(const float* __restrict__ out, int64_t ldo, int nJ, int64_t B, float n_total,
                                                       const TailSrc real /* rot 9 | transl 3 | joint nJ */, const float* __restrict__ g_pose,
                                                       float c_t, float c_r, float c_j, __nv_bfloat16* __restrict__ g_heads, int64_t ldg,
                                                       float* __restrict__ loss, float* __restrict__ dbias) {
    __shared__ float s_l[3][4];
    __shared__ float s_b[32];
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (threadIdx.x < 32) s_b[threadIdx.x] = 0.f;
    __syncthreads();
    float lt = 0.f, lr = 0.f, lj = 0.f;
    float gv[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) gv[i] = 0.f;
    const int T = 12 + nJ;
    if (m < B) {
        const float* h = out + m * ldo;
        const V3 a{h[nJ], h[nJ + 1], h[nJ + 2]}, b{h[nJ + 3], h[nJ + 4], h[nJ + 5]};
        V3 x, y, z;
        rot6d_columns(a, b, x, y, z);
        const float R[9] = {x.x, y.x, z.x, x.y, y.y, z.y, x.z, y.z, z.z};
        float gR[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) {
            const float d = R[i] - tail_value(real, m, i);
            lr = fmaf(d, d, lr);
            gR[i] = c_r * 2.f * d / (9.f * n_total) + (g_pose ? g_pose[m * T + i] : 0.f);
        }
        V3 ga, gb;
        rot6d_columns_bwd(a, b, V3{gR[0], gR[3], gR[6]}, V3{gR[1], gR[4], gR[7]}, V3{gR[2], gR[5], gR[8]}, ga, gb);
#pragma unroll
        for (int j = 0; j < 20; ++j) {
            if (j < nJ) {
                const float d = h[j] - tail_value(real, m, 12 + j);
                lj = fmaf(d, d, lj);
                gv[j] = c_j * 2.f * d / ((float)nJ * n_total) + (g_pose ? g_pose[m * T + 12 + j] : 0.f);
            }
        }
        const float g6[6] = {ga.x, ga.y, ga.z, gb.x, gb.y, gb.z};
        float g3[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const float d = h[nJ + 6 + j] - tail_value(real, m, 9 + j);
            lt = fmaf(d, d, lt);
            g3[j] = c_t * 2.f * d / (3.f * n_total) + (g_pose ? g_pose[m * T + 9 + j] : 0.f);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) {  // static indexing keeps gv in registers
            if (i >= nJ && i < nJ + 6) gv[i] = g6[i - nJ];
            if (i >= nJ + 6 && i < nJ + 9) gv[i] = g3[i - nJ - 6];
        }
        __nv_bfloat16* go = g_heads + m * ldg;
#pragma unroll
        for (int i = 0; i < 32; ++i)
            if (i < ldg) go[i] = __float2bfloat16_rn(gv[i]);
    }
    // head-bias gradients = column sums of gv: over the warp's 32 rows first, then one shared atomic per warp and column
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        float v = gv[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && i < nJ + 9) atomicAdd(&s_b[i], v);
    }
    lt *= c_t / (3.f * n_total);
    lr *= c_r / (9.f * n_total);
    lj *= c_j / ((float)nJ * n_total);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lt += __shfl_xor_sync(0xffffffffu, lt, o);
        lr += __shfl_xor_sync(0xffffffffu, lr, o);
        lj += __shfl_xor_sync(0xffffffffu, lj, o);
    }
    if ((threadIdx.x & 31) == 0) {
        s_l[0][threadIdx.x >> 5] = lt;
        s_l[1][threadIdx.x >> 5] = lr;
        s_l[2][threadIdx.x >> 5] = lj;
    }
    __syncthreads();
    if (threadIdx.x < 3) atomicAdd(loss + threadIdx.x, s_l[threadIdx.x][0] + s_l[threadIdx.x][1] + s_l[threadIdx.x][2] + s_l[threadIdx.x][3]);
    if (threadIdx.x < 32 && threadIdx.x < nJ + 9 && dbias) atomicAdd(dbias + threadIdx.x, s_b[threadIdx.x]);
}

// ------------------------------------------------------------------------------------------------ weights: bf16 operands <-> fp32 grads
// One ResBlock's first layer (fc1 [256, Fin] on top of fc3 [512, Fin] = 768 rows) in the layout the GEMMs use:
//   cmap[c] >= 0: input column c goes through the network's input BatchNorm (index into its gamma / beta), else -1
//   dmap[c] < 4096: column c is BPS column dmap[c] -> Wb [768, 4096];  else column dmap[c] - 4096 of Wxt [768, ldxt]
struct FirstLayer {
    const float *W1, *W3, *b1, *b3;  // fc1.weight [256, Fin], fc3.weight [512, Fin], biases
    int Fin;
    const int *cmap, *dmap;
    int bps_src0;                    // first input column of the 4096 BPS columns (they are contiguous; BPS column d has BatchNorm index d)
    const float *in_gamma, *in_beta;
    __nv_bfloat16 *Wb, *Wxt; int ldxt;
    float* bias13;                   // [768] = b + W[:, bn cols] . beta
    // gradients
    const float *dWb, *dWxt;         // fp32 [768, 4096], [768, ldxt]: gradients wrt the folded weights
    const float* db13;               // [768] column sums of the gradient wrt the layer output (fc1 part is exactly 0: BatchNorm follows)
    float *gW1, *gW3, *gb1, *gb3;    // .grad of fc1 / fc3 weight and bias
    float *g_in_gamma, *g_in_beta;   // .grad of the input BatchNorm (+=)
};
struct PlainJob {                    // dst[r, c] = bf16(src[r, c])  or  gdst[r, c] = gsrc[r, c] (fp32)
    const float* src; int64_t lds; int rows, cols;
    __nv_bfloat16* dst; int64_t ldd;
    const float* gsrc; int64_t ldgs; float* gdst; int64_t ldgd;
};
constexpr int TF_MAX_FL = 3, TF_MAX_PJ = 8;
struct WeightJobs {
    int n_fl, n_pj;
    FirstLayer fl[TF_MAX_FL];
    PlainJob pj[TF_MAX_PJ];
};

// grid.x = 768 * n_fl + sum of plain-job rows; one CTA per weight row
__global__ void __launch_bounds__(256) prep_weights_kernel(const WeightJobs J) {
    __shared__ float red[256];
    int row = blockIdx.x;
    if (row < 768 * J.n_fl) {
        const FirstLayer& L = J.fl[row / 768];
        const int n = row % 768;
        const float* w = n < 256 ? L.W1 + (int64_t)n * L.Fin : L.W3 + (int64_t)(n - 256) * L.Fin;
        float acc = 0.f;
        // the 4096 BPS columns: two adjacent destination columns per thread (bf16x2 stores), gamma / beta read as float2
        const float* wb = w + L.bps_src0;
        __nv_bfloat162* drow = reinterpret_cast<__nv_bfloat162*>(L.Wb + (int64_t)n * 4096);
#pragma unroll 4
        for (int i = threadIdx.x; i < 2048; i += 256) {
            const float w0 = wb[2 * i], w1 = wb[2 * i + 1];
            const float2 g = *reinterpret_cast<const float2*>(L.in_gamma + 2 * i), be = *reinterpret_cast<const float2*>(L.in_beta + 2 * i);
            acc = fmaf(w0, be.x, fmaf(w1, be.y, acc));
            drow[i] = __floats2bfloat162_rn(w0 * g.x, w1 * g.y);
        }
        // the other columns (previous block's X, per-row tail): through the maps
        for (int c = threadIdx.x; c < L.Fin - 4096; c += 256) {
            const int cc = c < L.bps_src0 ? c : c + 4096;
            const int bi = L.cmap[cc], d = L.dmap[cc];
            float v = w[cc];
            if (bi >= 0) {
                acc = fmaf(v, L.in_beta[bi], acc);
                v *= L.in_gamma[bi];
            }
            L.Wxt[(int64_t)n * L.ldxt + (d - 4096)] = __float2bfloat16_rn(v);
        }
        red[threadIdx.x] = acc;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) {
            if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x == 0) L.bias13[n] = (n < 256 ? L.b1[n] : L.b3[n - 256]) + red[0];
        return;
    }
    row -= 768 * J.n_fl;
    for (int j = 0; j < J.n_pj; ++j) {
        const PlainJob& P = J.pj[j];
        if (row < P.rows) {
            if (P.dst)
                for (int c = threadIdx.x; c < P.cols; c += 256) P.dst[(int64_t)row * P.ldd + c] = __float2bfloat16_rn(P.src[(int64_t)row * P.lds + c]);
            return;
        }
        row -= P.rows;
    }
}

// grid (ceil(Fin_max / 256), 24 row chunks of 32, n_fl) for the first layers + a tail of plain fp32 copies handled by blockIdx.z == n_fl
__global__ void __launch_bounds__(256) unfold_grads_kernel(const WeightJobs J) {
    if ((int)blockIdx.z == J.n_fl) {  // plain copies (head gradients out of the GEMM scratch)
        int row = blockIdx.y * gridDim.x + blockIdx.x;
        for (int j = 0; j < J.n_pj; ++j) {
            const PlainJob& P = J.pj[j];
            if (row < P.rows) {
                if (P.gdst)
                    for (int c = threadIdx.x; c < P.cols; c += 256) P.gdst[(int64_t)row * P.ldgd + c] = P.gsrc[(int64_t)row * P.ldgs + c];
                return;
            }
            row -= P.rows;
        }
        return;
    }
    const FirstLayer& L = J.fl[blockIdx.z];
    const int c = blockIdx.x * 256 + threadIdx.x;
    const int n0 = blockIdx.y * 32;
    if (c == 0 && blockIdx.x == 0) {  // biases: fc1's is exactly zero (its output is batch-normalised), fc3's is the column sum
        for (int n = n0; n < n0 + 32; ++n) {
            if (n < 256) L.gb1[n] = 0.f;
            else L.gb3[n - 256] = L.db13[n];
        }
    }
    if (c >= L.Fin) return;
    const int bi = L.cmap[c], d = L.dmap[c];
    const float gam = bi >= 0 ? L.in_gamma[bi] : 1.f, bet = bi >= 0 ? L.in_beta[bi] : 0.f;
    float dg = 0.f, dbt = 0.f;
#pragma unroll 8
    for (int n = n0; n < n0 + 32; ++n) {
        const float gf = d < 4096 ? L.dWb[(int64_t)n * 4096 + d] : L.dWxt[(int64_t)n * L.ldxt + (d - 4096)];
        const float dbn = n < 256 ? 0.f : L.db13[n];
        float* gw = n < 256 ? L.gW1 + (int64_t)n * L.Fin : L.gW3 + (int64_t)(n - 256) * L.Fin;
        gw[c] = gam * gf + bet * dbn;
        if (bi >= 0) {
            const float w = n < 256 ? L.W1[(int64_t)n * L.Fin + c] : L.W3[(int64_t)(n - 256) * L.Fin + c];
            dg = fmaf(w, gf, dg);
            dbt = fmaf(w, dbn, dbt);
        }
    }
    if (bi >= 0) {
        atomicAdd(L.g_in_gamma + bi, dg);
        atomicAdd(L.g_in_beta + bi, dbt);
    }
}

// ------------------------------------------------------------------------------------------------ running statistics
struct RunEntry {
    float *rm, *rv;
    const float *mean, *second;  // second: invstd (kind 0) or biased variance (kind 1), [n_groups, C] (stride C) or a single row reused
    long long* nbt;
    int C, n_groups, kind, same_row;  // same_row 1: every group uses row 0 (the shared BPS statistics applied `n_groups` times)
    float count, eps, momentum;
};
constexpr int TF_MAX_RUN = 24;
struct RunJobs {
    int n;
    RunEntry e[TF_MAX_RUN];
};
// nn.BatchNorm1d's buffer updates for every BatchNorm of one network pass in one launch (momentum 0.1, unbiased variance,
// num_batches_tracked += 1 per forward); groups (real batch, then fake batch) are applied in order.  grid (n entries, 8 chunks)
__global__ void bn_running_multi_kernel(const RunJobs J) {
    const RunEntry& E = J.e[blockIdx.x];
    for (int c = blockIdx.y * blockDim.x + threadIdx.x; c < E.C; c += gridDim.y * blockDim.x) {
        float rm = E.rm[c], rv = E.rv[c];
        for (int g = 0; g < E.n_groups; ++g) {
            const int64_t i = (E.same_row ? 0 : (int64_t)g * E.C) + c;
            const float mu = E.mean[i];
            const float var = E.kind ? E.second[i] : fmaxf(1.f / (E.second[i] * E.second[i]) - E.eps, 0.f);
            rm = (1.f - E.momentum) * rm + E.momentum * mu;
            rv = (1.f - E.momentum) * rv + E.momentum * (E.count > 1.f ? var * (E.count / (E.count - 1.f)) : var);
        }
        E.rm[c] = rm;
        E.rv[c] = rv;
    }
    if (blockIdx.y == 0 && threadIdx.x == 0 && E.nbt) *E.nbt += E.n_groups;
}

// merge per-rank (mean, M2) pairs gathered from `world` ranks (equal row counts n_local) with Chan's formula:
// gathered [world, n_sets, 2, C] -> mean, invstd [n_sets, C]
__global__ void bn_merge_kernel(const float* __restrict__ gathered, int world, int n_sets, int C, float n_local, float eps,
                                float* __restrict__ mean, float* __restrict__ invstd, float* __restrict__ var_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_sets * C) return;
    const int s = i / C, c = i % C;
    float mu = 0.f;
    for (int r = 0; r < world; ++r) mu += gathered[(((int64_t)r * n_sets + s) * 2 + 0) * C + c];
    mu /= (float)world;
    float m2 = 0.f;
    for (int r = 0; r < world; ++r) {
        const float d = gathered[(((int64_t)r * n_sets + s) * 2 + 0) * C + c] - mu;
        m2 += gathered[(((int64_t)r * n_sets + s) * 2 + 1) * C + c] + n_local * d * d;
    }
    const float var = m2 / (n_local * (float)world);
    mean[i] = mu;
    invstd[i] = rsqrtf(var + eps);
    if (var_out) var_out[i] = var;
}

// bumps a device-side 64-bit counter (dropout seed of the step)
__global__ void bump_counter_kernel(unsigned long long* c) { *c += 0x9e3779b97f4a7c15ull; }

static inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }

}  // namespace dexgg

using namespace dexgg;

#define TF_ST static_cast<cudaStream_t>(stream)

static int slab_comm(const dexgg_slab_comm& q, int mode, const float* acc, int n, SlabComm& c) {
    c.peers = reinterpret_cast<float* const*>(q.peers); c.world = q.world; c.rank = q.rank; c.slot = q.slot; c.parity = q.parity;
    if (mode == 0) {
        DEXGG_REQUIRE(q.peers && q.parity && q.world >= 1 && q.world <= SLAB_MAXW && q.rank >= 0 && q.rank < q.world && q.slot >= 0 &&
                      q.slot < SLAB_SLOTS && n <= SLAB_ACC, "slab kernel: bad exchange descriptor (world %d rank %d slot %d n %d)", q.world, q.rank, q.slot, n);
    } else {
        DEXGG_REQUIRE(acc, "slab kernel: accumulator scratch missing");
    }
    return DEXGG_OK;
}

// (column slabs, row splits, groups) with at most `per_sm` CTAs per SM in total (the occupancy __launch_bounds__ guarantees): every
// CTA of the grid is resident at once, which the in-kernel exchange needs; more CTAs = more loads in flight for these
// latency-bound streaming passes
static dim3 slab_grid(int C, int64_t rows_per_group, int n_groups, int per_sm) {
    const int slabs = C / 16;
    int rs = per_sm * sm_count() / (slabs * n_groups);
    const int64_t max_rs = (rows_per_group + 127) / 128;
    if (rs > max_rs) rs = (int)max_rs;
    if (rs < 1) rs = 1;
    return dim3(slabs, rs, n_groups);
}

extern "C" int dexgg_bn_fwd_slab(const dexgg_bn_fwd* q, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(q && q->x && q->C >= 16 && q->C % 16 == 0 && q->rows_per_group >= 1 && q->n_groups >= 1 && q->ldx % 8 == 0,
                  "bn_fwd_slab: bad arguments (C %% 16 == 0, row pitches %% 8 == 0)");
    DEXGG_REQUIRE(q->mode == 1 || (q->y && q->gamma && q->beta && q->mean && q->invstd && q->ldy % 8 == 0), "bn_fwd_slab: null pointer");
    DEXGG_REQUIRE(q->n_groups <= 2, "bn_fwd_slab: at most two batch groups");
    BnFwd a;
    a.x = static_cast<const __nv_bfloat16*>(q->x); a.ldx = q->ldx; a.C = q->C; a.rpg = q->rows_per_group; a.n_groups = q->n_groups;
    a.gamma = q->gamma; a.beta = q->beta; a.shift = q->shift; a.mean = q->mean; a.invstd = q->invstd; a.eps = q->eps;
    a.count = q->count_total > 0.f ? q->count_total : (float)q->rows_per_group;
    a.res = static_cast<const __nv_bfloat16*>(q->residual); a.ldr = q->ldr; a.act = q->act;
    a.mask = q->drop_mask; a.ldm = q->ldm; a.drop_p = q->drop_p; a.seed = q->drop_seed; a.stream = q->drop_stream;
    a.y = static_cast<__nv_bfloat16*>(q->y); a.ldy = q->ldy; a.mode = q->mode; a.acc = q->acc;
    if (int rc = slab_comm(q->comm, q->mode, q->acc, q->n_groups * 2 * q->C, a.comm)) return rc;
    bn_fwd_slab_kernel<<<slab_grid(q->C, q->rows_per_group, q->n_groups, (q->ctas_per_sm >= 1 && q->ctas_per_sm <= 3) ? q->ctas_per_sm : 3), 256, 0, TF_ST>>>(a);
    count_launches(1);
    DEXGG_LAUNCH_OK("bn_fwd_slab_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_bn_bwd_slab(const dexgg_bn_bwd* q, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(q && q->x && q->mean && q->invstd && q->gamma && q->C >= 16 && q->C % 16 == 0 && q->rows_per_group >= 1 && q->n_groups >= 1,
                  "bn_bwd_slab: bad arguments");
    DEXGG_REQUIRE(q->g || (q->dl && q->w), "bn_bwd_slab: no incoming gradient");
    DEXGG_REQUIRE(q->mode == 1 || q->gx, "bn_bwd_slab: null output");
    BnBwd a;
    a.g = static_cast<const __nv_bfloat16*>(q->g); a.ldg = q->ldg; a.dl = q->dl; a.w = q->w; a.dw_out = q->dw_out;
    a.out = static_cast<const __nv_bfloat16*>(q->out); a.ldo = q->ldo; a.act = q->act;
    a.mask = q->drop_mask; a.ldm = q->ldm; a.drop_p = q->drop_p; a.seed = q->drop_seed; a.stream = q->drop_stream;
    a.x = static_cast<const __nv_bfloat16*>(q->x); a.ldx = q->ldx; a.mean = q->mean; a.invstd = q->invstd; a.gamma = q->gamma;
    a.C = q->C; a.rpg = q->rows_per_group; a.n_groups = q->n_groups;
    a.count = q->count_total > 0.f ? q->count_total : (float)q->rows_per_group;
    a.dgamma = q->dgamma; a.dbeta = q->dbeta; a.gx = static_cast<__nv_bfloat16*>(q->gx); a.ldgx = q->ldgx;
    a.skip = static_cast<const __nv_bfloat16*>(q->skip); a.lds = q->lds; a.gskip = static_cast<__nv_bfloat16*>(q->gskip); a.ldgs = q->ldgs;
    a.dbias_skip = q->dbias_skip; a.mode = q->mode; a.acc = q->acc;
    if (int rc = slab_comm(q->comm, q->mode, q->acc, q->n_groups * 2 * q->C, a.comm)) return rc;
    bn_bwd_slab_kernel<<<slab_grid(q->C, q->rows_per_group, q->n_groups, (q->ctas_per_sm >= 1 && q->ctas_per_sm <= 2) ? q->ctas_per_sm : 2), 256, 0, TF_ST>>>(a);
    count_launches(1);
    DEXGG_LAUNCH_OK("bn_bwd_slab_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_bn_input_stats(const float* x, int64_t ldx, int64_t B, int C, float eps, float* mean, float* var, float* invstd,
                                    float* acc, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(x && mean && var && invstd && acc && B >= 1 && C >= 4 && C % 4 == 0 && ldx % 4 == 0, "bn_input_stats: bad arguments");
    const int gx = (C + 127) / 128;
    int rs = 4 * sm_count() / gx;
    if (rs > (B + 31) / 32) rs = (int)((B + 31) / 32);
    if (rs < 1) rs = 1;
    bn_input_stats_kernel<<<dim3(gx, rs), 256, 0, TF_ST>>>(x, ldx, B, C, acc);
    bn_input_finalize_kernel<<<nblk(C), 256, 0, TF_ST>>>(x, acc, B, C, eps, mean, var, invstd);
    count_launches(2);
    DEXGG_LAUNCH_OK("bn_input_stats kernels");
    return DEXGG_OK;
}

extern "C" int dexgg_xhat(const float* x, int64_t ldx, const float* mean, const float* invstd, int64_t B, int C, void* out_bf16, int copies,
                          void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(x && mean && invstd && out_bf16 && B >= 1 && C % 4 == 0 && ldx % 4 == 0 && copies >= 1, "xhat: bad arguments");
    xhat_kernel<<<nblk(B * (C / 4)), 256, 0, TF_ST>>>(x, ldx, mean, invstd, B, C, static_cast<__nv_bfloat16*>(out_bf16), copies);
    count_launches(1);
    DEXGG_LAUNCH_OK("xhat_kernel");
    return DEXGG_OK;
}

static TailSrc tail_src(const dexgg_tail_src* s) {
    TailSrc t;
    for (int i = 0; i < 3; ++i) { t.p[i] = s->p[i]; t.n[i] = s->n[i]; }
    t.f64 = s->f64;
    return t;
}

extern "C" int dexgg_tail_pack(const dexgg_tail_src* src, int64_t B, int bn, float eps, float count_total, float momentum, float* mean,
                               float* invstd, float* running_mean, float* running_var, void* const* dst, const int64_t* dst_ld,
                               const int* dst_col0, int mode, float* stat, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(src && dst && dst_ld && dst_col0 && B >= 1, "tail_pack: bad arguments");
    const int T = src->n[0] + src->n[1] + src->n[2];
    DEXGG_REQUIRE(T >= 1 && (!bn || (mean && invstd)), "tail_pack: bad sizes");
    TailDst3 d;
    for (int i = 0; i < 3; ++i) { d.p[i] = static_cast<__nv_bfloat16*>(dst[i]); d.ld[i] = dst_ld[i]; d.col0[i] = dst_col0[i]; }
    tail_pack_kernel<<<T, 256, 0, TF_ST>>>(tail_src(src), T, B, bn, eps, count_total > 0.f ? count_total : (float)B, momentum, mean, invstd,
                                           running_mean, running_var, d, mode, stat);
    count_launches(1);
    DEXGG_LAUNCH_OK("tail_pack_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_tail_bn_bwd(const dexgg_tail_src* src, int64_t B, float count_total, const float* mean, const float* invstd,
                                 const void* gA, int64_t lda, int a0, const void* gB, int64_t ldb, int b0, float* out, int mode, float* sums,
                                 void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(src && mean && invstd && gA && B >= 1 && (mode == 1 || out), "tail_bn_bwd: bad arguments");
    const int T = src->n[0] + src->n[1] + src->n[2];
    tail_bn_bwd_kernel<<<T, 256, 0, TF_ST>>>(tail_src(src), T, B, count_total > 0.f ? count_total : (float)B, mean, invstd,
                                             static_cast<const __nv_bfloat16*>(gA), lda, a0, static_cast<const __nv_bfloat16*>(gB), ldb, b0, out,
                                             mode, sums);
    count_launches(1);
    DEXGG_LAUNCH_OK("tail_bn_bwd_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_score_head(const dexgg_score_head_args* q, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(q && q->X && q->w && q->b && q->dl && q->loss && q->rows_per_group >= 1 && q->n_groups >= 1 && q->n_groups <= 2 && q->ldx % 8 == 0,
                  "score_head: bad arguments");
    ScoreHead a;
    a.X = static_cast<const __nv_bfloat16*>(q->X); a.ldx = q->ldx; a.w = q->w; a.b = q->b; a.rpg = q->rows_per_group; a.n_groups = q->n_groups;
    a.labels = q->labels;
    for (int i = 0; i < 2; ++i) { a.tconst[i] = q->target[i]; a.dscale[i] = q->dscale[i]; a.lscale[i] = q->lscale[i]; }
    a.p = q->p; a.dl = q->dl; a.loss = q->loss; a.db = q->db; a.wass = q->wass;
    score_head_kernel<<<nblk(q->rows_per_group * q->n_groups, 8), 256, 0, TF_ST>>>(a);
    count_launches(1);
    DEXGG_LAUNCH_OK("score_head_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_gen_pose(const float* out, int64_t ldo, int n_joints, int64_t B, float* rot, float* transl, float* joint, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(out && rot && transl && joint && B >= 1 && n_joints >= 1 && n_joints <= 20, "gen_pose: bad arguments");
    gen_pose_kernel<<<nblk(B, 128), 128, 0, TF_ST>>>(out, ldo, n_joints, B, rot, transl, joint);
    count_launches(1);
    DEXGG_LAUNCH_OK("gen_pose_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_gen_loss(const float* out, int64_t ldo, int n_joints, int64_t B, float n_total, const dexgg_tail_src* real,
                              const float* g_pose, float c_transl, float c_rot, float c_joint, void* g_heads, int64_t ldg, float* loss,
                              float* dbias, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(out && real && g_heads && loss && B >= 1 && n_joints >= 1 && n_joints <= 20 && ldg >= n_joints + 9 && ldg <= 32, "gen_loss: bad arguments");
    gen_loss_kernel<<<nblk(B, 128), 128, 0, TF_ST>>>(out, ldo, n_joints, B, n_total, tail_src(real), g_pose, c_transl, c_rot, c_joint,
                                                     static_cast<__nv_bfloat16*>(g_heads), ldg, loss, dbias);
    count_launches(1);
    DEXGG_LAUNCH_OK("gen_loss_kernel");
    return DEXGG_OK;
}

static int weight_jobs(const dexgg_weight_jobs* q, WeightJobs& J) {
    DEXGG_REQUIRE(q && q->n_first >= 0 && q->n_first <= TF_MAX_FL && q->n_plain >= 0 && q->n_plain <= TF_MAX_PJ, "weight jobs: bad counts");
    J.n_fl = q->n_first; J.n_pj = q->n_plain;
    for (int i = 0; i < q->n_first; ++i) {
        const dexgg_first_layer& s = q->first[i];
        FirstLayer& L = J.fl[i];
        L.W1 = s.W1; L.W3 = s.W3; L.b1 = s.b1; L.b3 = s.b3; L.Fin = s.Fin; L.cmap = s.cmap; L.dmap = s.dmap; L.bps_src0 = s.bps_src0; L.in_gamma = s.in_gamma;
        DEXGG_REQUIRE(s.bps_src0 >= 0 && s.bps_src0 + 4096 <= s.Fin, "weight jobs: bad bps_src0");
        L.in_beta = s.in_beta; L.Wb = static_cast<__nv_bfloat16*>(s.Wb); L.Wxt = static_cast<__nv_bfloat16*>(s.Wxt); L.ldxt = s.ldxt;
        L.bias13 = s.bias13; L.dWb = s.dWb; L.dWxt = s.dWxt; L.db13 = s.db13; L.gW1 = s.gW1; L.gW3 = s.gW3; L.gb1 = s.gb1; L.gb3 = s.gb3;
        L.g_in_gamma = s.g_in_gamma; L.g_in_beta = s.g_in_beta;
    }
    for (int i = 0; i < q->n_plain; ++i) {
        const dexgg_plain_job& s = q->plain[i];
        PlainJob& P = J.pj[i];
        P.src = s.src; P.lds = s.lds; P.rows = s.rows; P.cols = s.cols; P.dst = static_cast<__nv_bfloat16*>(s.dst); P.ldd = s.ldd;
        P.gsrc = s.gsrc; P.ldgs = s.ldgs; P.gdst = s.gdst; P.ldgd = s.ldgd;
    }
    return DEXGG_OK;
}

extern "C" int dexgg_prep_weights(const dexgg_weight_jobs* q, void* stream) {
    DEXGG_ARCH_OK();
    WeightJobs J;
    int rc = weight_jobs(q, J);
    if (rc) return rc;
    int rows = 768 * J.n_fl;
    for (int i = 0; i < J.n_pj; ++i) rows += J.pj[i].rows;
    if (rows == 0) return DEXGG_OK;
    prep_weights_kernel<<<rows, 256, 0, TF_ST>>>(J);
    count_launches(1);
    DEXGG_LAUNCH_OK("prep_weights_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_unfold_grads(const dexgg_weight_jobs* q, void* stream) {
    DEXGG_ARCH_OK();
    WeightJobs J;
    int rc = weight_jobs(q, J);
    if (rc) return rc;
    int fin_max = 256, prow = 0;
    for (int i = 0; i < J.n_fl; ++i) fin_max = J.fl[i].Fin > fin_max ? J.fl[i].Fin : fin_max;
    for (int i = 0; i < J.n_pj; ++i) prow += J.pj[i].rows;
    const unsigned gx = nblk(fin_max);
    DEXGG_REQUIRE(prow <= (int)gx * 24, "unfold_grads: too many plain rows");
    unfold_grads_kernel<<<dim3(gx, 24, J.n_fl + (J.n_pj ? 1 : 0)), 256, 0, TF_ST>>>(J);
    count_launches(1);
    DEXGG_LAUNCH_OK("unfold_grads_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_bn_running_multi(const dexgg_run_entry* e, int n, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(e && n >= 1 && n <= TF_MAX_RUN, "bn_running_multi: 1..%d entries", TF_MAX_RUN);
    RunJobs J;
    J.n = n;
    for (int i = 0; i < n; ++i) {
        RunEntry& E = J.e[i];
        E.rm = e[i].running_mean; E.rv = e[i].running_var; E.mean = e[i].mean; E.second = e[i].second; E.nbt = e[i].num_batches_tracked;
        E.C = e[i].C; E.n_groups = e[i].n_groups; E.kind = e[i].second_is_var; E.same_row = e[i].same_row; E.count = e[i].count;
        E.eps = e[i].eps; E.momentum = e[i].momentum;
        DEXGG_REQUIRE(E.rm && E.rv && E.mean && E.second && E.C >= 1 && E.n_groups >= 1, "bn_running_multi: bad entry %d", i);
    }
    bn_running_multi_kernel<<<dim3(n, 8), 256, 0, TF_ST>>>(J);
    count_launches(1);
    DEXGG_LAUNCH_OK("bn_running_multi_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_bn_merge(const float* gathered, int world, int n_sets, int C, float n_local, float eps, float* mean, float* invstd,
                              float* var, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(gathered && mean && invstd && world >= 1 && n_sets >= 1 && C >= 1, "bn_merge: bad arguments");
    bn_merge_kernel<<<nblk((int64_t)n_sets * C), 256, 0, TF_ST>>>(gathered, world, n_sets, C, n_local, eps, mean, invstd, var);
    count_launches(1);
    DEXGG_LAUNCH_OK("bn_merge_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_bump_counter(unsigned long long* counter, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(counter, "bump_counter: null");
    bump_counter_kernel<<<1, 1, 0, TF_ST>>>(counter);
    count_launches(1);
    DEXGG_LAUNCH_OK("bump_counter_kernel");
    return DEXGG_OK;
}

// ---- peer-mapped scratch of the in-kernel exchange (one per rank; handles travel through torch.distributed) ----
extern "C" size_t dexgg_slab_scratch_bytes(void) { return (size_t)2 * SLAB_SLOTS * SLAB_SLOT_FLOATS * sizeof(float); }

extern "C" int dexgg_peer_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(ptr && handle64 && bytes > 0, "peer_alloc: bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    DEXGG_CUDA_OK(cudaMalloc(ptr, bytes));
    DEXGG_CUDA_OK(cudaMemset(*ptr, 0, bytes));
    cudaIpcMemHandle_t h;
    DEXGG_CUDA_OK(cudaIpcGetMemHandle(&h, *ptr));
    memcpy(handle64, &h, 64);
    return DEXGG_OK;
}
extern "C" int dexgg_peer_open(const unsigned char* handle64, void** ptr) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(ptr && handle64, "peer_open: bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    DEXGG_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return DEXGG_OK;
}
extern "C" int dexgg_peer_close(void* ptr) {
    if (ptr) DEXGG_CUDA_OK(cudaIpcCloseMemHandle(ptr));
    return DEXGG_OK;
}
extern "C" int dexgg_peer_free(void* ptr) {
    if (ptr) DEXGG_CUDA_OK(cudaFree(ptr));
    return DEXGG_OK;
}
// sum (in place if sum_out == src) and / or gather n floats across the ranks through the peer scratch; uses ceil(n / 2048) slots
extern "C" int dexgg_p2p_small(const dexgg_slab_comm* q, const float* src, int n, float* sum_out, float* gather_out, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(q && src && n >= 1 && (sum_out || gather_out), "p2p_small: bad arguments");
    const int nb = (n + SLAB_ACC - 1) / SLAB_ACC;
    SlabComm c;
    if (int rc = slab_comm(*q, 0, nullptr, 1, c)) return rc;
    DEXGG_REQUIRE(q->slot + nb <= SLAB_SLOTS, "p2p_small: out of exchange slots");
    p2p_small_kernel<<<nb, 256, 0, TF_ST>>>(c, src, n, sum_out, gather_out);
    count_launches(1);
    DEXGG_LAUNCH_OK("p2p_small_kernel");
    return DEXGG_OK;
}

// in-place mean of n floats (n % 4 == 0) over the ranks; bufs: DEVICE array [world] of every rank's peer-mapped buffer
extern "C" int dexgg_grad_allreduce(const dexgg_slab_comm* q, void* const* bufs, long long n, float scale, int ctas, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(q && bufs && n >= 4 && n % 4 == 0, "grad_allreduce: bad arguments (n %lld)", n);
    SlabComm c;
    if (int rc = slab_comm(*q, 0, nullptr, 1, c)) return rc;
    DEXGG_REQUIRE(c.world >= 2, "grad_allreduce: world %d (nothing to exchange)", c.world);
    if (ctas <= 0) ctas = 64;
    grad_allreduce_kernel<<<ctas, 256, 0, TF_ST>>>(c, reinterpret_cast<float4* const*>(bufs), n / 4, scale);
    count_launches(1);
    DEXGG_LAUNCH_OK("grad_allreduce_kernel");
    return DEXGG_OK;
}

// start of a training step: flip the parity word, zero the other parity buffer for the next step
extern "C" int dexgg_slab_step_begin(void* scratch_base, unsigned* parity, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(scratch_base && parity, "slab_step_begin: null");
    slab_step_begin_kernel<<<64, 256, 0, TF_ST>>>(static_cast<float*>(scratch_base), parity);
    slab_parity_flip_kernel<<<1, 1, 0, TF_ST>>>(parity);
    count_launches(2);
    DEXGG_LAUNCH_OK("slab_step_begin kernels");
    return DEXGG_OK;
}
