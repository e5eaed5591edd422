// Stage 4 — score threshold + sort / top-k per object, and the gather of the kept grasps.
// Replaces `p.sort(descending=True)`, `indices[sorted_score > thresh]` and the per-key torch.gather of
// DexGanGrasp.filter_grasps (DexGanGrasp/models/dexgangrasp.py:412-435) — per segment (object) instead of one object per
// Python call.  Order: score descending, equal scores by the lower row index (torch leaves that order unspecified).
//
//   segments of <= 16384 rows   one CTA per segment: 64-bit keys (~orderable(score) << 32 | local index) bitonic-sorted in
//                               shared memory; the same CTA then gathers the kept rows' rot / transl / joint_conf into one
//                               packed [k, 14 + nJ] record block (dexgg_filter_topk_gather) — the whole stage is one launch
//   larger segments             keys in a caller-provided workspace: 16384-key chunks are sorted in shared memory, the
//                               bitonic merge steps with stride >= 8192 run in global memory (dexgg_filter_topk_ws)
// HBM-bound in principle (4 B read per grasp); at the sizes of this path it is launch/latency-bound and negligible.
#include "common.cuh"

namespace dexgg {

constexpr int FILTER_MAX_SEG = 16384;  // rows per segment that fit the shared-memory sort (128 KB of keys)

__device__ __forceinline__ uint32_t orderable(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float unorderable(uint32_t o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

struct GatherArgs {           // optional: gather of the kept rows (all pointers null = off)
    const float* rot;         // [M,9]
    const float* transl;      // [M,3]
    const float* joint;       // [M,nJ]
    int nJ;
    float* packed;            // [n_seg, k, 14 + nJ]: score | row id (int32 bits) | rot 9 | transl 3 | joint nJ; rows >= count: id -1, rest 0
};

__global__ void __launch_bounds__(256)
filter_sort_kernel(const float* __restrict__ p, int64_t rows_per_seg, int64_t M, float thresh, int k, int npow2,
                   int32_t* __restrict__ idx_out, int32_t* __restrict__ count_out, float* __restrict__ best_out,
                   float* __restrict__ score_out, const GatherArgs g) {
    extern __shared__ unsigned long long keys[];
    __shared__ int s_count;
    __shared__ uint32_t s_best;
    const int seg = blockIdx.x;
    const int64_t r0 = (int64_t)seg * rows_per_seg;
    const int n = (int)((M - r0 < rows_per_seg) ? (M - r0) : rows_per_seg);
    if (threadIdx.x == 0) {
        s_count = 0;
        s_best = 0;
    }
    __syncthreads();
    int local = 0;
    uint32_t lbest = 0;
    for (int i = threadIdx.x; i < npow2; i += blockDim.x) {
        unsigned long long key = ~0ull;
        if (i < n) {
            const float v = p[r0 + i];
            const uint32_t o = orderable(v);
            lbest = o > lbest ? o : lbest;
            if (v > thresh) {  // strict, like the reference
                key = ((unsigned long long)(~o) << 32) | (uint32_t)i;
                ++local;
            }
        }
        keys[i] = key;
    }
    atomicAdd(&s_count, local);
    atomicMax(&s_best, lbest);
    __syncthreads();
    for (int size = 2; size <= npow2; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = threadIdx.x; i < npow2 / 2; i += blockDim.x) {
                const int lo = 2 * i - (i & (stride - 1));
                const int hi = lo + stride;
                const bool up = (lo & size) == 0;
                const unsigned long long a = keys[lo], b = keys[hi];
                if ((a > b) == up) {
                    keys[lo] = b;
                    keys[hi] = a;
                }
            }
            __syncthreads();
        }
    }
    const int kept = s_count < k ? s_count : k;
    for (int i = threadIdx.x; i < k; i += blockDim.x) {
        int32_t id = -1;
        float sc = 0.f;
        if (i < kept) {
            const int li = (int)(keys[i] & 0xffffffffu);
            id = (int32_t)(r0 + li);
            sc = p[r0 + li];
        }
        idx_out[(int64_t)seg * k + i] = id;
        if (score_out) score_out[(int64_t)seg * k + i] = sc;
    }
    if (g.packed) {  // gather: one thread per (kept row, record column)
        const int W = 14 + g.nJ;
        float* out = g.packed + (int64_t)seg * k * W;
        for (int e = threadIdx.x; e < k * W; e += blockDim.x) {
            const int i = e / W, c = e % W;
            float v = 0.f;
            if (i < kept) {
                const int64_t row = r0 + (int64_t)(keys[i] & 0xffffffffu);
                if (c == 0) v = p[row];
                else if (c == 1) v = __int_as_float((int32_t)row);
                else if (c < 11) v = g.rot[row * 9 + (c - 2)];
                else if (c < 14) v = g.transl[row * 3 + (c - 11)];
                else v = g.joint[row * g.nJ + (c - 14)];
            } else if (c == 1) {
                v = __int_as_float(-1);
            }
            out[e] = v;
        }
    }
    if (threadIdx.x == 0) {
        count_out[seg] = kept;
        if (best_out) best_out[seg] = n > 0 ? unorderable(s_best) : -__int_as_float(0x7f800000);
    }
}

// ---------------------------------------------------------------- large segments: keys in global memory
// keys[seg, npow2]; counts[seg] = rows above thresh; best[seg] = orderable(best score)
__global__ void big_keys_kernel(const float* __restrict__ p, int64_t rows_per_seg, int64_t M, float thresh, int64_t npow2,
                                unsigned long long* __restrict__ keys, int* __restrict__ counts, uint32_t* __restrict__ best) {
    const int seg = blockIdx.y;
    const int64_t r0 = (int64_t)seg * rows_per_seg;
    const int64_t n = (M - r0 < rows_per_seg) ? (M - r0) : rows_per_seg;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npow2) return;
    unsigned long long key = ~0ull;
    if (i < n) {
        const float v = p[r0 + i];
        const uint32_t o = orderable(v);
        atomicMax(best + seg, o);
        if (v > thresh) {
            key = ((unsigned long long)(~o) << 32) | (uint32_t)i;
            atomicAdd(counts + seg, 1);
        }
    }
    keys[(int64_t)seg * npow2 + i] = key;
}
// one (size, stride) compare-exchange step over the whole key array, stride >= FILTER_MAX_SEG / 2
__global__ void big_step_kernel(unsigned long long* __restrict__ keys, int64_t npow2, int64_t size, int64_t stride) {
    unsigned long long* k = keys + (int64_t)blockIdx.y * npow2;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npow2 / 2) return;
    const int64_t lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
    const bool up = (lo & size) == 0;
    const unsigned long long a = k[lo], b = k[hi];
    if ((a > b) == up) {
        k[lo] = b;
        k[hi] = a;
    }
}
// all steps with stride < FILTER_MAX_SEG / 2 ... 1 of bitonic stage `size` inside one 16384-key chunk (first = 1: the whole
// local sort, sizes 2 .. 16384)
__global__ void __launch_bounds__(256) big_chunk_kernel(unsigned long long* __restrict__ keys, int64_t npow2, int64_t size, int first) {
    extern __shared__ unsigned long long sk[];
    const int64_t c0 = (int64_t)blockIdx.x * FILTER_MAX_SEG;
    unsigned long long* k = keys + (int64_t)blockIdx.y * npow2 + c0;
    for (int i = threadIdx.x; i < FILTER_MAX_SEG; i += blockDim.x) sk[i] = k[i];
    __syncthreads();
    auto steps = [&](int64_t sz, int s_hi) {
        for (int stride = s_hi; stride > 0; stride >>= 1) {
            for (int i = threadIdx.x; i < FILTER_MAX_SEG / 2; i += blockDim.x) {
                const int lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
                const bool up = ((c0 + lo) & sz) == 0;
                const unsigned long long a = sk[lo], b = sk[hi];
                if ((a > b) == up) {
                    sk[lo] = b;
                    sk[hi] = a;
                }
            }
            __syncthreads();
        }
    };
    if (first) {
        for (int sz = 2; sz <= FILTER_MAX_SEG; sz <<= 1) steps(sz, sz >> 1);
    } else {
        steps(size, FILTER_MAX_SEG / 2);
    }
    for (int i = threadIdx.x; i < FILTER_MAX_SEG; i += blockDim.x) k[i] = sk[i];
}
__global__ void big_out_kernel(const float* __restrict__ p, int64_t rows_per_seg, const unsigned long long* __restrict__ keys,
                               int64_t npow2, const int* __restrict__ counts, const uint32_t* __restrict__ best, int k,
                               int32_t* __restrict__ idx_out, int32_t* __restrict__ count_out, float* __restrict__ best_out,
                               float* __restrict__ score_out) {
    const int seg = blockIdx.y;
    const int64_t r0 = (int64_t)seg * rows_per_seg;
    const int kept = counts[seg] < k ? counts[seg] : k;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < k) {
        int32_t id = -1;
        float sc = 0.f;
        if (i < kept) {
            const int64_t li = (int64_t)(keys[(int64_t)seg * npow2 + i] & 0xffffffffu);
            id = (int32_t)(r0 + li);
            sc = p[r0 + li];
        }
        idx_out[(int64_t)seg * k + i] = id;
        if (score_out) score_out[(int64_t)seg * k + i] = sc;
    }
    if (i == 0) {
        count_out[seg] = kept;
        if (best_out) best_out[seg] = unorderable(best[seg]);
    }
}

static int64_t pow2_at_least(int64_t n) {
    int64_t v = 2;
    while (v < n) v <<= 1;
    return v;
}

static int launch_small(const float* p, int64_t M, int64_t rows_per_seg, float thresh, int k, int32_t* idx_out, int32_t* count_out,
                        float* best_out, float* score_out, const GatherArgs& g, cudaStream_t st) {
    const int64_t seg_rows = rows_per_seg < M ? rows_per_seg : M;
    const int64_t nseg = (M + rows_per_seg - 1) / rows_per_seg;
    const int npow2 = (int)pow2_at_least(seg_rows);
    const size_t smem = (size_t)npow2 * sizeof(unsigned long long);
    static std::atomic<unsigned long long> attr_mask{0};
    DEXGG_SMEM_OPT_IN(filter_sort_kernel, FILTER_MAX_SEG * 8, attr_mask);
    filter_sort_kernel<<<(unsigned)nseg, 256, smem, st>>>(p, rows_per_seg, M, thresh, k, npow2, idx_out, count_out, best_out, score_out, g);
    count_launches(1);
    DEXGG_LAUNCH_OK("filter_sort_kernel");
    return DEXGG_OK;
}

}  // namespace dexgg

using namespace dexgg;

extern "C" size_t dexgg_filter_topk_workspace_bytes(int64_t M, int64_t rows_per_seg) {
    if (M <= 0 || rows_per_seg < 1) return 0;
    const int64_t seg_rows = rows_per_seg < M ? rows_per_seg : M;
    if (seg_rows <= FILTER_MAX_SEG) return 0;
    const int64_t nseg = (M + rows_per_seg - 1) / rows_per_seg;
    return (size_t)nseg * (size_t)pow2_at_least(seg_rows) * 8 + align_up((size_t)nseg * 8, 256);
}

extern "C" int dexgg_filter_topk_ws(const float* p, int64_t M, int64_t rows_per_seg, float thresh, int k, int32_t* idx_out,
                                    int32_t* count_out, float* best_out, float* score_out, void* workspace, size_t workspace_bytes,
                                    void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && rows_per_seg >= 1 && k >= 1, "filter_topk: bad sizes");
    if (M == 0) return DEXGG_OK;
    DEXGG_REQUIRE(p && idx_out && count_out, "filter_topk: null pointer");
    DEXGG_REQUIRE(k <= rows_per_seg || k <= M, "filter_topk: k larger than a segment");
    DEXGG_REQUIRE(M < (1ll << 31), "filter_topk: row ids are int32 (M = %lld)", (long long)M);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t seg_rows = rows_per_seg < M ? rows_per_seg : M;
    if (seg_rows <= FILTER_MAX_SEG) return launch_small(p, M, rows_per_seg, thresh, k, idx_out, count_out, best_out, score_out, GatherArgs{}, st);
    const size_t need = dexgg_filter_topk_workspace_bytes(M, rows_per_seg);
    if (!workspace || workspace_bytes < need) {
        set_error("filter_topk: segments of %lld rows need a workspace of %zu bytes, %zu given", (long long)seg_rows, need, workspace_bytes);
        return DEXGG_E_WORKSPACE;
    }
    const int64_t nseg = (M + rows_per_seg - 1) / rows_per_seg, npow2 = pow2_at_least(seg_rows);
    DEXGG_REQUIRE(nseg <= 65535, "filter_topk: too many large segments");
    int* counts = static_cast<int*>(workspace);
    uint32_t* best = reinterpret_cast<uint32_t*>(counts + nseg);
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(static_cast<char*>(workspace) + align_up((size_t)nseg * 8, 256));
    DEXGG_CUDA_OK(cudaMemsetAsync(workspace, 0, (size_t)nseg * 8, st));
    static std::atomic<unsigned long long> attr_mask{0};
    DEXGG_SMEM_OPT_IN(big_chunk_kernel, FILTER_MAX_SEG * 8, attr_mask);
    const dim3 g_all((unsigned)((npow2 + 255) / 256), (unsigned)nseg), g_half((unsigned)((npow2 / 2 + 255) / 256), (unsigned)nseg);
    const dim3 g_chunk((unsigned)(npow2 / FILTER_MAX_SEG), (unsigned)nseg);
    int launches = 2;
    big_keys_kernel<<<g_all, 256, 0, st>>>(p, rows_per_seg, M, thresh, npow2, keys, counts, best);
    big_chunk_kernel<<<g_chunk, 256, FILTER_MAX_SEG * 8, st>>>(keys, npow2, 0, 1);
    for (int64_t size = 2 * FILTER_MAX_SEG; size <= npow2; size <<= 1) {
        for (int64_t stride = size >> 1; stride >= FILTER_MAX_SEG; stride >>= 1) {
            big_step_kernel<<<g_half, 256, 0, st>>>(keys, npow2, size, stride);
            ++launches;
        }
        big_chunk_kernel<<<g_chunk, 256, FILTER_MAX_SEG * 8, st>>>(keys, npow2, size, 0);
        ++launches;
    }
    big_out_kernel<<<dim3((unsigned)((k + 255) / 256), (unsigned)nseg), 256, 0, st>>>(p, rows_per_seg, keys, npow2, counts, best, k, idx_out,
                                                                                     count_out, best_out, score_out);
    count_launches(launches + 1);
    DEXGG_LAUNCH_OK("filter_topk large-segment kernels");
    return DEXGG_OK;
}

extern "C" int dexgg_filter_topk(const float* p, int64_t M, int64_t rows_per_seg, float thresh, int k, int32_t* idx_out,
                                 int32_t* count_out, float* best_out, float* score_out, void* stream) {
    const int64_t seg_rows = rows_per_seg < M ? rows_per_seg : M;
    if (M > 0 && rows_per_seg >= 1 && seg_rows > FILTER_MAX_SEG) {
        DEXGG_ARCH_OK();
        set_error("filter_topk: segments of more than %d rows need dexgg_filter_topk_ws (workspace)", FILTER_MAX_SEG);
        return DEXGG_E_WORKSPACE;
    }
    return dexgg_filter_topk_ws(p, M, rows_per_seg, thresh, k, idx_out, count_out, best_out, score_out, nullptr, 0, stream);
}

// The whole stage in ONE launch for segments of at most 16384 rows: threshold + per-segment top-k + gather of the kept rows'
// poses into packed [n_seg, k, 14 + nJ] records (score | row id as int32 bits | rot 9 | transl 3 | joint nJ).
extern "C" int dexgg_filter_topk_gather(const float* p, int64_t M, int64_t rows_per_seg, float thresh, int k, const float* rot,
                                        const float* transl, const float* joint, int n_joints, int32_t* idx_out,
                                        int32_t* count_out, float* packed_out, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && rows_per_seg >= 1 && k >= 1 && n_joints >= 0, "filter_topk_gather: bad sizes");
    if (M == 0) return DEXGG_OK;
    DEXGG_REQUIRE(p && idx_out && count_out && packed_out && rot && transl && (n_joints == 0 || joint), "filter_topk_gather: null pointer");
    DEXGG_REQUIRE(k <= rows_per_seg || k <= M, "filter_topk_gather: k larger than a segment");
    DEXGG_REQUIRE(M < (1ll << 31), "filter_topk_gather: row ids are int32");
    const int64_t seg_rows = rows_per_seg < M ? rows_per_seg : M;
    DEXGG_REQUIRE(seg_rows <= FILTER_MAX_SEG, "filter_topk_gather: at most %d rows per segment", FILTER_MAX_SEG);
    GatherArgs g{rot, transl, joint, n_joints, packed_out};
    return launch_small(p, M, rows_per_seg, thresh, k, idx_out, count_out, nullptr, nullptr, g, static_cast<cudaStream_t>(stream));
}
