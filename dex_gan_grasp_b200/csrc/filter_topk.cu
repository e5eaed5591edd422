// Stage 4 — score threshold + sort / top-k per object.
// Replaces `p.sort(descending=True)`, `indices[sorted_score > thresh]` of DexGanGrasp.filter_grasps
// (DexGanGrasp/models/dexgangrasp.py:412-419) — per segment (object) instead of one object per Python call.
// Order: score descending, equal scores by the lower row index (torch leaves that order unspecified).
// One CTA per segment: 64-bit keys (~orderable(score) << 32 | local index) are bitonic-sorted in shared memory.
// HBM-bound in principle (4 B read per grasp); at the sizes of this path it is launch/latency-bound and negligible.
#include "common.cuh"

namespace dexgg {

constexpr int FILTER_MAX_SEG = 16384;  // rows per segment that fit the shared-memory sort (128 KB of keys)

__device__ __forceinline__ uint32_t orderable(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__global__ void __launch_bounds__(256)
filter_sort_kernel(const float* __restrict__ p, int64_t rows_per_seg, int64_t M, float thresh, int k, int npow2,
                   int32_t* __restrict__ idx_out, int32_t* __restrict__ count_out, float* __restrict__ best_out,
                   float* __restrict__ score_out) {
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
    if (threadIdx.x == 0) {
        count_out[seg] = kept;
        if (best_out) {
            const uint32_t o = s_best;
            const uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
            best_out[seg] = n > 0 ? __uint_as_float(u) : -__int_as_float(0x7f800000);
        }
    }
}

}  // namespace dexgg

using namespace dexgg;

extern "C" int dexgg_filter_topk(const float* p, int64_t M, int64_t rows_per_seg, float thresh, int k, int32_t* idx_out,
                                 int32_t* count_out, float* best_out, float* score_out, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && rows_per_seg >= 1 && k >= 1, "filter_topk: bad sizes");
    if (M == 0) return DEXGG_OK;
    DEXGG_REQUIRE(p && idx_out && count_out, "filter_topk: null pointer");
    DEXGG_REQUIRE(rows_per_seg <= FILTER_MAX_SEG || M <= FILTER_MAX_SEG, "filter_topk: at most %d rows per segment", FILTER_MAX_SEG);
    DEXGG_REQUIRE(k <= rows_per_seg || k <= M, "filter_topk: k larger than a segment");
    const int64_t seg_rows = rows_per_seg < M ? rows_per_seg : M;
    const int64_t nseg = (M + rows_per_seg - 1) / rows_per_seg;
    int npow2 = 2;
    while (npow2 < seg_rows) npow2 <<= 1;
    const size_t smem = (size_t)npow2 * sizeof(unsigned long long);
    static bool attr = false;
    if (!attr) {
        DEXGG_CUDA_OK(cudaFuncSetAttribute(filter_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FILTER_MAX_SEG * 8));
        attr = true;
    }
    filter_sort_kernel<<<(unsigned)nseg, 256, smem, static_cast<cudaStream_t>(stream)>>>(p, rows_per_seg, M, thresh, k, npow2, idx_out,
                                                                                         count_out, best_out, score_out); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("filter_sort_kernel");
    return DEXGG_OK;
}
