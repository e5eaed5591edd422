// Part-conditioned grasp ranking: centroid of the cropped object-part cloud, squared distance of every
// grasp translation to it, ascending order.  Reference: DexGanGrasp/utils/filter_grasps_given_mask.py:12-30 (pairwise squared
// distance), :84-138 (`part_pcd_np -= pc_center`, three |coord| < 0.2 crops, `mean(axis=0)`, `np.min(dist, axis=1)`,
// `np.argsort`); caller dexafford_prompt.py:191-198 keeps the first 30 and sorts them with the evaluator (thresh = -1).
// The ordering itself is dexgg_filter_topk on the score -d2 (best first = nearest first, equal keys by lower row).
#include "common.cuh"

namespace dexgg {

__global__ void part_centroid_kernel(const float* __restrict__ pts, int64_t n, float cx, float cy, float cz, float crop,
                                     double* __restrict__ acc /* [4]: sum x, y, z, count */) {
    double sx = 0.0, sy = 0.0, sz = 0.0, c = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        // the reference subtracts in the array's own precision (float64 from Open3D); here f32 inputs, f64 arithmetic
        const double x = (double)pts[3 * i] - (double)cx, y = (double)pts[3 * i + 1] - (double)cy, z = (double)pts[3 * i + 2] - (double)cz;
        if (fabs(z) < (double)crop && fabs(y) < (double)crop && fabs(x) < (double)crop) {
            sx += x; sy += y; sz += z; c += 1.0;
        }
    }
    for (int o = 16; o; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o);
        sy += __shfl_xor_sync(0xffffffffu, sy, o);
        sz += __shfl_xor_sync(0xffffffffu, sz, o);
        c += __shfl_xor_sync(0xffffffffu, c, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&acc[0], sx);
        atomicAdd(&acc[1], sy);
        atomicAdd(&acc[2], sz);
        atomicAdd(&acc[3], c);
    }
}

__global__ void part_score_kernel(const float* __restrict__ transl, int64_t n, const double* __restrict__ acc,
                                  float* __restrict__ score, float* __restrict__ centroid_out) {
    const double c = acc[3];
    const double mx = acc[0] / c, my = acc[1] / c, mz = acc[2] / c;  // c == 0 -> NaN like numpy's mean of an empty array
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0 && centroid_out) {
        centroid_out[0] = (float)mx; centroid_out[1] = (float)my; centroid_out[2] = (float)mz;
    }
    if (i >= n) return;
    const double dx = fabs(mx - (double)transl[3 * i]), dy = fabs(my - (double)transl[3 * i + 1]), dz = fabs(mz - (double)transl[3 * i + 2]);
    score[i] = -(float)(dx * dx + dy * dy + dz * dz);
}

}  // namespace dexgg

using namespace dexgg;

extern "C" int dexgg_filter_topk(const float* p, int64_t M, int64_t rows_per_seg, float thresh, int k, int32_t* idx_out,
                                 int32_t* count_out, float* best_out, float* score_out, void* stream);

extern "C" size_t dexgg_part_rank_workspace_bytes(int64_t n_grasps) {
    return align_up(4 * sizeof(double), 256) + align_up((size_t)(n_grasps > 0 ? n_grasps : 1) * sizeof(float), 256) + 256;
}

extern "C" int dexgg_part_rank(const float* transl, int64_t n_grasps, const float* part_pts, int64_t n_part, const float* pc_center,
                               float crop, int k, int32_t* idx_out, float* sqdist_out, float* centroid_out, void* workspace,
                               size_t workspace_bytes, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(transl && part_pts && pc_center && idx_out && n_grasps >= 1 && n_grasps <= 16384 && n_part >= 1 && k >= 1 && k <= n_grasps,
                  "part_rank: bad arguments (n_grasps=%lld n_part=%lld k=%d; at most 16384 grasps per call)", (long long)n_grasps,
                  (long long)n_part, k);
    DEXGG_REQUIRE(workspace && workspace_bytes >= dexgg_part_rank_workspace_bytes(n_grasps), "part_rank: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned char* ws = static_cast<unsigned char*>(workspace);
    double* acc = reinterpret_cast<double*>(ws);
    float* score = reinterpret_cast<float*>(ws + align_up(4 * sizeof(double), 256));
    int32_t* cnt = reinterpret_cast<int32_t*>(ws + align_up(4 * sizeof(double), 256) + align_up((size_t)n_grasps * sizeof(float), 256));
    DEXGG_CUDA_OK(cudaMemsetAsync(acc, 0, 4 * sizeof(double), st));
    const int blocks = (int)((n_part + 255) / 256 < 4 * sm_count() ? (n_part + 255) / 256 : 4 * sm_count());
    part_centroid_kernel<<<blocks, 256, 0, st>>>(part_pts, n_part, pc_center[0], pc_center[1], pc_center[2], crop, acc);
    count_launches(1);
    DEXGG_LAUNCH_OK("part_centroid_kernel");
    part_score_kernel<<<(unsigned)((n_grasps + 255) / 256), 256, 0, st>>>(transl, n_grasps, acc, score, centroid_out);
    count_launches(1);
    DEXGG_LAUNCH_OK("part_score_kernel");
    int rc = dexgg_filter_topk(score, n_grasps, n_grasps, -INFINITY, k, idx_out, cnt, nullptr, sqdist_out, stream);
    if (rc) return rc;
    return DEXGG_OK;  // sqdist_out holds the NEGATED squared distances of the kept grasps, nearest first (see dexgg.h)
}
