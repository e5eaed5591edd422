// Mean Absolute Grasp Deviation on the device: the consumer of the generator output in eval.py.
// Reference: magd_for_grasp_distribution (eval.py:101-137) over geodesic_distance_rotmats_pairwise_np (eval.py:48-61) and the
// two SQUARED-distance helpers (eval.py:64-99).  The reference materialises three [N,M] matrices in python loops; here one
// thread owns one predicted grasp, scans the ground-truth translations staged in shared memory (direct-difference squared
// distance, first index on ties like np.argmin) and evaluates the rotation angle / joint distance only for its match.
#include "common.cuh"

namespace dexgg {

constexpr int MAGD_THREADS = 128, MAGD_TILE = 1024;

__global__ void __launch_bounds__(MAGD_THREADS) magd_match_kernel(const float* __restrict__ pt, const float* __restrict__ pr,
                                                                   const float* __restrict__ pj, const float* __restrict__ gt,
                                                                   const float* __restrict__ gr, const float* __restrict__ gj, int64_t N,
                                                                   int M, int J, int32_t* __restrict__ match, float* __restrict__ per,
                                                                   int32_t* __restrict__ flags, double* __restrict__ sums) {
    __shared__ float s_t[MAGD_TILE * 3];
    __shared__ double s_red[3][MAGD_THREADS / 32];
    const int64_t i = (int64_t)blockIdx.x * MAGD_THREADS + threadIdx.x;
    const bool live = i < N;
    const int64_t ii = live ? i : N - 1;
    const float x = pt[ii * 3], y = pt[ii * 3 + 1], z = pt[ii * 3 + 2];
    float best = INFINITY;
    int bj = 0;
    for (int j0 = 0; j0 < M; j0 += MAGD_TILE) {
        const int n = min(MAGD_TILE, M - j0);
        __syncthreads();
        for (int e = threadIdx.x; e < n * 3; e += MAGD_THREADS) s_t[e] = gt[(int64_t)j0 * 3 + e];
        __syncthreads();
        for (int j = 0; j < n; ++j) {
            const float dx = s_t[3 * j] - x, dy = s_t[3 * j + 1] - y, dz = s_t[3 * j + 2] - z;
            const float d = dx * dx + dy * dy + dz * dz;
            if (d < best) {  // strict: the first index wins a tie (np.argmin)
                best = d;
                bj = j0 + j;
            }
        }
    }
    double vt = 0.0, vr = 0.0, vj = 0.0;
    if (live) {
        float tr = 0.f;
#pragma unroll
        for (int k = 0; k < 9; ++k) tr = fmaf(pr[ii * 9 + k], gr[(int64_t)bj * 9 + k], tr);  // trace(R1 R2^T)
        const float ang = acosf(fminf(fmaxf((tr - 1.f) * 0.5f, -1.f), 1.f));
        float dj = 0.f;
        for (int k = 0; k < J; ++k) {
            const float d = gj[(int64_t)bj * J + k] - pj[ii * J + k];
            dj = fmaf(d, d, dj);
        }
        if (match) match[i] = bj;
        if (per) {
            per[i * 3] = best;
            per[i * 3 + 1] = ang;
            per[i * 3 + 2] = dj;
        }
        flags[bj] = 1;
        vt = best; vr = ang; vj = dj;
    }
    // block sums in f64 (np.sum of the reference runs in f64 whenever the ground truth is f64)
    for (int o = 16; o; o >>= 1) {
        vt += __shfl_xor_sync(0xffffffffu, vt, o);
        vr += __shfl_xor_sync(0xffffffffu, vr, o);
        vj += __shfl_xor_sync(0xffffffffu, vj, o);
    }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { s_red[0][w] = vt; s_red[1][w] = vr; s_red[2][w] = vj; }
    __syncthreads();
    if (threadIdx.x < 3) {
        double s = 0.0;
        for (int k = 0; k < MAGD_THREADS / 32; ++k) s += s_red[threadIdx.x][k];
        atomicAdd(&sums[threadIdx.x], s);
    }
}

__global__ void magd_coverage_kernel(const int32_t* __restrict__ flags, int M, double* __restrict__ sums) {
    __shared__ int s_cnt;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    int c = 0;
    for (int j = threadIdx.x; j < M; j += blockDim.x) c += flags[j] != 0;
    atomicAdd(&s_cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) sums[3] = (double)s_cnt / (double)M;
}

}  // namespace dexgg

using namespace dexgg;

extern "C" size_t dexgg_magd_workspace_bytes(int M) { return align_up((size_t)(M > 0 ? M : 1) * sizeof(int32_t), 256); }

extern "C" int dexgg_magd(const float* pred_transl, const float* pred_rot, const float* pred_joint, int64_t N, const float* gt_transl,
                          const float* gt_rot, const float* gt_joint, int M, int n_joints, int32_t* match_out, float* per_grasp_out,
                          double* sums_out, void* workspace, size_t workspace_bytes, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(N >= 1 && M >= 1 && n_joints >= 0 && pred_transl && pred_rot && gt_transl && gt_rot && sums_out &&
                      (n_joints == 0 || (pred_joint && gt_joint)),
                  "magd: bad arguments (N=%lld M=%d J=%d)", (long long)N, M, n_joints);
    DEXGG_REQUIRE(workspace && workspace_bytes >= dexgg_magd_workspace_bytes(M), "magd: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int32_t* flags = static_cast<int32_t*>(workspace);
    DEXGG_CUDA_OK(cudaMemsetAsync(flags, 0, (size_t)M * sizeof(int32_t), st));
    DEXGG_CUDA_OK(cudaMemsetAsync(sums_out, 0, 4 * sizeof(double), st));
    magd_match_kernel<<<(unsigned)((N + MAGD_THREADS - 1) / MAGD_THREADS), MAGD_THREADS, 0, st>>>(
        pred_transl, pred_rot, pred_joint, gt_transl, gt_rot, gt_joint, N, M, n_joints, match_out, per_grasp_out, flags, sums_out);
    count_launches(1);
    DEXGG_LAUNCH_OK("magd_match_kernel");
    magd_coverage_kernel<<<1, 256, 0, st>>>(flags, M, sums_out);
    count_launches(1);
    DEXGG_LAUNCH_OK("magd_coverage_kernel");
    return DEXGG_OK;
}
