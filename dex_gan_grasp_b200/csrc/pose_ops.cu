// rot6d -> SO(3) (Gram-Schmidt) forward and backward.
// Reference: utils.rot_matrix_from_ortho6d / normalize_vector / cross_product (DexGanGrasp/utils/utils.py:380-394,323-332,68-78).
//   x = a / max(|a|,1e-8);  z = (x × b) / max(|x × b|,1e-8);  y = z × x;  R = [x y z] as columns (row-major [M,3,3]).
#include "common.cuh"
#include "pose.cuh"

namespace dexgg {

__global__ void rot6d_forward_kernel(const float* __restrict__ r6, int64_t M, float* __restrict__ rot) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    const V3 a{r6[m * 6], r6[m * 6 + 1], r6[m * 6 + 2]}, b{r6[m * 6 + 3], r6[m * 6 + 4], r6[m * 6 + 5]};
    const V3 x = scale(a, 1.f / fmaxf(sqrtf(dot(a, a)), 1e-8f));
    const V3 u = cross(x, b);
    const V3 z = scale(u, 1.f / fmaxf(sqrtf(dot(u, u)), 1e-8f));
    const V3 y = cross(z, x);
    float* r = rot + m * 9;
    r[0] = x.x; r[1] = y.x; r[2] = z.x;
    r[3] = x.y; r[4] = y.y; r[5] = z.y;
    r[6] = x.z; r[7] = y.z; r[8] = z.z;
}

// g_rot [M,9] -> g_r6 [M,6]
__global__ void rot6d_backward_kernel(const float* __restrict__ r6, const float* __restrict__ g_rot, int64_t M,
                                      float* __restrict__ g_r6) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    const V3 a{r6[m * 6], r6[m * 6 + 1], r6[m * 6 + 2]}, b{r6[m * 6 + 3], r6[m * 6 + 4], r6[m * 6 + 5]};
    const float la = sqrtf(dot(a, a)), na = fmaxf(la, 1e-8f);
    const V3 x = scale(a, 1.f / na);
    const V3 u = cross(x, b);
    const float lu = sqrtf(dot(u, u)), nu = fmaxf(lu, 1e-8f);
    const V3 z = scale(u, 1.f / nu);
    const float* g = g_rot + m * 9;
    V3 gx{g[0], g[3], g[6]}, gy{g[1], g[4], g[7]}, gz{g[2], g[5], g[8]};
    // y = z × x
    gz = add(gz, cross(x, gy));
    gx = add(gx, cross(gy, z));
    // z = u / max(|u|, eps)
    V3 gu = scale(gz, 1.f / nu);
    if (lu > 1e-8f) gu = add(gu, scale(z, -dot(z, gz) / nu));
    // u = x × b
    gx = add(gx, cross(b, gu));
    const V3 gb = cross(gu, x);
    // x = a / max(|a|, eps)
    V3 ga = scale(gx, 1.f / na);
    if (la > 1e-8f) ga = add(ga, scale(x, -dot(x, gx) / na));
    float* o = g_r6 + m * 6;
    o[0] = ga.x; o[1] = ga.y; o[2] = ga.z; o[3] = gb.x; o[4] = gb.y; o[5] = gb.z;
}

}  // namespace dexgg

using namespace dexgg;

extern "C" int dexgg_rot6d_forward(const float* r6, int64_t M, float* rot, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && (M == 0 || (r6 && rot)), "rot6d_forward: bad arguments");
    if (M == 0) return DEXGG_OK;
    rot6d_forward_kernel<<<(unsigned)((M + 127) / 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(r6, M, rot); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("rot6d_forward_kernel");
    return DEXGG_OK;
}

extern "C" int dexgg_rot6d_backward(const float* r6, const float* g_rot, int64_t M, float* g_r6, void* stream) {
    DEXGG_ARCH_OK();
    DEXGG_REQUIRE(M >= 0 && (M == 0 || (r6 && g_rot && g_r6)), "rot6d_backward: bad arguments");
    if (M == 0) return DEXGG_OK;
    rot6d_backward_kernel<<<(unsigned)((M + 127) / 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(r6, g_rot, M, g_r6); ::dexgg::count_launches(1);
    DEXGG_LAUNCH_OK("rot6d_backward_kernel");
    return DEXGG_OK;
}
