// 3-vector helpers shared by the rot6d kernels (pose_ops.cu) and the fused generator-loss kernel (train_fused.cu).
#pragma once
namespace dexgg {

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 cross(V3 p, V3 q) { return {p.y * q.z - p.z * q.y, p.z * q.x - p.x * q.z, p.x * q.y - p.y * q.x}; }
__device__ __forceinline__ float dot(V3 p, V3 q) { return p.x * q.x + p.y * q.y + p.z * q.z; }
__device__ __forceinline__ V3 add(V3 p, V3 q) { return {p.x + q.x, p.y + q.y, p.z + q.z}; }
__device__ __forceinline__ V3 scale(V3 p, float s) { return {p.x * s, p.y * s, p.z * s}; }

// rot6d -> (x, y, z) columns (utils.rot_matrix_from_ortho6d, DexGanGrasp/utils/utils.py:380-394)
__device__ __forceinline__ void rot6d_columns(V3 a, V3 b, V3& x, V3& y, V3& z) {
    x = scale(a, 1.f / fmaxf(sqrtf(dot(a, a)), 1e-8f));
    const V3 u = cross(x, b);
    z = scale(u, 1.f / fmaxf(sqrtf(dot(u, u)), 1e-8f));
    y = cross(z, x);
}
// gradient of the above: (gx, gy, gz) wrt the columns -> (ga, gb)
__device__ __forceinline__ void rot6d_columns_bwd(V3 a, V3 b, V3 gx, V3 gy, V3 gz, V3& ga, V3& gb) {
    const float la = sqrtf(dot(a, a)), na = fmaxf(la, 1e-8f);
    const V3 x = scale(a, 1.f / na);
    const V3 u = cross(x, b);
    const float lu = sqrtf(dot(u, u)), nu = fmaxf(lu, 1e-8f);
    const V3 z = scale(u, 1.f / nu);
    gz = add(gz, cross(x, gy));
    gx = add(gx, cross(gy, z));
    V3 gu = scale(gz, 1.f / nu);
    if (lu > 1e-8f) gu = add(gu, scale(z, -dot(z, gz) / nu));
    gx = add(gx, cross(b, gu));
    gb = cross(gu, x);
    ga = scale(gx, 1.f / na);
    if (la > 1e-8f) ga = add(ga, scale(x, -dot(x, gx) / na));
}

}  // namespace dexgg
