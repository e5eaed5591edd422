/*
 * dexgg.h — C ABI of libdexgg.so, the B200 (sm_100a) grasp-synthesis hot path of DexGANGrasp:
 *   point cloud -> BPS encode -> DexGenerator -> rot6d->SO(3) (+palm offset) -> DexEvaluator -> threshold / top-k filter,
 *   plus the building blocks of the DexGenerator / Discriminator / DexEvaluator training step.
 *
 * The reference has no FFI layer: its boundary is the Python class surface.  Each entry point below
 * names the reference code it replaces (paths relative to the reference repo).  INTEGRATION.md shows the ctypes
 * binding a maintainer of the reference would add; dex_gan_grasp_b200/_native.py is that binding in this repo.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no torch types.  Unless a function name ends in `_host`, every data
 *     pointer is a DEVICE pointer owned by the caller; every call is stream-ordered on `stream` (a cudaStream_t
 *     passed as void*; NULL = the legacy default stream) and does not synchronise.
 *   - return value: 0 = ok, <0 = DEXGG_E_*.  dexgg_last_error() returns a thread-local message for the last failure.
 *   - no allocation inside hot calls: the caller passes a workspace sized by the matching *_workspace_bytes().
 *   - there is NO CPU fallback and no other backend: on a device that is not compute capability 10.x every compute
 *     call fails with DEXGG_E_ARCH.
 */
#ifndef DEXGG_H
#define DEXGG_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DEXGG_VERSION 100

enum {
    DEXGG_OK = 0,
    DEXGG_E_ARG = -1,       /* bad argument (null pointer, negative size, unsupported shape) */
    DEXGG_E_ARCH = -2,      /* device is not sm_100 (B200) */
    DEXGG_E_WORKSPACE = -3, /* workspace missing or too small */
    DEXGG_E_CUDA = -4       /* a CUDA runtime / driver call failed (message has the CUDA error string) */
};

int dexgg_version(void);
const char* dexgg_last_error(void);
/* 0 when the CURRENT device can run this library (compute capability 10.x), else DEXGG_E_ARCH / DEXGG_E_CUDA. */
int dexgg_device_check(void);

/* ------------------------------------------------------------------------------------------------------------
 * Stage 1 — Basis Point Set encoding.
 * Replaces bps_torch.encode(x)['dists'] (+ the chamfer_distance NmDistanceKernel underneath it) as called from
 * DexGanGrasp/data/bps_encoder.py:28-55, DexGanGrasp/utils/utils.py:132-138, dexgangrasp_real_robot.py:139.
 *   basis    [P,3] f32 (models/basis_point_set.npy, P = 4096 in the reference; any P >= 1 here)
 *   pts      [total_points,3] f32 AoS, all clouds back to back; must be 16-byte aligned
 *   offsets  [B+1] i64 DEVICE array of point offsets (cloud b = pts[offsets[b] : offsets[b+1]]), or NULL for B
 *            uniform clouds of max_cloud_points points each
 *   dists    [B,P] f32 out: sqrt((dx^2 + dz^2) + dy^2) of the nearest point (the summation order that reproduces
 *            the reference's stored encodings data/real_objects/bps/<name>.npy bit-exactly)
 *   idx      [B,P] i32 out (may be NULL): index of the nearest point inside its cloud; fp32 direct-difference
 *            squared distance fma(dz,dz,fma(dy,dy,dx*dx)), lowest index wins an exact tie.
 *            Empty cloud -> dists = +inf, idx = -1.
 *   flags    DEXGG_BPS_* below.
 * ---------------------------------------------------------------------------------------------------------- */
#define DEXGG_BPS_DEFAULT 0
#define DEXGG_BPS_SCALAR 1     /* simple scalar compare-select kernel (cross-check of the packed kernel) */
#define DEXGG_BPS_NSPLIT(n) (((n) & 0xff) << 8) /* force the number of point-range splits per cloud (0 = auto) */

size_t dexgg_bps_workspace_bytes(int B, int P);
int dexgg_bps_encode(const float* basis, int P, const float* pts, const int64_t* offsets, int B,
                     int64_t max_cloud_points, int64_t total_points, float* dists, int32_t* idx, int flags,
                     void* workspace, size_t workspace_bytes, void* stream);
/* Same call with HOST buffers (copies in, encodes, copies back, synchronises): the shape of
 * BPSEncoder.encode_pcd_with_bps (bps_encoder.py:28-55), which hands back a numpy array. offsets is a HOST array. */
int dexgg_bps_encode_host(const float* basis, int P, const float* pts, const int64_t* offsets, int B, float* dists,
                          int32_t* idx, int flags);

/* FP32-pipe micro-benchmark used for the BPS roofline denominator: runs `iters` rounds of 8
 * independent FFMA (kind 0), FFMA2 (kind 1), FADD2 (kind 2) or FMNMX3 (kind 3) chains per thread on a full grid and
 * returns the elapsed milliseconds in *ms and the number of lane-operations executed in *lane_ops. */
int dexgg_fp32_pipe_probe(int kind, int iters, float* ms, double* lane_ops);

#ifdef __cplusplus
}
#endif
#endif /* DEXGG_H */
