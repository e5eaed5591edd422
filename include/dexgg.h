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
/* Number of CUDA kernels this library has launched since it was loaded (bench.py reports the per-step difference). */
unsigned long long dexgg_launch_count(void);

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
#define DEXGG_BPS_SCALAR 1     /* simple scalar compare-select kernel (cross-check of the packed kernels) */
#define DEXGG_BPS_DIRECT 2     /* packed kernel that scans with the direct-difference form itself (6 FP32 ops / pair) instead of
                                  the default expanded-form filter (3 FFMA / pair) + exact evaluation of the candidates */
#define DEXGG_BPS_FMA 4        /* the filter scan on the FP32 pipe (bps_nn_filter_kernel, 3 FFMA / pair) instead of the default
                                  tensor-core scan (bps_umma.cu); same results, kept as its cross-check */
#define DEXGG_BPS_NSPLIT(n) (((n) & 0xff) << 8) /* force the number of point-range splits per cloud (0 = auto) */

size_t dexgg_bps_workspace_bytes(int B, int P);
int dexgg_bps_encode(const float* basis, int P, const float* pts, const int64_t* offsets, int B,
                     int64_t max_cloud_points, int64_t total_points, float* dists, int32_t* idx, int flags,
                     void* workspace, size_t workspace_bytes, void* stream);
/* dexgg_bps_encode WITHOUT a caller workspace and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The scratch (arg-min keys, packed basis operand) is a static device array inside the library, held by every GPU that
 * loads it until exit (2.5 MiB).  Limit: B * P <= dexgg_bps_encode_static_max_keys() = 262144 (e.g. 64 clouds x 4096 basis points);
 * beyond it the call fails with DEXGG_E_ARG and a message, and dexgg_bps_encode with a caller workspace is the path.  Same arguments,
 * results (bit-identical) and flags otherwise.  The call itself calls no cudaMalloc / cudaFree.  All calls on a device SHARE the
 * array: each call re-initialises what it needs in stream order, so calls ordered on one stream (or by events) are safe; calls that
 * overlap on the GPU race on it. */
size_t dexgg_bps_encode_static_max_keys(void);
int dexgg_bps_encode_static(const float* basis, int P, const float* pts, const int64_t* offsets, int B,
                            int64_t max_cloud_points, int64_t total_points, float* dists, int32_t* idx, int flags, void* stream);
/* The step in front of the encoder, on the device: obj_pcd.translate(-obj_pcd.get_center()) as the callers do it with Open3D
 * (dexgangrasp_real_robot.py:133-134, DexGanGrasp/data/bps_encoder.py:69).  Open3D keeps points as doubles, so per cloud:
 * centers[b] = float64 mean of the points, pts_out = float32(double(x) - centre) (pts_out may alias pts).  Same cloud layout as
 * dexgg_bps_encode (offsets NULL = B uniform clouds).  The caller keeps `centers` to move the grasps back (transl + pc_center, :167). */
int dexgg_center_clouds(const float* pts, const int64_t* offsets, int B, int64_t max_cloud_points, int64_t total_points,
                        float* pts_out, double* centers, void* stream);
/* Same call with HOST buffers (copies in, encodes, copies back, synchronises): the shape of
 * BPSEncoder.encode_pcd_with_bps (bps_encoder.py:28-55), which hands back a numpy array. offsets is a HOST array. */
int dexgg_bps_encode_host(const float* basis, int P, const float* pts, const int64_t* offsets, int B, float* dists,
                          int32_t* idx, int flags);

/* ------------------------------------------------------------------------------------------------------------
 * Stages 2-3 — DexGenerator / Discriminator / DexEvaluator forward (eval-mode BatchNorm).
 * Replace Generator.forward (DexGanGrasp/models/networks.py:88-103) + utils.rot_matrix_from_ortho6d
 * (utils/utils.py:380-394) + utils.translate_along_axis (utils/utils.py:19-48) as driven by
 * DexGANGrasp.generate_poses (networks.py:223-255); DexEvaluator.forward (networks.py:315-337) as driven by
 * DexGanGrasp.evaluate_grasps / filter_grasps (models/dexgangrasp.py:367-442); Discriminator.forward
 * (networks.py:148-159) as driven by filter_grasps_discriminator (dexgangrasp.py:444-493).
 *
 * A net handle owns packed copies of the weights (input BatchNorm folded into the BPS columns, fc1 and fc3 of every
 * ResBlock fused into one 768-row matrix, bn1/bn2 folded into per-column affines).  `params` lists the module's
 * float tensors in state_dict order WITHOUT num_batches_tracked:
 *   generator (38):     gen_bn1.{weight,bias,running_mean,running_var}, gen_rb1.*, gen_rb2.*, gen_joint_conf.{weight,bias},
 *                       gen_rot.{weight,bias}, gen_transl.{weight,bias}
 *   discriminator (34): disc_bn1.*, disc_rb1.*, disc_rb2.*, out_success.{weight,bias}
 *   evaluator (48):     bn1.*, rb1.*, rb2.*, rb3.*, out_success.{weight,bias}
 *   with ResBlock = fc1.{weight,bias}, bn1.{weight,bias,running_mean,running_var}, fc2.{weight,bias}, bn2.{...},
 *   fc3.{weight,bias} (14 tensors).  Call dexgg_net_update after the parameters change (optimizer step,
 *   load_state_dict); the handle keeps no reference to the caller's tensors.
 *
 * Rows and objects: the BPS encoding is given once per OBJECT (bps [n_obj,4096]); row m belongs to object
 * m / rows_per_obj (the reference tiles the encoding instead, np.tile at dexgangrasp.py:403-404,509-510).  The 4096-wide part
 * of every first-layer product is computed once per object.  rows_per_obj = 1 is the row-wise case (every row its
 * own encoding: eval_batch, training-style batches).  M must equal n_obj * rows_per_obj.
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct dexgg_net dexgg_net;
enum { DEXGG_NET_GENERATOR = 0, DEXGG_NET_DISCRIMINATOR = 1, DEXGG_NET_EVALUATOR = 2 };
enum {
    DEXGG_PRECISION_FP32 = 0, /* CUDA-core fp32: within 1e-4 abs of the reference's torch fp32 path */
    DEXGG_PRECISION_BF16 = 1, /* tcgen05 bf16 operands, fp32 accumulation in TMEM */
    DEXGG_PRECISION_FP32_TC = 2 /* the 1e-4 contract ON the tensor cores: every activation and weight carried as a bf16 hi + lo
                                   pair, three products per MAC (hi*hi + lo*hi + hi*lo) in one tcgen05 GEMM of 3x the K,
                                   fp32 accumulation, fp32 skip branch */
};

int dexgg_net_create(int kind, const float* const* params, int n_params, int latent_dim, int n_joints, int precision,
                     void* stream, dexgg_net** out);
int dexgg_net_update(dexgg_net* net, const float* const* params, int n_params, void* stream);
void dexgg_net_destroy(dexgg_net* net);
size_t dexgg_net_workspace_bytes(const dexgg_net* net, int64_t M, int64_t n_obj);

/* z [M,latent_dim] (torch.randn at networks.py:240 — the caller draws it, "same z" contract) ->
 * rot [M,9] row-major 3x3 with columns (x,y,z), transl [M,3] already moved by -z_offset along the palm x axis,
 * joint [M,n_joints]; rot6d [M,6] optional raw head output (NULL to skip). */
int dexgg_generator_forward(const dexgg_net* net, const float* bps, int64_t n_obj, int64_t rows_per_obj, const float* z,
                            int64_t M, float z_offset, float* rot, float* transl, float* joint, float* rot6d,
                            void* workspace, size_t workspace_bytes, void* stream);
/* p [M] = sigmoid(out_success(...)); joint_conf is not an evaluator input (networks.py:307-313). */
int dexgg_evaluator_forward(const dexgg_net* net, const float* bps, int64_t n_obj, int64_t rows_per_obj, const float* rot,
                            const float* transl, int64_t M, float* p, void* workspace, size_t workspace_bytes,
                            void* stream);
int dexgg_discriminator_forward(const dexgg_net* net, const float* bps, int64_t n_obj, int64_t rows_per_obj,
                                const float* rot, const float* transl, const float* joint, int64_t M, float* p,
                                void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Stage 4 — threshold + sort / top-k, per segment of rows_per_seg consecutive rows (one object each).
 * Replaces p.sort(descending=True) / indices[sorted_score > thresh] of DexGanGrasp.filter_grasps
 * (models/dexgangrasp.py:412-419): keeps p > thresh (strict), best first, equal scores by lower row index.
 *   idx_out   [n_seg,k] i32 global row ids, -1 padded;  count_out [n_seg] i32 = min(#kept, k)
 *   best_out  [n_seg] f32 (may be NULL) best score of the segment — the caller raises the reference's ValueError
 *             when best < thresh (dexgangrasp.py:414-416);  score_out [n_seg,k] f32 (may be NULL).
 * k = rows_per_seg gives the reference's full sorted order.
 * dexgg_filter_topk: segments of at most 16384 rows (one CTA sorts a segment in shared memory).  dexgg_filter_topk_ws: any
 * segment length (eval.py samples 5 x n_gt grasps per object, :317-322) — longer segments are sorted in the caller's
 * workspace (dexgg_filter_topk_workspace_bytes, 0 for short segments).  Row ids are int32: M < 2^31.
 * dexgg_filter_topk_gather: the batched pipeline's whole stage in ONE launch (segments <= 16384 rows): top-k as above plus the
 * gather of the kept rows (the per-key torch.gather of dexgangrasp.py:421-435, without the tiled [k,4096] BPS block) into
 *   packed_out [n_seg,k,14+n_joints] f32: score | global row id (int32 BITS) | rot 9 | transl 3 | joint_conf n_joints;
 *   records beyond count hold id -1 and zeros.
 * ---------------------------------------------------------------------------------------------------------- */
int dexgg_filter_topk(const float* p, int64_t M, int64_t rows_per_seg, float thresh, int k, int32_t* idx_out,
                      int32_t* count_out, float* best_out, float* score_out, void* stream);
size_t dexgg_filter_topk_workspace_bytes(int64_t M, int64_t rows_per_seg);
int dexgg_filter_topk_ws(const float* p, int64_t M, int64_t rows_per_seg, float thresh, int k, int32_t* idx_out,
                         int32_t* count_out, float* best_out, float* score_out, void* workspace, size_t workspace_bytes,
                         void* stream);
int dexgg_filter_topk_gather(const float* p, int64_t M, int64_t rows_per_seg, float thresh, int k, const float* rot,
                             const float* transl, const float* joint, int n_joints, int32_t* idx_out, int32_t* count_out,
                             float* packed_out, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * One fused linear layer, fp32 CUDA cores:  C = act(scale * (A·Wᵀ + bias) + shift + residual)  [+ C if accumulate].
 * Building block of the training step (nn.Linear forward, dgrad with trans_w, wgrad with trans_a) and of the nets above.
 * A(m,k) = A[m*lda+k] (trans_a: A[k*lda+m]);  W(n,k) = W[n*ldw+k] (trans_w: W[k*ldw+n]);  act 1 = LeakyReLU(0.2).
 * ---------------------------------------------------------------------------------------------------------- */
int dexgg_linear_f32(const float* A, int64_t lda, int trans_a, const float* W, int64_t ldw, int trans_w, int M, int N,
                     int K, const float* bias, const float* scale, const float* shift, const float* residual,
                     int64_t ldr, int act, int accumulate, float* C, int64_t ldc, void* stream);

/* One fused linear layer on the 5th-generation tensor cores (tcgen05, bf16 operands, fp32 accumulation in TMEM):
 * C = act(scale * (A·Wᵀ + bias) + shift + residual).  A [M,K] and W [N,K] are bf16, row-major, row pitch a multiple of
 * 16 bytes; K any length.  Output bf16 (C_bf16) and / or fp32 (C_f32), row pitch ldc elements.
 * dexgg_f32_to_bf16 converts a strided fp32 matrix to bf16 (round to nearest even); dexgg_f32_to_bf16_t writes the
 * TRANSPOSE (dst[c*ldd + r] = src[r*lds + c], ldd >= rows) - the K-major operand of a product that contracts over the rows
 * of src, as the training step's wgrad (dW = dY^T X) and dgrad (dX = dY W) need (models/dexgangrasp.py:739-823). */
int dexgg_linear_bf16(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, const float* bias,
                      const float* scale, const float* shift, const void* residual_bf16, int64_t ldr, int act, void* C_bf16,
                      float* C_f32, int64_t ldc, void* stream);
/* ------------------------------------------------------------------------------------------------------------
 * General bf16 tensor-core GEMM of the training step (csrc/gemm_umma.cu):
 *     C[M,N] = act(sum_s op(A_s)[M,K_s] . op(B_s)[N,K_s]^T + bias)
 * Up to three K segments accumulate into one TMEM tile (the reference's torch.cat of ResBlock inputs, networks.py:97,153,325,
 * is never materialised).  Operand layouts, bf16, row pitch (ld, elements) a multiple of 8:
 *     K-major   element (i, k) at p[i*ld + k]    (forward: activations [rows,K], nn.Linear weights [N_out,K_in])
 *     MN-major  element (i, k) at p[k*ld + i]    (dgrad reads W [N_out,K_in] as the [K_in x N_out] operand; wgrad reads
 *                                                 dY [rows,N_out] and X [rows,K_in] as they lie: no transposed copies)
 * Replaces the cuBLAS calls behind nn.Linear forward / backward in DexGanGrasp.update_dexgangrasp / update_dexevaluator
 * (models/dexgangrasp.py:739-823).  act_from: columns >= act_from get LeakyReLU(0.2) (>= N: none).  split_k > 1 splits the
 * contraction over CTAs and accumulates raw fp32 partial sums into C32 with atomic adds (C32 zeroed by the caller; no bias /
 * act / bf16 output): for outputs of few tiles and a long contraction (weight gradients of the small layers).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct {
    const void* A;      /* bf16 */
    int64_t lda;
    int a_mn_major;
    const void* B;      /* bf16 */
    int64_t ldb;
    int b_mn_major;
    int K;
} dexgg_gemm_seg;
typedef struct {
    int M, N, n_seg;
    dexgg_gemm_seg seg[3];
    const float* bias;  /* [N] or NULL */
    int act_from;
    void* C16;          /* bf16 [M, ldc] or NULL */
    float* C32;         /* f32  [M, ldc] or NULL */
    int64_t ldc;
    int accumulate;     /* 1: C32 += result (split_k == 1) */
    int split_k;        /* 0 / 1: no split */
    int tile_n;         /* 0: chosen by the library; else 64 / 128 / 192 / 256 */
    int64_t split_slice_stride; /* split_k > 1: 0 = atomic accumulation; > 0 = split s writes its partial sums at C32 + s * stride
                                   (plain stores, deterministic; the caller sums the slices) */
} dexgg_gemm;
int dexgg_gemm_bf16(const dexgg_gemm* g, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Fused non-GEMM kernels of the bf16 training step (csrc/train_fused.cu; driven by dex_gan_grasp_b200/train_engine.py).
 * They replace the ATen kernels autograd runs for nn.BatchNorm1d / LeakyReLU / Dropout / torch.cat / BCELoss / MSELoss /
 * rot_matrix_from_ortho6d forward AND backward in DexGanGrasp.update_dexgangrasp / update_dexevaluator
 * (models/dexgangrasp.py:739-823; networks.py:43-56, 88-103, 148-159, 315-337; losses.py:61-76).  All matrices bf16 unless noted,
 * row pitches (ld, elements) multiples of 8; statistics, losses and parameter gradients fp32.
 * ---------------------------------------------------------------------------------------------------------- */
/* Exchange descriptor of the slab kernels (mode 0): the column sums are combined across the CTAs of the launch AND across the
 * ranks of a data-parallel job INSIDE the kernel, through peer-mapped scratch buffers (cudaIpc over NVLink) - one exchange per
 * BatchNorm, no NCCL launch, bit-identical totals on every rank.  peers: DEVICE array [world] of every rank's scratch base
 * (dexgg_peer_alloc(dexgg_slab_scratch_bytes()) on each rank, handles exchanged by the host, dexgg_peer_open on the others);
 * slot: index of this launch inside the step (same on every rank, < 64); parity: device word flipped by dexgg_slab_step_begin
 * at the start of every step.  world = 1: peers = {own scratch}. */
typedef struct {
    void* const* peers;
    int world, rank, slot;
    const unsigned* parity;
} dexgg_slab_comm;
size_t dexgg_slab_scratch_bytes(void);
int dexgg_peer_alloc(size_t bytes, void** ptr, unsigned char* handle64);
int dexgg_peer_open(const unsigned char* handle64, void** ptr);
int dexgg_peer_close(void* ptr);
int dexgg_peer_free(void* ptr);
int dexgg_slab_step_begin(void* scratch_base, unsigned* parity, void* stream);
/* small cross-rank sum (sum_out, may alias src) and / or gather (gather_out [world, n]) of n floats through the same peer scratch:
 * one launch, no NCCL; takes ceil(n / 2048) slots starting at comm->slot */
int dexgg_p2p_small(const dexgg_slab_comm* comm, const float* src, int n, float* sum_out, float* gather_out, void* stream);
/* gradient exchange of a data-parallel step without NCCL: in-place MEAN (scale = 1 / world) of n floats (n % 4 == 0) held by
 * every rank in peer-mapped memory (dexgg_peer_alloc + dexgg_peer_open; bufs: DEVICE array [world], bufs[rank] = own buffer).
 * Two-shot over NVLink: rank r reduces chunk r of every rank in rank order and stores the result to every rank, so the ranks
 * end bit-identical.  One launch of `ctas` (0: 64) narrow CTAs, one exchange slot (comm->slot) for its two flag barriers.
 * Replaces dataparallel.allreduce_gradients' NCCL call for the training engine (reference: single-process training,
 * train.py:249-276; the data-parallel step is this repo's multi-GPU row e2). */
int dexgg_grad_allreduce(const dexgg_slab_comm* comm, void* const* bufs, long long n, float scale, int ctas, void* stream);

/* Train-mode BatchNorm1d forward as ONE kernel: batch statistics of x per group of rows_per_group rows (the discriminator's
 * real and fake batches are two groups of one call), y = [dropout] LeakyReLU(gamma * xhat + beta [+ residual]).  A CTA owns
 * 16 columns x a row range; the grid (at most one CTA per SM) exchanges its column sums between the statistics pass and the
 * apply pass (mode 0: in-kernel, across CTAs and ranks - dexgg_slab_comm).  NCCL fallback: mode 1 adds the local shifted sums to
 * acc and returns; after ONE allreduce of acc, mode 2 applies (count_total = rows of a group over all ranks).  Dropout: drop_mask (fp32, pre-scaled keep mask) or, if NULL and
 * drop_p > 0, a counter-based Bernoulli mask from *drop_seed and drop_stream (the backward regenerates it). */
typedef struct {
    const void* x; int64_t ldx;
    int C; int64_t rows_per_group; int n_groups;
    const float* gamma; const float* beta;
    const float* shift;                  /* [C] common shift of the one-pass variance (the layer's running mean) or NULL */
    float* mean; float* invstd;          /* [n_groups, C] saved for the backward */
    float eps; float count_total;        /* rows of a group over ALL ranks (0: rows_per_group) */
    const void* residual; int64_t ldr;
    int act;
    const float* drop_mask; int64_t ldm; float drop_p; const unsigned long long* drop_seed; unsigned drop_stream;
    void* y; int64_t ldy;
    int mode;
    float* acc;                          /* modes 1 / 2 (NCCL fallback): [n_groups, 2, C] scratch, ZEROED by the caller */
    dexgg_slab_comm comm;                /* mode 0 */
    int ctas_per_sm;                     /* 0: library default (3 forward / 2 backward); 1 when two step chains run concurrently on
                                            two streams, so that the grids of two spinning slab kernels always fit the GPU together */
} dexgg_bn_fwd;
int dexgg_bn_fwd_slab(const dexgg_bn_fwd* args, void* stream);

/* Its backward: g_v = g [* dropout] * LeakyReLU'(out);  dgamma += sum g_v xhat;  dbeta += sum g_v;
 * gx = gamma invstd (g_v - mean(g_v) - xhat mean(g_v xhat));  gskip = g_v * LeakyReLU'(skip), dbias_skip += sum gskip.
 * The incoming gradient is a matrix g or the rank-1 form dl[m] * w[c] of a score head (then dw_out += sum_m dl[m] out[m,:]). */
typedef struct {
    const void* g; int64_t ldg;
    const float* dl; const float* w; float* dw_out;
    const void* out; int64_t ldo; int act;
    const float* drop_mask; int64_t ldm; float drop_p; const unsigned long long* drop_seed; unsigned drop_stream;
    const void* x; int64_t ldx;
    const float* mean; const float* invstd; const float* gamma;
    int C; int64_t rows_per_group; int n_groups; float count_total;
    float* dgamma; float* dbeta;
    void* gx; int64_t ldgx;
    const void* skip; int64_t lds; void* gskip; int64_t ldgs; float* dbias_skip;
    int mode;                            /* data parallel: 1 = local column sums into acc; 2 = apply with the (allreduced) acc */
    float* acc;                          /* modes 1 / 2: [n_groups, 2, C] scratch, ZEROED by the caller */
    dexgg_slab_comm comm;                /* mode 0 */
    int ctas_per_sm;
} dexgg_bn_bwd;
int dexgg_bn_bwd_slab(const dexgg_bn_bwd* args, void* stream);

/* Input BatchNorm over the BPS columns, shared by the three networks of a step: statistics of x f32 [B,C], then
 * xhat = (x - mean) * invstd as bf16 [copies * B, C] (gamma / beta are folded into the first-layer weights). */
int dexgg_bn_input_stats(const float* x, int64_t ldx, int64_t B, int C, float eps, float* mean, float* var, float* invstd,
                         float* acc /* [2,C] zeroed scratch */, void* stream);
int dexgg_xhat(const float* x, int64_t ldx, const float* mean, const float* invstd, int64_t B, int C, void* out_bf16, int copies,
               void* stream);

/* The per-row inputs that are not BPS ("tail": rot 9 | transl 3 | joint nJ, or z): up to three column blocks, f32 or f64. */
typedef struct {
    const void* p[3];
    int n[3];
    int f64;
} dexgg_tail_src;
/* bn 1: train-mode BatchNorm statistics per tail column, normalised value (no gamma / beta) written as bf16 to up to three
 * destinations (dst[i] + r*dst_ld[i] + dst_col0[i] + j), running statistics of those columns updated;  bn 0: plain copy. */
int dexgg_tail_pack(const dexgg_tail_src* src, int64_t B, int bn, float eps, float count_total, float momentum, float* mean,
                    float* invstd, float* running_mean, float* running_var, void* const* dst, const int64_t* dst_ld,
                    const int* dst_col0, int mode, float* stat, void* stream);
/* gradient wrt the raw tail inputs from the gradient wrt the normalised columns (gA[:, a0 + j] + gB[:, b0 + j]) -> out f32 [B,T] */
int dexgg_tail_bn_bwd(const dexgg_tail_src* src, int64_t B, float count_total, const float* mean, const float* invstd, const void* gA,
                      int64_t lda, int a0, const void* gB, int64_t ldb, int b0, float* out, int mode, float* sums, void* stream);

/* Score head + loss (out_success + sigmoid + BCELoss, networks.py:158,336; dexgangrasp.py:142-147,213-232,293):
 * p = sigmoid(X . w + b) over K = 512; target labels[row] or target[group]; loss[group] += lscale[group] * BCE; dl[row] =
 * dscale[group] * (p - t); db += sum dl.  wass 1: loss = lscale * p, dl = dscale * p (1 - p) (dexgangrasp.py:204-218,258). */
typedef struct {
    const void* X; int64_t ldx;
    const float* w; const float* b;
    int64_t rows_per_group; int n_groups;
    const float* labels;
    float target[2], dscale[2], lscale[2];
    float* p; float* dl; float* loss; float* db;
    int wass;
} dexgg_score_head_args;
int dexgg_score_head(const dexgg_score_head_args* args, void* stream);

/* Generator heads (joint nJ | rot6d 6 | transl 3, f32 [B,ldo]) -> fake grasp (rot [B,9] by Gram-Schmidt, transl, joint), and
 * the generator's reconstruction losses + gradient wrt the heads (dexgangrasp.py:272-307, losses.py:61-76, utils.py:380-394). */
int dexgg_gen_pose(const float* out, int64_t ldo, int n_joints, int64_t B, float* rot, float* transl, float* joint, void* stream);
int dexgg_gen_loss(const float* out, int64_t ldo, int n_joints, int64_t B, float n_total, const dexgg_tail_src* real, const float* g_pose,
                   float c_transl, float c_rot, float c_joint, void* g_heads, int64_t ldg, float* loss, float* dbias, void* stream);

/* fp32 master weights -> bf16 GEMM operands (input BatchNorm gamma / beta folded into the first layers), and the gradients of
 * the folded weights back to the .grad tensors of fc1 / fc3 and of the input BatchNorm. */
typedef struct {
    const float *W1, *W3, *b1, *b3; int Fin;
    const int *cmap, *dmap;
    int bps_src0;       /* first input column of the 4096 contiguous BPS columns */
    const float *in_gamma, *in_beta;
    void *Wb, *Wxt; int ldxt; float* bias13;
    const float *dWb, *dWxt, *db13;
    float *gW1, *gW3, *gb1, *gb3, *g_in_gamma, *g_in_beta;
} dexgg_first_layer;
typedef struct {
    const float* src; int64_t lds; int rows, cols; void* dst; int64_t ldd;
    const float* gsrc; int64_t ldgs; float* gdst; int64_t ldgd;
} dexgg_plain_job;
typedef struct {
    int n_first, n_plain;
    dexgg_first_layer first[3];
    dexgg_plain_job plain[8];
} dexgg_weight_jobs;
int dexgg_prep_weights(const dexgg_weight_jobs* jobs, void* stream);
int dexgg_unfold_grads(const dexgg_weight_jobs* jobs, void* stream);

/* nn.BatchNorm1d buffer updates (running_mean / running_var / num_batches_tracked) of every BatchNorm of a network pass. */
typedef struct {
    float *running_mean, *running_var;
    const float *mean, *second;
    long long* num_batches_tracked;
    int C, n_groups, second_is_var, same_row;
    float count, eps, momentum;
} dexgg_run_entry;
int dexgg_bn_running_multi(const dexgg_run_entry* entries, int n, void* stream);
/* data parallel: merge per-rank (mean, M2) [world, n_sets, 2, C] with Chan's formula */
int dexgg_bn_merge(const float* gathered, int world, int n_sets, int C, float n_local, float eps, float* mean, float* invstd, float* var,
                   void* stream);
int dexgg_bump_counter(unsigned long long* counter, void* stream);

int dexgg_f32_to_bf16(const float* src, int64_t lds, void* dst, int64_t ldd, int64_t rows, int cols, void* stream);
int dexgg_f32_to_bf16_t(const float* src, int64_t lds, void* dst, int64_t ldd, int64_t rows, int cols, void* stream);

/* rot6d -> rotation matrix (utils.rot_matrix_from_ortho6d, DexGanGrasp/utils/utils.py:380-394) and its backward
 * (used by the generator training step, dexgangrasp.py:776 and losses.py:67).  r6 [M,6], rot / g_rot [M,9], g_r6 [M,6]. */
int dexgg_rot6d_forward(const float* r6, int64_t M, float* rot, void* stream);
int dexgg_rot6d_backward(const float* r6, const float* g_rot, int64_t M, float* g_r6, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Training-step building blocks (DexGanGrasp.update_dexgangrasp / update_dexevaluator, models/dexgangrasp.py:739-823).
 * Linear forward / dgrad / wgrad are dexgg_linear_f32 (trans flags); below are train-mode BatchNorm1d, LeakyReLU backward,
 * bias gradients and Adam.  Row-major fp32 [M,N] activations with explicit row pitches.
 *   BatchNorm forward (nn.BatchNorm1d in training mode, networks.py:33,36,71,117,289):
 *     zero sum; dexgg_col_sum(x -> sum); [allreduce sum over ranks]; dexgg_bn_finalize(pass 0: mean = sum / count);
 *     zero sumsq; dexgg_col_centered_sumsq(x, mean -> sumsq); [allreduce]; dexgg_bn_finalize(pass 1: invstd, running stats
 *     with momentum and the unbiased variance); dexgg_bn_apply: y = act(gamma*(x-mean)*invstd + beta + residual).
 *   BatchNorm backward: g' = dexgg_lrelu_backward(g, y) when an activation was fused; dexgg_bn_backward_reduce ->
 *     sum_g (= d beta), sum_gx (= d gamma); [allreduce]; dexgg_bn_backward_apply -> gx.
 *   dexgg_adam_step: torch.optim.Adam semantics (dexgangrasp.py:97-112), one launch for all tensors of an optimizer.
 * ---------------------------------------------------------------------------------------------------------- */
int dexgg_col_sum(const float* x, int64_t ldx, int64_t M, int N, float* sums, void* stream);
int dexgg_col_centered_sumsq(const float* x, int64_t ldx, const float* mean, int64_t M, int N, float* sumsq, void* stream);
int dexgg_bn_finalize(const float* sum_or_sumsq, int pass, float count, int N, float eps, float momentum, float* mean,
                      float* invstd, float* running_mean, float* running_var, void* stream);
int dexgg_bn_apply(const float* x, int64_t ldx, const float* mean, const float* invstd, const float* gamma, const float* beta,
                   const float* residual, int64_t ldr, int act, int64_t M, int N, float* y, int64_t ldy, void* stream);
int dexgg_lrelu_backward(const float* g, int64_t ldg, const float* y, int64_t ldy, int64_t M, int N, float* out, int64_t ldo,
                         void* stream);
int dexgg_bn_backward_reduce(const float* g, int64_t ldg, const float* x, int64_t ldx, const float* mean, const float* invstd,
                             int64_t M, int N, float* sum_g, float* sum_gx, void* stream);
int dexgg_bn_backward_apply(const float* g, int64_t ldg, const float* x, int64_t ldx, const float* mean, const float* invstd,
                            const float* gamma, const float* sum_g, const float* sum_gx, float count, int64_t M, int N,
                            float* gx, int64_t ldo, void* stream);
int dexgg_adam_step(const void* chunks_dev, int n_chunks, float lr, float beta1, float beta2, float eps, float weight_decay,
                    int step, void* stream);
/* Same update with the step count in DEVICE memory (`*step_dev` = steps taken so far; incremented by the call), so that a
 * whole training step can be captured in a CUDA graph and replayed (graph_step.py). */
int dexgg_adam_step_dev(const void* chunks_dev, int n_chunks, float lr, float beta1, float beta2, float eps,
                        float weight_decay, int* step_dev, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Mean Absolute Grasp Deviation between a predicted grasp set and a ground-truth set.
 * Replaces magd_for_grasp_distribution (eval.py:101-137) and its pairwise helpers (eval.py:48-99): every predicted
 * grasp is matched to the ground-truth grasp with the smallest SQUARED translation distance (first index on ties);
 *   sums_out[0] = sum of those squared translation distances     sums_out[1] = sum of geodesic rotation angles (rad)
 *   sums_out[2] = sum of squared joint-configuration distances   sums_out[3] = coverage = distinct matches / M
 * pred_* [N,3] [N,9] [N,n_joints], gt_* [M,3] [M,9] [M,n_joints], all f32 row-major on the device; match_out [N] i32 and
 * per_grasp_out [N,3] f32 (transl, rot, joint terms) may be NULL; sums_out 4 doubles on the device.
 * ---------------------------------------------------------------------------------------------------------- */
size_t dexgg_magd_workspace_bytes(int M);
int dexgg_magd(const float* pred_transl, const float* pred_rot, const float* pred_joint, int64_t N, const float* gt_transl,
               const float* gt_rot, const float* gt_joint, int M, int n_joints, int32_t* match_out, float* per_grasp_out,
               double* sums_out, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Part-conditioned ranking: the k grasps whose translation is nearest to the centroid of an object
 * part.  Replaces filter_grasps_given_mask + sort_grasps (DexGanGrasp/utils/filter_grasps_given_mask.py:84-138;
 * dexafford_prompt.py:191-192): part points are shifted by -pc_center, cropped to |x|,|y|,|z| < crop (0.2 in the reference),
 * averaged; grasps are ordered by ascending squared distance of `transl` to that centroid.
 *   transl [n_grasps,3], part_pts [n_part,3] f32 on the device; pc_center 3 floats on the HOST; n_grasps <= 16384
 *   idx_out [k] i32 grasp rows, nearest first; neg_sqdist_out [k] f32 (may be NULL) = MINUS the squared distance of each;
 *   centroid_out [3] f32 on the device (may be NULL; NaN when the crop leaves no point, like numpy's empty mean).
 * ---------------------------------------------------------------------------------------------------------- */
size_t dexgg_part_rank_workspace_bytes(int64_t n_grasps);
int dexgg_part_rank(const float* transl, int64_t n_grasps, const float* part_pts, int64_t n_part, const float* pc_center,
                    float crop, int k, int32_t* idx_out, float* neg_sqdist_out, float* centroid_out, void* workspace,
                    size_t workspace_bytes, void* stream);

/* FP32-pipe micro-benchmark used for the BPS roofline denominator: runs `iters` rounds of 8
 * independent FFMA (kind 0), FFMA2 (kind 1), FADD2 (kind 2) or FMNMX3 (kind 3) chains per thread on a full grid and
 * returns the elapsed milliseconds in *ms and the number of lane-operations executed in *lane_ops. */
int dexgg_fp32_pipe_probe(int kind, int iters, float* ms, double* lane_ops);

#ifdef __cplusplus
}
#endif
#endif /* DEXGG_H */
