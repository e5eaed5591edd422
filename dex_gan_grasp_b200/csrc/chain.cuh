// Fused whole-network tensor-core kernel (chain_umma.cu): interface used by nets.cu.
//
// One CTA takes a tile of 128 grasps through EVERY layer of the DexGenerator, the DexEvaluator or the Discriminator (eval-mode BatchNorm,
// per-object hoisted BPS projection): activations never leave the SM (shared memory as the tcgen05 A operand, TMEM for the
// accumulators and the packed residual), the weights are streamed from L2 as a pre-swizzled tile stream by bulk TMA
// copies, and the heads (rot6d -> SO(3) + palm offset + joints, or sigmoid score) are applied from TMEM in the last
// epilogue.  Reference: ResBlock.forward / Generator.forward / DexEvaluator.forward, DexGanGrasp/models/networks.py:43-56,
// 88-103, 148-159, 315-337; utils.rot_matrix_from_ortho6d (utils/utils.py:380-394); utils.translate_along_axis (utils/utils.py:19-48).
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"

namespace dexgg {

constexpr int CH_MAX_BLOCKS = 3;

struct ChainPackSrc {          // fp32 packed weights of nets.cu (device pointers)
    int kind;                  // DEXGG_NET_GENERATOR, DEXGG_NET_EVALUATOR or DEXGG_NET_DISCRIMINATOR
    int n_blocks;
    int tailw;                 // 16 (generator, evaluator) or 32 (discriminator: n_tail <= 24 pose + joint inputs)
    int n_tail;                // real tail inputs
    int nh;                    // head rows: nJ + 9 (generator) or 1 (evaluator)
    const float* Wt[CH_MAX_BLOCKS];  // [768, KT]: cols [X 512 (if has_x) | tail (tailw)]
    int KT[CH_MAX_BLOCKS];
    int has_x[CH_MAX_BLOCKS];
    const float* W2[CH_MAX_BLOCKS];  // [512, 256]
    const float* sc[CH_MAX_BLOCKS];  // [768] first-layer affine scale (bn1 on the fc1 rows, 1 on the fc3 rows): folded into Wt
    const float* sc2[CH_MAX_BLOCKS]; // [512] bn2 scale: folded into W2
    const float* Wh;                 // [nh, 512]
};

struct ChainArgs {
    int kind, n_blocks, nJ, nh;
    const __nv_bfloat16* stream;     // weight tile stream built by chain_pack
    int has_x[CH_MAX_BLOCKS];
    const float* sh2[CH_MAX_BLOCKS];  // [512] bn2 shift (fc2 bias folded); the bn1 / bn2 scales live in the weight stream
    const float* base;               // [n_obj, 768 * n_blocks] per-object hoisted projection WITH the first-layer affine
                                     // applied (sc * (bps.W + b) + sh)
    int64_t n_obj, rows_per_obj, M;
    const float* bh;                 // head bias
    // generator
    const float* z;                  // [M, zD]
    int zD;
    float z_offset;
    float *rot, *transl, *joint, *rot6d;
    // evaluator
    const float *in_rot, *in_transl; // [M,9], [M,3]
    const float *in_joint;           // discriminator: [M,nJ]
    const float *tail_a, *tail_c;    // input BatchNorm affine of the pose (+ joint) columns
    float* p;                        // [M]
    long long* debug_clock;          // timing experiment only: [512] clock64 samples of CTA 0 (MMA issuer | epilogue warp)
    int debug_no_stream;             // timing experiment only (DEXGG_CHAIN_NOSTREAM=1): weights are not loaded
};

size_t chain_stream_bytes(int kind, int n_blocks, const int* has_x);
int chain_pack(const ChainPackSrc& s, __nv_bfloat16* stream, cudaStream_t st);
int chain_forward(const ChainArgs& a, cudaStream_t st);

}  // namespace dexgg
