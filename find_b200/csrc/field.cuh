// find_b200/csrc/field.cuh — shared definitions of the neural displacement field kernels (sm_100a).
//
// Replaces NeuralDisplacementField.forward + autograd (/root/reference/src/model/model.py:393-453) and the
// registration transform that follows it (model.py:481-491).  Layer sizes: model.py:224-259,350-371.
//
// Data layout in HBM ("tile images")
//   Every activation / gradient matrix [rows x 256] lives as 128-row TILES of 64 KB; a tile is 4 IMAGES of
//   [128 rows x 64 cols] fp16 stored exactly as tcgen05 SWIZZLE_128B wants them in shared memory (row r at
//   r*128 B, its eight 16-B chunks permuted by chunk ^ (r & 7)).  One cp.async.bulk moves a tile either way, and
//   the same bytes serve as K-major operand (forward / dX GEMMs) and as MN-major operand (dW GEMMs).
//   Weights are packed once per step into [256 n x 64 k] fp16 chunk images (32 KB), W for the forward GEMMs and
//   W^T for the dX GEMMs.
#pragma once
#include "common.cuh"
#include "../../include/find_b200.h"

namespace fb {
namespace field {

constexpr int kW = 256;                 // hidden width (model.py:207 width=256)
constexpr int kMap = 256;               // Fourier mapping size (model.py:241)
constexpr int kEnc = 515;               // 3 + 2*256 (fourier_feature_transform.py:52)
constexpr int kEncChunks = 9;           // 515 padded to 9 x 64 = 576
constexpr int kTile = 128;              // rows per tile (= UMMA M)
constexpr int kImgBytes = 16384;        // [128 x 64] fp16
constexpr int kTileBytes = 65536;       // 4 images
constexpr int kEncTileBytes = kEncChunks * kImgBytes;
constexpr int kMaskTileBytes = 4096;     // ReLU sign bits of one tile: [128 rows][8 words], see epi_relu / epi_mask
constexpr int kWChunkBytes = 32768;     // [256 x 64] fp16
constexpr int kWSmallChunkBytes = 2048; // [16 x 64] fp16 (the 256 -> 3 output layers, N padded to 16)
constexpr int kRing = 3;                // weight ring stages in the chain kernels

// internal K order of the encoded input (9 chunks of 64): chunk c < 8 = [sin(f), f = 32c..32c+31 | cos(f), same f],
// chunk 8 = [x, y, z, 0...]; the reference order is [x, y, z, sin(0..255), cos(0..255)]
// (fourier_feature_transform.py:52).  Returns the reference column of internal column kk, -1 for padding.
__host__ __device__ inline int enc_ref_col(int kk) {
    if (kk < 512) {
        const int c = kk >> 6, w = kk & 63;
        return 3 + c * 32 + (w & 31) + ((w >> 5) ? 256 : 0);
    }
    if (kk < 515) return kk - 512;
    return -1;
}

// parameter order everywhere: base.0,2,4,6,8 | mlp_disp.0,2,4,6 | mlp_col.0,2,4,6   (model.py:250-257,351-371)
enum { P_T0 = 0, P_T1, P_T2, P_T3, P_T4, P_D1, P_D2, P_D3, P_D4, P_C1, P_C2, P_C3, P_C4, P_COUNT };

struct GemmW {            // one streamed weight operand (a run of chunk images inside the packed blob)
    uint32_t off;         // byte offset in wpack
    uint32_t chunk_bytes; // 32768, or 2048 for the N=16 output layers
    uint16_t n_chunks;    // K / 64
    uint16_t k16_last;    // 16-wide K steps in the last chunk (1..4); all other chunks use 4
    uint16_t N;           // UMMA N (256 or 16)
    uint16_t group;       // chunks that travel in ONE ring slot (1; 4 for the 2 KB chunks of the N=16 output layers)
};

struct PackLayout {
    GemmW fwd[P_COUNT];        // W   images: forward GEMMs (for P_D1 / P_C1 only the 256 trunk-feature columns)
    GemmW bwd[P_COUNT];        // W^T images: dX GEMMs (bwd[P_T0] unused: the input has no gradient)
    uint32_t total_bytes;
};

// ---- workspace carve-up (all offsets in bytes from the workspace base, 1024-aligned) -----------------------
struct Workspace {
    size_t wpack;        // packed weights
    size_t biasfoot;     // float [2][B][256]: layer-1 bias of each head with the latent term folded in
    size_t enc;          // trunk: encoded input tiles          [trunk_tiles][9 images]
    size_t tact;         // trunk: a0..a4 tiles                 [5][trunk_tiles]
    size_t hact;         // heads: a1..a3 tiles                 [2][3][B*tpf2]
    size_t hdz;          // heads: dZ3,dZ2,dZ1 tiles            [2][3][B*tpf2]
    size_t hsum;         // heads: sum over feet of dZ1         [2][trunk_tiles] (shared template only)
    size_t hmask;        // heads: ReLU bit masks of a1..a3     [2][3][B*tpf2] x 4 KB
    size_t tmask;        // trunk: ReLU bit masks of a0..a4     [5][trunk_tiles] x 4 KB
    size_t doimg;        // heads: d(pre-tanh) images           [2][B*tpf2] x 16 KB
    size_t tdz;          // trunk: dZ4..dZ0 tiles               [5][trunk_tiles]
    size_t dovals;       // float [2][B][V][4]: unscaled d(pre-tanh) values
    size_t scal;         // uint32 maxabs bits, + spare
    size_t regsums;      // float [B][12]: sum g (3), Q[i][j] = sum u_i g_j (9)
    size_t colsum;       // float: [2][B][256] layer-1 per-foot sums | [2][2][256] head layers 2,3 | [2][16] out | [5][256] trunk
    size_t partial;      // float [n_partials][256][256]
    size_t total;
    int tpf, tpf2, trunk_feet, trunk_tiles, n_partials;
};

constexpr int kMaxDwCtas = 192;

// step.cu -> field.cu (inside the library, not part of the C ABI): the region of the field workspace that
// findb200_field_backward clears before it accumulates (global max | per-foot reg sums | column sums), so that a fused step
// can have an earlier kernel clear it and tell the backward so.
void field_bwd_zero_region(int nfeet, int V, int pos_shared, size_t* offset, size_t* bytes);
extern thread_local bool g_field_bwd_prezeroed;
// d loss / d verts handed over as fixed point (see fx_add): grad_verts[b][v][k] = fx[b][v][k] * scale[b].  Set by step.cu around
// findb200_field_backward when the fused Chamfer tail produced the gradient; null = the float grad_verts argument is used.
extern thread_local const long long* g_grad_verts_fx;
extern thread_local const float* g_grad_verts_fx_scale;

// Order-independent accumulation: sums that many CTAs build with atomics (column sums of dZ, registration sums, the Chamfer
// scatter) are kept as 64-bit fixed point, because integer addition is associative — two runs of a step give bit-identical
// gradients whatever the arrival order — and at these scales the quantum is far below an fp32 ulp of the result.
constexpr float kColsumFx = 16777216.f;        // 2^24: |sum of scaled dZ| < 2^39, quantum 6e-8
constexpr float kRegFx = 4294967296.f;         // 2^32: registration sums, quantum 2.3e-10
FB_DEVINL void fx_add(long long* dst, float v, float scale) {
    atomicAdd(reinterpret_cast<unsigned long long*>(dst), (unsigned long long)__double2ll_rn((double)v * (double)scale));
}
FB_DEVINL float fx_get(const long long* src, float scale) { return (float)((double)(*src) / (double)scale); }

}  // namespace field
}  // namespace fb
