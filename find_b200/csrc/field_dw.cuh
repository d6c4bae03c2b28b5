// find_b200/csrc/field_dw.cuh — descriptors of the weight-gradient GEMM list (see field_dw.cu).
#pragma once
#include "field.cuh"

namespace fb {
namespace field {

struct DwGemm {
    const uint8_t* A;        // M' operand tiles: dZ (or the d(pre-tanh) images for the 256 -> 3 layers)
    const uint8_t* Bm;       // N' operand tiles: the layer's input activations
    uint32_t a_tile_bytes, b_tile_bytes, b_img_off;
    int a_imgs, b_imgs;      // 64-column images per tile that are used (A: 4 or 1, B: 4 or 1)
    int mhalves;             // 2: M' = 256, 1: M' = 128
    int N;                   // UMMA N (256 or 64)
    int feet, tpf;           // units = feet * tpf tiles
    int a_foot_stride, b_foot_stride;   // in tiles
    long long* colsum;       // column sums of A (nullable), fixed point (kColsumFx)
    int colsum_mode;         // 0 none, 1 one vector [256], 2 per foot [feet][256]
    int colsum_cols;         // 256 or 16
    int cta0, ncta;          // partial slots [cta0, cta0 + ncta) hold the row-slices of this GEMM (one per segment, see DwSeg)
    float* partial;          // [ncta][256][256] fp32
};

// One contiguous run of units (tiles) of one GEMM, accumulated into one partial.  The host cuts the concatenated, cost-weighted
// unit list of all GEMMs into equal pieces, one piece per CTA; a piece that crosses a GEMM border is two (or more) segments.
struct DwSeg {
    short gemm, slot;        // slot = index of the partial inside the GEMM
    int u0, u1;
};
constexpr int kMaxDwSegs = 192;      // = kMaxDwCtas partial slots

constexpr int kMaxDwGemms = 24;
struct DwParams {
    DwGemm g[kMaxDwGemms];
    int n_gemms;
    DwSeg seg[kMaxDwSegs];
    unsigned short cta_seg[kMaxDwCtas + 1];   // CTA b owns segments [cta_seg[b], cta_seg[b + 1])
    unsigned long long* dbg;   // diagnostics (findb200_chain_trace): %globaltimer at the start / end of every CTA, [2 * cta]
    int cta_base;              // blockIdx.x + cta_base = the CTA's number in the plan (the plan may be launched in two parts)
};

int launch_dw(const DwParams& P, int n_ctas, cudaStream_t st, bool tpc_pairs = false);

}  // namespace field
}  // namespace fb
