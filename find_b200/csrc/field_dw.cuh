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
    float* colsum;           // column sums of A (nullable)
    int colsum_mode;         // 0 none, 1 one vector [256], 2 per foot [feet][256]
    int colsum_cols;         // 256 or 16
    int cta0, ncta;          // CTAs [cta0, cta0 + ncta) split the units of this GEMM
    float* partial;          // [ncta][256][256] fp32
};

constexpr int kMaxDwGemms = 24;
struct DwParams {
    DwGemm g[kMaxDwGemms];
    int n_gemms;
};

int launch_dw(const DwParams& P, int n_ctas, cudaStream_t st);

}  // namespace field
}  // namespace fb
