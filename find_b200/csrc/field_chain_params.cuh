// find_b200/csrc/field_chain_params.cuh — kernel parameters of the chain kernels (field_chain.cu), filled by field.cu.
#pragma once
#include "field.cuh"

namespace fb {
namespace field {

enum { TRUNK_FWD = 0, HEAD_FWD = 1, HEAD_BWD = 2, TRUNK_BWD = 3 };

struct ChainParams {
    const uint8_t* wpack;
    PackLayout lay;
    int B, V, tpf, tpf2, trunk_feet, trunk_tiles, shared;
    int n_items, save, heads_mask, n_heads;
    int enc_cached;              // TRUNK_FWD: load the encoded tiles from `enc` instead of computing them
    const float* pos;            // [trunk_feet][V][3]
    const float* Bmat;           // [3][256]
    const float* bias[P_COUNT];
    const float* biasfoot;       // [2][B][256]
    uint8_t *enc, *tact, *hact, *hdz, *doimg, *tdz;
    const uint8_t* tsrc[2];      // TRUNK_BWD input per head: [trunk_tiles] tiles = sum over feet of dZ1 (or dZ1 itself)
    uint32_t *hmask, *tmask;     // ReLU sign bits saved by the forward epilogues, [..tile..][128 rows][8 words]
    float *disp, *col, *verts;   // [B][V][3]
    const float* reg;            // [B][9] or null
    const float* avg_col;        // [3] or null
    const float* dovals;         // [2][B][V][4]
    const uint32_t* maxabs;      // bits of max |d(pre-tanh)|
    unsigned long long* dbg;     // optional timeline of CTA 0's first item (findb200_chain_trace; diagnostics only)
};

extern unsigned long long* g_chain_dbg;


int launch_chain(int mode, int nt, const ChainParams& P, int sms, cudaStream_t st);

}  // namespace field
}  // namespace fb
