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
    int dbg_flags;               // bit 0: per-chunk %globaltimer stamps (slow); cycle accounting of the issuing thread is always on with dbg
};

extern unsigned long long* g_chain_dbg;
extern int g_chain_dbg_flags;


int launch_chain(int mode, int nt, const ChainParams& P, int sms, cudaStream_t st);
// CTA-pair build (field_chain2.cu, tcgen05 cta_group::2): chooses its own tiles-per-item and item count from the shapes in P
int launch_chain2(int mode, const ChainParams& P, int sms, cudaStream_t st);
bool chain_pairs_enabled();      // FINDB200_CHAIN=1 selects the single-CTA kernels of field_chain.cu (A/B)

}  // namespace field
}  // namespace fb
