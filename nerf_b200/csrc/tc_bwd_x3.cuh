// Arguments of the fp32-grade data-gradient kernel (tc_bwd_x3.cu), shared with its host entry point (tc_train.cu).
#pragma once
#include "tc_layout.cuh"

struct BwdX3Args {
    const uint8_t* stages_hi; const uint8_t* stages_lo;   // transposed weight stages, unit l in 1..9 -> 8 stages at (l-1)*8 (tc_train.cu layout)
    const float* vec;                                     // forward epilogue vectors (head weights with gamma folded)
    const uint8_t* act;                                   // [10][tr_tiles] hidden images (hi and lo interleaved, tc_layout.cuh)
    const uint32_t* mask;                                 // [10][P_pad][8] words: ReLU mask, bit = column active
    const float2* stats;                                  // [10][P_pad] (rstd, -mean*rstd)
    const float* d_sigma; const float* d_rgb;
    uint8_t* dz;                                          // [10][tr_tiles] hidden images out
    uint8_t* dh;                                          // [tr_tiles] small-operand images out: (d_sigma, d_rgb)
    float* hsum;                                          // [4] += sums of (d_sigma, d_rgb)
    int64_t P, P_pad; int n_tiles, tr_tiles;
    uint32_t wait_hint;
    long long* prof;                                      // NERF_X3_PROF: clock64 stamps of CTA 0, its second tile: [unit][8]
};

int launch_bwd_x3(BwdX3Args& a, cudaStream_t st);
