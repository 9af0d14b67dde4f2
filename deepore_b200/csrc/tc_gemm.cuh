// Tensor-core kernels of the model forward (sm_100a): declarations shared by tc_gemm.cu and model.cu.
//
//   gemm_bf16     persistent, warp-specialised tcgen05 GEMM  D[M,N] = sum_chunks A_src[m0+drow.., col..+64] . B[n0.., bcol..+64]^T
//                 (bf16 operands staged by TMA with 128-byte swizzle, fp32 accumulators in TMEM, double-buffered so the
//                 epilogue of tile i overlaps the MMAs of tile i+1).  The K loop is TABLE DRIVEN: every 64-wide K chunk
//                 names its A source (up to 4 tensor maps), A column, A row shift and B column.  That one mechanism gives
//                 plain GEMMs, the 3-row sliding window of the folded CNN operator (row shift 0/1/2), (1x3) convolutions as
//                 implicit GEMM (row shift -1/0/+1), skip-path fusion (second A source) and split-bf16 (hi/lo) precision.
//   gru_step      one time step of a bidirectional GRU layer for both directions: [x_t | h_prev] . [W_ih | W_hh]^T with the
//                 r/z/n gate accumulators side by side in TMEM and the gate math, state update and direction sum in the
//                 epilogue (replaces torch.nn.GRU at /root/reference/src/modules/model.py:83-87,124-135).
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

namespace dp {

struct KChunk {          // one 64-element slice of the contraction
    int32_t a_src;       // bits 0-7: A tensor map index (0..3); bits 8-15: kind (GRU: 0 = input chunk, 1 = hidden chunk)
    int32_t a_col;       // A column (elements)
    int32_t a_drow;      // added to the tile's first row (may be negative: TMA zero-fills out of bounds)
    int32_t b_col;       // B column (elements)
};

struct GemmEpilogue {
    const float *bias;                 // [N] fp32 or null
    const __nv_bfloat16 *residual;     // optional [M][res_ld] bf16 added before the store
    long long res_ld;
    void *out;                         // bf16 or fp32, row pitch ldc elements
    __nv_bfloat16 *out_lo;             // optional bf16 remainder (x - bf16(x)) for split precision, same layout as out
    long long ldc;
    int out_fp32;
    int nblk;                          // >0: column n is stored at (n / nblk) * nblk_stride + n % nblk
    long long nblk_stride;
    int mask_period, mask_lo, mask_hi; // period>0: rows with (m % period) outside [lo,hi) are stored as ZERO (conv padding)
    float acc_scale;                   // accumulators are multiplied by this before the bias (0 = 1): undoes the power-of-two
                                       // scale the fp16 weights carry
    int compact_h, compact_keep;       // compact_h > 0: rows with (m % compact_h) >= compact_keep are NOT stored and the others move
                                       // up: row m lands at m - (compact_h - compact_keep) * (m / compact_h)  (the calls of a
                                       // pileup whose 3-row window would run into the next pileup are dropped here)
};

struct GemmArgs {
    const KChunk *chunks;
    int n_chunks;
    int M, N;
    int tiles_m, tiles_n;
    int f16;                           // 16-bit operand / output format: 0 = bf16, 1 = fp16
    GemmEpilogue epi;
};

struct GruProblem {
    const KChunk *chunks;
    int n_chunks;
    int b_row_base;                    // first row of this (layer, direction) in the stacked weight matrix
    const float *bias;                 // [4][512]: b_ir+b_hr, b_iz+b_hz, b_in, b_hn (zero padded)
    const float *h_prev;               // fp32 state (blocked layout) or null
    const __nv_bfloat16 *h_prev_bf;    // bf16 [M][512] rows: previous state when no fp32 state is kept (one-pass bf16 mode);
                                       // both null: h_prev == 0
    float *h_out;                      // fp32 state (blocked layout); null in one-pass bf16 mode (the bf16 rows ARE the state)
    __nv_bfloat16 *h_out_bf;           // bf16 [M][512]: A operand of the next step
    __nv_bfloat16 *h_out_lo;           // optional split-precision remainder
    const float *partner;              // optional: xsum = h_new + partner   (sum of directions, model.py:131-133)
    const __nv_bfloat16 *partner_bf, *partner2_bf;   // the same from bf16 rows (one-pass bf16 mode)
    __nv_bfloat16 *xsum, *xsum_lo;
    const float *partner2;             // optional: xsum2 = h_prev + partner2
    __nv_bfloat16 *xsum2, *xsum2_lo;
};

struct GruArgs {
    GruProblem prob[2];
    int n_prob;
    int M;
    int tiles_m;                       // per problem: tiles_m * 8 unit blocks
    int debug;                         // timing experiments only (env DP_GRU_DEBUG): 1 skip epilogue math, 2 one MMA per k-step
    int fast_act;                      // 1: MUFU.TANH activations; 0: exp-based
    int f16;                           // 16-bit operand / state-row format: 0 = bf16, 1 = fp16
    float acc_scale;                   // accumulators are multiplied by this before the biases (0 = 1)
};

int launch_gemm_bf16(const CUtensorMap *tmA /*[4]*/, const CUtensorMap *tmB, const GemmArgs &g, int block_n,
                     cudaStream_t st);
int launch_gru_step(const CUtensorMap *tmA /*[4]*/, const CUtensorMap *tmB, const GruArgs &g, cudaStream_t st);
// CTA-pair (cta_group::2) version: tmB96 is a map of the weight matrix with 96-row boxes (W_ih gates packed n, r, z);
// np = CTA pairs per cluster sharing their A tiles by TMA multicast (1, 2 or 4), tmA boxes are 128 / np rows
int launch_gru_step2(const CUtensorMap *tmA /*[4]*/, const CUtensorMap *tmB96, const GruArgs &g, int np, cudaStream_t st);
// wide variant: 256 rows x 128 hidden units per tile, one accumulator stage (tmA boxes of 128 rows)
int launch_gru_step2w(const CUtensorMap *tmA /*[4]*/, const CUtensorMap *tmB96, const GruArgs &g, cudaStream_t st);
// staggered wide variant: the two 64-unit halves of the 128-unit tile run their K loops GRUS_LAG chunks apart, so one half's
// accumulator drains while the tensor pipe works on the other (tmA boxes of 128 rows)
int launch_gru_step2s(const CUtensorMap *tmA /*[4]*/, const CUtensorMap *tmB96, const GruArgs &g, cudaStream_t st);
int device_sm_count();

}  // namespace dp
