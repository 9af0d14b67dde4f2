// deePore model forward on B200: C-ABI (include/deepore_b200.h, dp_model_*) over the tcgen05 kernels of tc_gemm.cu.
//
// Replaces /root/reference/src/modules/model.py Model.forward (:137-156) and the decode loop of
// /root/reference/src/test.py (:59-89) for whole pileup images:
//
//   u8 image rows [n_rows][W*4]  --convert-->  bf16 (values 0..255 are exact in bf16)
//   folded CNN+Linear operator   y[row][t][:] = sum_s Weff[t][s] . img[row+s] + beff[t]       (one tcgen05 GEMM, K = 3*W*4)
//        The five residual blocks contain convolutions and eval-mode BatchNorm only (model.py:93-105: no activation),
//        so CNN + reinterpreting view (:144-149) + cnn_fc_rnn (:151) is ONE affine map of the three input rows; the host
//        composes it once per checkpoint (deepore_b200/fold.py) and the /255 of ToTensor is folded in as well.
//   3 x bidirectional GRU        9 gru_step launches (3 layers x 3 time steps, both directions per launch), h0 of layer k =
//        h_n of layer k-1 (model.py:124-135), directions summed
//   fc1 + argmax                 logits[row][t][c] (model.py:154-155), first maximum wins (torch.topk, test.py:68)
//   vote                         3-way majority over (row-2,t=2),(row-1,t=1),(row,t=0), smallest class on a 3-way tie
//                                (test.py:23-24,71-82)
#include <cuda_fp16.h>
#include <math.h>
#include <stdlib.h>

#include <vector>

#include "tc_gemm.cuh"

namespace dp {

constexpr int HP = 512;            // padded hidden size
constexpr int NSLOT_BF = 30;       // bf16 arena slots  (y: 0..2; layer k: 3+9k + {hf 0..2, hb 3..5, xn 6..8})
constexpr int NSLOT_F32 = 18;      // fp32 state slots  (layer k: 6k + {hf 0..2, hb 3..5})

static inline uint16_t f2bf(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
static inline float bf2f(uint16_t h) {
    uint32_t u = (uint32_t)h << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}
// 16-bit operand formats of the tensor-core path: bf16 (8-bit significand) or fp16 (11-bit significand, range kept by
// power-of-two weight scales that the epilogues divide out again)
static inline uint16_t f2h16(float f, bool f16) {
    if (!f16) return f2bf(f);
    if (f > 65504.f) f = 65504.f;
    if (f < -65504.f) f = -65504.f;
    const __half h = __float2half_rn(f);
    uint16_t u;
    memcpy(&u, &h, 2);
    return u;
}
static inline float h162f(uint16_t u, bool f16) {
    if (!f16) return bf2f(u);
    __half h;
    memcpy(&h, &u, 2);
    return __half2float(h);
}
// largest power of two s with max|w| * s <= 16384 (fp16: headroom of 4 below the largest finite value); 1 for bf16
static inline float pow2_scale(float maxabs, bool f16) {
    if (!f16 || !(maxabs > 0.f)) return 1.f;
    int e = 0;
    frexpf(16384.f / maxabs, &e);          // 16384 / maxabs = m * 2^e, m in [0.5, 1)
    return ldexpf(1.f, e - 1);
}

struct Model {
    int in_ch, width, hidden, n_class, seq_len, n_layers;
    int row_elems;                 // width * in_ch (800)
    int kfold_pad;                 // per-row-slot K of the folded operator padded to 64 (832)
    int f16;                       // 16-bit operand format: 0 = bf16, 1 = fp16
    int fold_split;                // folded-operator weights as hi + lo (two passes; the u8 image is exact in either format)
    int gru_split;                 // bit k: GRU layer k runs the three split passes (hi.hi, hi.lo, lo.hi)
    int alloc_lo;                  // lo copies of the activations exist (gru_split != 0)
    int fast_act;                  // MUFU.TANH activations instead of exp-based ones
    float fold_scale = 1.f, gru_scale = 1.f;   // power-of-two scales folded into the 16-bit weights (fp16 range)
    long long mcap;                // row capacity (multiple of 128)
    int device;
    // device buffers
    __nv_bfloat16 *img_bf = nullptr;      // [mcap + 128][row_elems]
    __nv_bfloat16 *arena = nullptr;       // [NSLOT_BF][mcap][HP]
    __nv_bfloat16 *arena_lo = nullptr;    // split modes only
    float *state = nullptr;               // [NSLOT_F32][mcap][HP]
    int bf_state = 0;                     // one-pass bf16 mode on the CTA-pair kernel: number of leading GRU layers whose state
                                          // lives in the bf16 rows only (no fp32 copy); DP_GRU_STATE = f32 | mixed | bf16
    __nv_bfloat16 *w_fold = nullptr;      // [3*HP][kcols_fold]
    float *b_fold = nullptr;              // [3*HP]
    __nv_bfloat16 *w_gru = nullptr;       // [n_layers*2*3*HP][kcols_gru]
    float *b_gru = nullptr;               // [n_layers*2][4][HP]
    float *w_fc1 = nullptr, *b_fc1 = nullptr;   // [n_class][HP], [n_class]
    KChunk *tables = nullptr;
    int fold_tab_off = 0, fold_tab_n = 0;
    int fold3_tab_off = 0;         // the same chunks addressing the image as [calls][seq_len * row_elems] (forward_triples)
    int gru_tab_off[3][3][2], gru_tab_n[3][3][2];
    long long kcols_fold = 0, kcols_gru = 0;
    CUtensorMap tm_img3;           // image buffer viewed as one row of seq_len * row_elems elements per forward call
    CUtensorMap tm_img, tm_arena, tm_arena_lo, tm_wfold, tm_wgru, tm_wgru96, tm_arena_mc, tm_arena_lo_mc;
    int gru_np = 1;                // CTA pairs per cluster of the pair kernel (A tiles multicast between them); env DP_GRU_NP
    int gru_wide = 0;              // env DP_GRU=wide: 256 x 128-unit tiles (gru_step2w_kernel); DP_GRU=stag: the same with
                                   // staggered halves (gru_step2s_kernel, value 2)
    int gru_2cta = 1;              // CTA-pair GRU kernel (env DP_GRU=1cta selects the single-CTA kernel)
    // optional per-kernel timing (bench.py roofline): one event per launch boundary, harvested by dp_model_profile_read
    int profile = 0;
    std::vector<cudaEvent_t> prof_events;      // PROF_MARKS events per profiled forward call
};
constexpr int PROF_MARKS = 14;                 // start, convert, fold, 9 x gru_step (up to), fc1, (spare)

static void free_model(Model *m) {
    if (!m) return;
    cudaFree(m->img_bf); cudaFree(m->arena); cudaFree(m->arena_lo); cudaFree(m->state); cudaFree(m->w_fold);
    cudaFree(m->b_fold); cudaFree(m->w_gru); cudaFree(m->b_gru); cudaFree(m->w_fc1); cudaFree(m->b_fc1);
    cudaFree(m->tables);
    delete m;
}

static inline int bf_slot_y(int t) { return t; }
static inline int bf_slot_hf(int k, int t) { return 3 + 9 * k + t; }
static inline int bf_slot_hb(int k, int t) { return 3 + 9 * k + 3 + t; }
static inline int bf_slot_xn(int k, int t) { return 3 + 9 * k + 6 + t; }
static inline int f32_slot_hf(int k, int t) { return 6 * k + t; }
static inline int f32_slot_hb(int k, int t) { return 6 * k + 3 + t; }

// ------------------------------------------------------------------------------------------------
// small CUDA-core kernels around the GEMMs

// u8 -> bf16, 16 bytes in / 32 bytes out per thread (exact: integers <= 255 fit the 8-bit bf16 significand)
__global__ void __launch_bounds__(256) u8_to_bf16_kernel(const uint4 *__restrict__ in, uint4 *__restrict__ out, long long n16,
                                                         int f16) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n16) return;
    const uint4 v = __ldg(in + i);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t o[8];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float b0 = (float)(w[k] & 0xff), b1 = (float)((w[k] >> 8) & 0xff);
        const float b2 = (float)((w[k] >> 16) & 0xff), b3 = (float)(w[k] >> 24);
        o[2 * k] = tc::pack16(b0, b1, f16 != 0);       // integers <= 255 are exact in bf16 and in fp16
        o[2 * k + 1] = tc::pack16(b2, b3, f16 != 0);
    }
    out[2 * i] = make_uint4(o[0], o[1], o[2], o[3]);
    out[2 * i + 1] = make_uint4(o[4], o[5], o[6], o[7]);
}

// fc1 (model.py:90,154) + top-1 (test.py:68).  out = hf + hb of the last GRU layer (fp32, blocked layout of
// gru_state_off in tc_gemm.cu).  One warp per (32 consecutive rows, t): lane = row, so every load instruction moves 512
// contiguous bytes of the blocked state; each lane accumulates its own row's dot products (no shuffles), the fc1
// weights are broadcast from shared memory.
constexpr int FC1_CMAX = 4;                            // classes per pass over the state
__global__ void __launch_bounds__(128) fc1_argmax_kernel(const float *__restrict__ hf, const float *__restrict__ hb,
                                                         long long slot_stride, const float *__restrict__ w,
                                                         const float *__restrict__ b, int n_class, int seq_len,
                                                         long long n_rows, float *__restrict__ logits,
                                                         uint8_t *__restrict__ pred, int ckeep, int cgap) {
    extern __shared__ float s_w[];                     // [n_class][HP]
    for (int i = threadIdx.x; i < n_class * HP; i += blockDim.x) s_w[i] = w[i];
    __syncthreads();
    const long long wid = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    const long long n_rb = (n_rows + 31) >> 5;
    if (wid >= n_rb * seq_len) return;
    const long long rb = wid / seq_len;
    const int t = (int)(wid % seq_len);
    const long long m = rb * 32 + lane;
    const long long base = (((m >> 7) * 8) * 16 * 128 + (m & 127)) * 4;     // offset of (m, u = 0)
    const long long mo = ckeep > 0 ? m + (long long)cgap * (m / ckeep) : m;   // output row: compacted calls go back to their row
    const float *pf = hf + t * slot_stride + base, *pb = hb + t * slot_stride + base;
    float best = 0.f;
    int besti = 0;
    for (int c0 = 0; c0 < n_class; c0 += FC1_CMAX) {
        float acc[FC1_CMAX];
#pragma unroll
        for (int c = 0; c < FC1_CMAX; c++) acc[c] = 0.f;
#pragma unroll 4
        for (int q = 0; q < HP / 4; q++) {             // q = (ub*4 + cb)*4 + k: consecutive float4 groups are 128 rows apart
            const float4 a = *reinterpret_cast<const float4 *>(pf + (long long)q * 512);
            const float4 b4 = *reinterpret_cast<const float4 *>(pb + (long long)q * 512);
            const float x0 = a.x + b4.x, x1 = a.y + b4.y, x2 = a.z + b4.z, x3 = a.w + b4.w;
#pragma unroll
            for (int c = 0; c < FC1_CMAX; c++) {
                if (c0 + c < n_class) {
                    const float4 wv = *reinterpret_cast<const float4 *>(s_w + (c0 + c) * HP + 4 * q);
                    acc[c] = fmaf(x0, wv.x, acc[c]); acc[c] = fmaf(x1, wv.y, acc[c]);
                    acc[c] = fmaf(x2, wv.z, acc[c]); acc[c] = fmaf(x3, wv.w, acc[c]);
                }
            }
        }
#pragma unroll
        for (int c = 0; c < FC1_CMAX; c++) {
            if (c0 + c < n_class) {
                const float v = acc[c] + __ldg(b + c0 + c);
                if (m < n_rows && logits) logits[(mo * seq_len + t) * n_class + c0 + c] = v;
                if (c0 + c == 0 || v > best) { best = v; besti = c0 + c; }
            }
        }
    }
    if (m < n_rows && pred) pred[mo * seq_len + t] = (uint8_t)besti;
}

// The same over the bf16 rows [row][512] of the arena (one-pass bf16 mode keeps no fp32 state): lane handles 16
// consecutive units (32 bytes of hf and of hb).
__global__ void __launch_bounds__(256) fc1_argmax_bf_kernel(const __nv_bfloat16 *__restrict__ hf,
                                                            const __nv_bfloat16 *__restrict__ hb, long long slot_stride,
                                                            const float *__restrict__ w, const float *__restrict__ b,
                                                            int n_class, int seq_len, long long n_rows,
                                                            float *__restrict__ logits, uint8_t *__restrict__ pred, int f16,
                                                            int ckeep, int cgap) {
    const long long wid = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (wid >= n_rows * seq_len) return;
    const long long m = wid / seq_len;
    const int t = (int)(wid % seq_len);
    const long long mo = ckeep > 0 ? m + (long long)cgap * (m / ckeep) : m;
    const uint4 *pf = reinterpret_cast<const uint4 *>(hf + t * slot_stride + m * HP + lane * 16);
    const uint4 *pb = reinterpret_cast<const uint4 *>(hb + t * slot_stride + m * HP + lane * 16);
    const uint4 f0 = pf[0], f1 = pf[1], b0 = pb[0], b1 = pb[1];
    const uint32_t fw[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
    const uint32_t bw[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    float x[16];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        float a0, a1, c0, c1;
        tc::unpack16(fw[i], a0, a1, f16 != 0);
        tc::unpack16(bw[i], c0, c1, f16 != 0);
        x[2 * i] = a0 + c0;
        x[2 * i + 1] = a1 + c1;
    }
    float best = 0.f;
    int besti = 0;
    for (int c = 0; c < n_class; c++) {
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float4 wv = __ldg(reinterpret_cast<const float4 *>(w + c * HP + lane * 16 + 4 * j));
            acc = fmaf(x[4 * j], wv.x, acc); acc = fmaf(x[4 * j + 1], wv.y, acc);
            acc = fmaf(x[4 * j + 2], wv.z, acc); acc = fmaf(x[4 * j + 3], wv.w, acc);
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
        acc += __ldg(b + c);
        if (lane == 0 && logits) logits[(mo * seq_len + t) * n_class + c] = acc;
        if (c == 0 || acc > best) { best = acc; besti = c; }
    }
    if (lane == 0 && pred) pred[mo * seq_len + t] = (uint8_t)besti;
}

// 3-way vote (test.py:71-82, most_common :23-24).  pred [n_windows][H][3]; calls [n_windows][H] int8, -1 where the
// reference scores nothing (rows 0,1 and the last rows without a full 3-row window).
__global__ void __launch_bounds__(256) vote_kernel(const uint8_t *__restrict__ pred, int n_windows, int H,
                                                   int8_t *__restrict__ calls, const uint8_t *__restrict__ labels,
                                                   unsigned long long *__restrict__ confusion, int n_class) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n_windows * H) return;
    const int row = (int)(i % H);
    const long long wbase = (i / H) * H;
    int call = -1;
    if (row >= 2 && row + 3 <= H) {
        const int a = pred[(wbase + row - 2) * 3 + 2], b = pred[(wbase + row - 1) * 3 + 1], c = pred[(wbase + row) * 3 + 0];
        // mode; all distinct -> max(set(l), key=l.count) walks the set of small ints in ascending order -> smallest
        if (a == b || a == c) call = a;
        else if (b == c) call = b;
        else call = min(a, min(b, c));
        if (labels && confusion) {
            const int t = labels[i];
            if (t < n_class && call < n_class) atomicAdd(&confusion[t * n_class + call], 1ull);
        }
    }
    calls[i] = (int8_t)call;
}

// Cross-entropy of raw logits against integer labels, one thread per (call, t) element: loss = logsumexp(logits) - logits[label]
// (nn.CrossEntropyLoss of the reference's validation pass, src/train.py:36,75 -- the reduction over the batch is the caller's).
// A label outside [0, n_class) gives +inf and is counted (torch raises on it).
__global__ void __launch_bounds__(256) cross_entropy_kernel(const float *__restrict__ logits, const uint8_t *__restrict__ labels,
                                                            long long n, int n_class, float *__restrict__ loss,
                                                            double *__restrict__ sum, int *__restrict__ bad) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    float l = 0.f;
    if (i < n) {
        const float *x = logits + i * n_class;
        float mx = x[0];
        for (int c = 1; c < n_class; c++) mx = fmaxf(mx, x[c]);
        float se = 0.f;
        for (int c = 0; c < n_class; c++) se += expf(x[c] - mx);
        const int y = labels[i];
        if (y < n_class) l = logf(se) + mx - x[y];
        else { l = INFINITY; if (bad) atomicAdd(bad, 1); }
        if (loss) loss[i] = l;
    }
    if (sum) {                                   // block sum in double, one atomic per block (order-dependent in the last bits)
        __shared__ double s_part[8];
        double v = (i < n && isfinite(l)) ? (double)l : 0.0;
#pragma unroll
        for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int w = 0; w < 8; w++) t += s_part[w];
            atomicAdd(sum, t);
        }
    }
}

}  // namespace dp

// ================================================================================================
// host side: weight packing, chunk tables, launches
namespace dp {

template <typename T>
static int upload(T **dst, const std::vector<T> &src) {
    DP_CUDA_CHECK(cudaMalloc((void **)dst, src.size() * sizeof(T)));
    DP_CUDA_CHECK(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
    return DP_OK;
}

// bf16 hi part, and the bf16 of the remainder when `lo` is wanted (split precision)
static inline void split16(float v, uint16_t &hi, uint16_t &lo, bool f16) {
    hi = f2h16(v, f16);
    lo = f2h16(v - h162f(hi, f16), f16);
}

static int build_model(Model *m, const dp_model_weights *w) {
    const int S = m->seq_len, Hd = m->hidden, RE = m->row_elems, KP = m->kfold_pad;
    const bool f16 = m->f16 != 0;
    // ---- folded operator: rows t*HP + h, columns s*KP + (w*in_ch + ch); lo copy (precise) after the hi columns
    const long long kfold = (long long)S * KP;
    m->kcols_fold = kfold * (m->fold_split ? 2 : 1);
    {
        float mx = 0.f;
        for (size_t i = 0; i < (size_t)S * Hd * S * RE; i++) mx = fmaxf(mx, fabsf(w->fold_w[i]));
        m->fold_scale = pow2_scale(mx, f16);
        std::vector<uint16_t> B((size_t)S * HP * m->kcols_fold, 0);
        std::vector<float> bias((size_t)S * HP, 0.f);
        for (int t = 0; t < S; t++)
            for (int h = 0; h < Hd; h++) {
                const float *src = w->fold_w + ((size_t)t * Hd + h) * S * RE;
                uint16_t *row = &B[((size_t)t * HP + h) * m->kcols_fold];
                for (int s = 0; s < S; s++)
                    for (int e = 0; e < RE; e++) {
                        uint16_t hi, lo;
                        split16(src[(size_t)s * RE + e] * m->fold_scale, hi, lo, f16);
                        row[(size_t)s * KP + e] = hi;
                        if (m->fold_split) row[kfold + (size_t)s * KP + e] = lo;
                    }
                bias[(size_t)t * HP + h] = w->fold_b[(size_t)t * Hd + h];
            }
        int rc = upload((uint16_t **)&m->w_fold, B);
        if (rc) return rc;
        rc = upload(&m->b_fold, bias);
        if (rc) return rc;
    }
    // ---- GRU: rows ((k*2+d)*8 + ub)*192 + gate*64 + j  <- unit u = ub*64+j; columns [W_ih | W_hh] (512 each).
    // For the CTA-pair kernel the W_ih columns hold the gates in the order (n, r, z) instead of (r, z, n), so that the
    // input chunks fill accumulator columns [W_in x | r | z] and the hidden chunks [r | z | W_hn h] one to the right:
    // both are ONE contiguous N=192 MMA (gru_step2_kernel).
    {
        const char *e = getenv("DP_GRU");
        m->gru_2cta = (e && !strcmp(e, "1cta")) ? 0 : 1;
        m->gru_wide = (e && !strcmp(e, "wide")) ? 1 : (e && !strcmp(e, "stag")) ? 2 : 0;
        const char *e2 = getenv("DP_GRU_NP");
        if (e2 && (atoi(e2) == 1 || atoi(e2) == 2 || atoi(e2) == 4)) m->gru_np = atoi(e2);
    }
    const long long kgru = 2 * HP;
    m->kcols_gru = kgru * (m->gru_split ? 2 : 1);
    {
        const int L = m->n_layers;
        float mx = 0.f;
        for (int k = 0; k < L; k++)
            for (int d = 0; d < 2; d++)
                for (size_t i = 0; i < (size_t)3 * Hd * Hd; i++)
                    mx = fmaxf(mx, fmaxf(fabsf(w->gru_w_ih[k][d][i]), fabsf(w->gru_w_hh[k][d][i])));
        m->gru_scale = pow2_scale(mx, f16);
        std::vector<uint16_t> B((size_t)L * 2 * 3 * HP * m->kcols_gru, 0);
        std::vector<float> bias((size_t)L * 2 * 4 * HP, 0.f);
        for (int k = 0; k < L; k++)
            for (int d = 0; d < 2; d++) {
                const float *wih = w->gru_w_ih[k][d], *whh = w->gru_w_hh[k][d];
                const float *bih = w->gru_b_ih[k][d], *bhh = w->gru_b_hh[k][d];
                for (int u = 0; u < Hd; u++) {
                    const int ub = u / 64, j = u % 64;
                    for (int gate = 0; gate < 3; gate++) {
                        const size_t r0 = ((size_t)(k * 2 + d) * 8 + ub) * 192 + j;
                        uint16_t *row = &B[(r0 + gate * 64) * m->kcols_gru];                                   // W_hh part
                        uint16_t *rowi = &B[(r0 + (m->gru_2cta ? (gate + 1) % 3 : gate) * 64) * m->kcols_gru];  // W_ih part
                        const float *si = wih + ((size_t)gate * Hd + u) * Hd, *sh = whh + ((size_t)gate * Hd + u) * Hd;
                        for (int c = 0; c < Hd; c++) {
                            uint16_t hi, lo;
                            split16(si[c] * m->gru_scale, hi, lo, f16);
                            rowi[c] = hi;
                            if (m->gru_split) rowi[kgru + c] = lo;
                            split16(sh[c] * m->gru_scale, hi, lo, f16);
                            row[HP + c] = hi;
                            if (m->gru_split) row[kgru + HP + c] = lo;
                        }
                    }
                    float *bb = &bias[(size_t)(k * 2 + d) * 4 * HP];
                    bb[0 * HP + u] = bih[u] + bhh[u];
                    bb[1 * HP + u] = bih[Hd + u] + bhh[Hd + u];
                    bb[2 * HP + u] = bih[2 * Hd + u];
                    bb[3 * HP + u] = bhh[2 * Hd + u];
                }
            }
        int rc = upload((uint16_t **)&m->w_gru, B);
        if (rc) return rc;
        rc = upload(&m->b_gru, bias);
        if (rc) return rc;
    }
    {
        std::vector<float> W((size_t)m->n_class * HP, 0.f), b(m->n_class);
        for (int c = 0; c < m->n_class; c++) {
            for (int u = 0; u < Hd; u++) W[(size_t)c * HP + u] = w->fc1_w[(size_t)c * Hd + u];
            b[c] = w->fc1_b[c];
        }
        int rc = upload(&m->w_fc1, W);
        if (rc) return rc;
        rc = upload(&m->b_fc1, b);
        if (rc) return rc;
    }
    // ---- K-chunk tables
    std::vector<KChunk> tab;
    const int nfc = KP / 64;
    m->fold_tab_off = 0;
    for (int pass = 0; pass < (m->fold_split ? 2 : 1); pass++)      // the u8 image is exact: only the weights are split
        for (int s = 0; s < S; s++)
            for (int i = 0; i < nfc; i++) tab.push_back(KChunk{0, 64 * i, s, (int)(pass * kfold) + s * KP + 64 * i});
    m->fold_tab_n = (int)tab.size();
    // Model.forward(x, hidden) of the reference API hands over ONE 3-row window per sample (test.py:63-66): the image buffer
    // is then [calls][3 * row_elems] and slot s of call m sits at column s * row_elems of row m.  The padded tail of a slot's
    // last chunk reads the first pixels of the next slot (or TMA zero fill); the weight columns there are zero.
    m->fold3_tab_off = (int)tab.size();
    for (int pass = 0; pass < (m->fold_split ? 2 : 1); pass++)
        for (int s = 0; s < S; s++)
            for (int i = 0; i < nfc; i++) tab.push_back(KChunk{3, s * RE + 64 * i, 0, (int)(pass * kfold) + s * KP + 64 * i});
    const long long mc = m->mcap;
    for (int k = 0; k < m->n_layers; k++)
        for (int st = 0; st < 3; st++)
            for (int d = 0; d < 2; d++) {
                const int t = d == 0 ? st : 2 - st;
                const int xslot = k == 0 ? bf_slot_y(t) : bf_slot_xn(k - 1, t);
                int hslot = -1;                      // -1: h_prev is zero (layer 0, first step)
                if (st > 0) hslot = d == 0 ? bf_slot_hf(k, t - 1) : bf_slot_hb(k, t + 1);
                else if (k > 0) hslot = d == 0 ? bf_slot_hf(k - 1, 2) : bf_slot_hb(k - 1, 0);
                m->gru_tab_off[k][st][d] = (int)tab.size();
                // input chunks first (the MMA issuer relies on it), then hidden chunks.  a_src: 1 = arena, 2 = arena_lo
                const int npass = ((m->gru_split >> k) & 1) ? 3 : 1;
                for (int pass = 0; pass < npass; pass++) {       // (hi,hi) (hi,lo) (lo,hi)
                    const int asrc = pass == 2 ? 2 : 1, boff = pass == 1 ? (int)kgru : 0;
                    for (int i = 0; i < HP / 64; i++)
                        tab.push_back(KChunk{asrc | (0 << 8), 64 * i, (int)(xslot * mc), boff + 64 * i});
                }
                if (hslot >= 0)
                    for (int pass = 0; pass < npass; pass++) {
                        const int asrc = pass == 2 ? 2 : 1, boff = pass == 1 ? (int)kgru : 0;
                        for (int i = 0; i < HP / 64; i++)
                            tab.push_back(KChunk{asrc | (1 << 8), 64 * i, (int)(hslot * mc), boff + HP + 64 * i});
                    }
                m->gru_tab_n[k][st][d] = (int)tab.size() - m->gru_tab_off[k][st][d];
            }
    int rc = upload(&m->tables, tab);
    if (rc) return rc;
    // ---- activations
    DP_CUDA_CHECK(cudaMalloc((void **)&m->img_bf, (size_t)(mc + 128) * RE * 2));
    DP_CUDA_CHECK(cudaMemset(m->img_bf, 0, (size_t)(mc + 128) * RE * 2));
    DP_CUDA_CHECK(cudaMalloc((void **)&m->arena, (size_t)NSLOT_BF * mc * HP * 2));
    DP_CUDA_CHECK(cudaMemset(m->arena, 0, (size_t)NSLOT_BF * mc * HP * 2));
    if (m->alloc_lo) {
        DP_CUDA_CHECK(cudaMalloc((void **)&m->arena_lo, (size_t)NSLOT_BF * mc * HP * 2));
        DP_CUDA_CHECK(cudaMemset(m->arena_lo, 0, (size_t)NSLOT_BF * mc * HP * 2));
    }
    DP_CUDA_CHECK(cudaMalloc((void **)&m->state, (size_t)NSLOT_F32 * mc * HP * 4));
    DP_CUDA_CHECK(cudaMemset(m->state, 0, (size_t)NSLOT_F32 * mc * HP * 4));
    // ---- tensor maps
    rc = make_tmap_bf16_2d(&m->tm_img, m->img_bf, (uint64_t)(mc + 128), (uint64_t)RE, (uint64_t)RE * 2, 128);
    if (rc) return rc;
    rc = make_tmap_bf16_2d(&m->tm_img3, m->img_bf, (uint64_t)((mc + 128) / S), (uint64_t)S * RE, (uint64_t)S * RE * 2, 128);
    if (rc) return rc;
    rc = make_tmap_bf16_2d(&m->tm_arena, m->arena, (uint64_t)NSLOT_BF * mc, HP, HP * 2, 128);
    if (rc) return rc;
    if (m->alloc_lo) rc = make_tmap_bf16_2d(&m->tm_arena_lo, m->arena_lo, (uint64_t)NSLOT_BF * mc, HP, HP * 2, 128);
    else m->tm_arena_lo = m->tm_arena;
    if (rc) return rc;
    rc = make_tmap_bf16_2d(&m->tm_arena_mc, m->arena, (uint64_t)NSLOT_BF * mc, HP, HP * 2, 128 / m->gru_np);
    if (rc) return rc;
    if (m->alloc_lo) rc = make_tmap_bf16_2d(&m->tm_arena_lo_mc, m->arena_lo, (uint64_t)NSLOT_BF * mc, HP, HP * 2, 128 / m->gru_np);
    else m->tm_arena_lo_mc = m->tm_arena_mc;
    if (rc) return rc;
    rc = make_tmap_bf16_2d(&m->tm_wfold, m->w_fold, (uint64_t)S * HP, (uint64_t)m->kcols_fold, (uint64_t)m->kcols_fold * 2, 256);
    if (rc) return rc;
    rc = make_tmap_bf16_2d(&m->tm_wgru, m->w_gru, (uint64_t)m->n_layers * 2 * 3 * HP, (uint64_t)m->kcols_gru,
                           (uint64_t)m->kcols_gru * 2, 192);
    if (rc) return rc;
    rc = make_tmap_bf16_2d(&m->tm_wgru96, m->w_gru, (uint64_t)m->n_layers * 2 * 3 * HP, (uint64_t)m->kcols_gru,
                           (uint64_t)m->kcols_gru * 2, 96);
    if (rc) return rc;
    const char *e3 = getenv("DP_GRU_STATE");            // "f32": keep the fp32 state copy also in one-pass bf16 mode
    m->bf_state = 0;
    if (m->gru_2cta && !m->alloc_lo && e3) {
        if (!strcmp(e3, "bf16")) m->bf_state = m->n_layers;
        else if (!strcmp(e3, "mixed")) m->bf_state = m->n_layers - 1;
    }
    return rc;
}

// triples == 0: img is a stack of n_rows image rows and call r reads rows r .. r+seq_len-1 (every sliding window);
// triples == 1: img is n_rows independent windows of seq_len rows each (the reference's Model.forward call)
// win_h > 0 (sliding mode only): the stack is n_rows / win_h pileups of win_h rows; the calls at the last seq_len-1 rows of a
// pileup -- whose rows would run into the NEXT pileup, and which test.py never makes (its row loop stops at H - seq_len,
// src/test.py:59) -- are dropped behind the folded GEMM: the GRU steps and fc1 run on win_h-seq_len+1 calls per pileup and
// fc1 writes each call back at its own row; the skipped rows of logits / pred are left untouched
static int forward_rows(Model *m, const uint8_t *img, long long n_rows, float *logits, uint8_t *pred, cudaStream_t st,
                        int triples = 0, int win_h = 0) {
    const int RE = m->row_elems, S = m->seq_len;
    const int ckeep = win_h > 0 ? win_h - (S - 1) : 0, cgap = win_h > 0 ? S - 1 : 0;
    const long long n_calls = win_h > 0 ? n_rows / win_h * ckeep : n_rows;       // rows the GRU and fc1 work on
    const long long mc = m->mcap;
    size_t prof_base = 0;
    int prof_i = 0;
    if (m->profile) {
        prof_base = m->prof_events.size();
        for (int i = 0; i < PROF_MARKS; i++) {
            cudaEvent_t e;
            DP_CUDA_CHECK(cudaEventCreate(&e));
            m->prof_events.push_back(e);
        }
    }
    auto mark = [&]() {
        if (m->profile && prof_i < PROF_MARKS) cudaEventRecord(m->prof_events[prof_base + prof_i++], st);
    };
    mark();
    // 1. u8 -> bf16 (plus seq_len-1 zero rows after the last one so trailing windows see zeros, not stale data)
    {
        const long long n_img_rows = triples ? n_rows * S : n_rows;
        const long long n16 = n_img_rows * RE / 16;
        u8_to_bf16_kernel<<<(unsigned)((n16 + 255) / 256), 256, 0, st>>>((const uint4 *)img, (uint4 *)m->img_bf, n16, m->f16);
        DP_CUDA_CHECK(cudaGetLastError());
        if (!triples) DP_CUDA_CHECK(cudaMemsetAsync(m->img_bf + n_rows * RE, 0, (size_t)(S - 1) * RE * 2, st));
    }
    mark();
    const int tiles_m = (int)((n_rows + 127) / 128);
    const int tiles_mc = (int)((n_calls + 127) / 128);
    const CUtensorMap tmA[4] = {m->tm_img, m->tm_arena, m->tm_arena_lo, m->tm_img3};
    const CUtensorMap tmA_mc[4] = {m->tm_img, m->tm_arena_mc, m->tm_arena_lo_mc, m->tm_arena_mc};   // GRU pair kernel
    // 2. folded CNN + Linear
    {
        GemmArgs g;
        memset(&g, 0, sizeof(g));
        g.chunks = m->tables + (triples ? m->fold3_tab_off : m->fold_tab_off);
        g.n_chunks = m->fold_tab_n;
        g.M = (int)n_rows; g.N = S * HP;
        g.tiles_m = tiles_m; g.tiles_n = (S * HP) / 256;
        g.f16 = m->f16;
        g.epi.acc_scale = 1.f / m->fold_scale;
        g.epi.bias = m->b_fold;
        g.epi.out = m->arena;
        g.epi.out_lo = m->alloc_lo ? m->arena_lo : nullptr;
        g.epi.ldc = HP; g.epi.out_fp32 = 0;
        g.epi.nblk = HP; g.epi.nblk_stride = mc * HP;
        g.epi.compact_h = win_h; g.epi.compact_keep = ckeep;
        int rc = launch_gemm_bf16(tmA, &m->tm_wfold, g, 256, st);
        if (rc) return rc;
    }
    mark();
    // 3. GRU layers
    for (int k = 0; k < m->n_layers; k++)
        for (int sg = 0; sg < 3; sg++) {
            GruArgs g;
            memset(&g, 0, sizeof(g));
            g.n_prob = 2; g.M = (int)n_calls; g.tiles_m = tiles_mc; g.fast_act = m->fast_act;
            g.f16 = m->f16; g.acc_scale = 1.f / m->gru_scale;
            for (int d = 0; d < 2; d++) {
                GruProblem &P = g.prob[d];
                const int t = d == 0 ? sg : 2 - sg;
                P.chunks = m->tables + m->gru_tab_off[k][sg][d];
                P.n_chunks = m->gru_tab_n[k][sg][d];
                P.b_row_base = (k * 2 + d) * 3 * HP;
                P.bias = m->b_gru + (size_t)(k * 2 + d) * 4 * HP;
                int hs = -1;
                if (sg > 0) hs = d == 0 ? f32_slot_hf(k, t - 1) : f32_slot_hb(k, t + 1);
                else if (k > 0) hs = d == 0 ? f32_slot_hf(k - 1, 2) : f32_slot_hb(k - 1, 0);
                int hsb = -1;                            // the same state in the bf16 arena
                if (sg > 0) hsb = d == 0 ? bf_slot_hf(k, t - 1) : bf_slot_hb(k, t + 1);
                else if (k > 0) hsb = d == 0 ? bf_slot_hf(k - 1, 2) : bf_slot_hb(k - 1, 0);
                const int os = d == 0 ? f32_slot_hf(k, t) : f32_slot_hb(k, t);
                const int ob = d == 0 ? bf_slot_hf(k, t) : bf_slot_hb(k, t);
                // layer k reads h0 from layer k-1 (sg == 0): fp32 only if that layer kept one
                const bool prev_f32 = sg > 0 ? k >= m->bf_state : (k - 1) >= m->bf_state;
                if (hs >= 0 && prev_f32) P.h_prev = m->state + (size_t)hs * mc * HP;
                else if (hsb >= 0) P.h_prev_bf = m->arena + (size_t)hsb * mc * HP;
                P.h_out = k >= m->bf_state ? m->state + (size_t)os * mc * HP : nullptr;
                // the bf16 copy is the next step's / next layer's A operand: nobody reads the last one of the last layer
                // (fc1 takes the fp32 state)
                const bool bf_unused = sg == 2 && k + 1 == m->n_layers && k >= m->bf_state;
                P.h_out_bf = bf_unused ? nullptr : m->arena + (size_t)ob * mc * HP;
                P.h_out_lo = (m->alloc_lo && !bf_unused) ? m->arena_lo + (size_t)ob * mc * HP : nullptr;
                if (sg == 2 && k + 1 < m->n_layers) {
                    // sum of directions (model.py:131-133) -> bf16 input of the next layer
                    if (d == 0) {
                        if (k < m->bf_state) {
                            P.partner_bf = m->arena + (size_t)bf_slot_hb(k, 2) * mc * HP;
                            P.partner2_bf = m->arena + (size_t)bf_slot_hb(k, 1) * mc * HP;
                        } else {
                            P.partner = m->state + (size_t)f32_slot_hb(k, 2) * mc * HP;
                            P.partner2 = m->state + (size_t)f32_slot_hb(k, 1) * mc * HP;
                        }
                        P.xsum = m->arena + (size_t)bf_slot_xn(k, 2) * mc * HP;
                        P.xsum2 = m->arena + (size_t)bf_slot_xn(k, 1) * mc * HP;
                        if (m->alloc_lo) {
                            P.xsum_lo = m->arena_lo + (size_t)bf_slot_xn(k, 2) * mc * HP;
                            P.xsum2_lo = m->arena_lo + (size_t)bf_slot_xn(k, 1) * mc * HP;
                        }
                    } else {
                        if (k < m->bf_state) P.partner_bf = m->arena + (size_t)bf_slot_hf(k, 0) * mc * HP;
                        else P.partner = m->state + (size_t)f32_slot_hf(k, 0) * mc * HP;
                        P.xsum = m->arena + (size_t)bf_slot_xn(k, 0) * mc * HP;
                        if (m->alloc_lo) P.xsum_lo = m->arena_lo + (size_t)bf_slot_xn(k, 0) * mc * HP;
                    }
                }
            }
            int rc = m->gru_wide == 2 ? launch_gru_step2s(tmA, &m->tm_wgru96, g, st)
                   : m->gru_wide ? launch_gru_step2w(tmA, &m->tm_wgru96, g, st)
                   : m->gru_2cta ? launch_gru_step2(tmA_mc, &m->tm_wgru96, g, m->gru_np, st)
                                 : launch_gru_step(tmA, &m->tm_wgru, g, st);
            if (rc) return rc;
            mark();
        }
    // 4. fc1 + argmax on the fp32 states of the last layer
    if (logits || pred) {
        const int k = m->n_layers - 1;
        const long long warps = n_calls * S;
        const unsigned blocks = (unsigned)((warps * 32 + 255) / 256);
        if (k < m->bf_state)
            fc1_argmax_bf_kernel<<<blocks, 256, 0, st>>>(m->arena + (size_t)bf_slot_hf(k, 0) * mc * HP,
                                                         m->arena + (size_t)bf_slot_hb(k, 0) * mc * HP, mc * HP, m->w_fc1,
                                                         m->b_fc1, m->n_class, S, n_calls, logits, pred, m->f16, ckeep, cgap);
        else
            fc1_argmax_kernel<<<(unsigned)((((n_calls + 31) / 32) * S + 3) / 4), 128, (size_t)m->n_class * HP * 4, st>>>(
                                                      m->state + (size_t)f32_slot_hf(k, 0) * mc * HP,
                                                      m->state + (size_t)f32_slot_hb(k, 0) * mc * HP, mc * HP, m->w_fc1,
                                                      m->b_fc1, m->n_class, S, n_calls, logits, pred, ckeep, cgap);
        DP_CUDA_CHECK(cudaGetLastError());
    }
    while (m->profile && prof_i < PROF_MARKS) mark();     // unused marks collapse onto the end
    return DP_OK;
}

}  // namespace dp

// ================================================================================================
struct dp_model {
    dp::Model *m;
};

extern "C" {

int dp_model_create(const dp_model_weights *w, int64_t max_rows, int32_t precision, dp_model **out) {
    DP_REQUIRE(w && out, "null argument");
    DP_REQUIRE(w->fold_w && w->fold_b && w->fc1_w && w->fc1_b, "null weight pointer");
    DP_REQUIRE(w->seq_len == 3, "seq_len must be 3 (test.py:48)");
    DP_REQUIRE(w->hidden >= 1 && w->hidden <= dp::HP, "hidden size must be <= 512");
    DP_REQUIRE(w->n_layers >= 1 && w->n_layers <= DP_MAX_GRU_LAYERS, "1..3 GRU layers");
    DP_REQUIRE(w->n_class >= 1 && w->n_class <= 16, "n_class must be in [1,16]");
    DP_REQUIRE(w->in_ch >= 1 && w->width >= 1 && (w->in_ch * w->width) % 16 == 0, "width*in_ch must be a multiple of 16");
    DP_REQUIRE(max_rows >= 1 && max_rows <= (1ll << 24), "max_rows out of range");
    DP_REQUIRE(precision >= 0 && precision <= 3, "precision must be 0 (bf16), 1 (split bf16), 2 (fp16) or 3 (split fp16)");
    for (int k = 0; k < w->n_layers; k++)
        for (int d = 0; d < 2; d++)
            DP_REQUIRE(w->gru_w_ih[k][d] && w->gru_w_hh[k][d] && w->gru_b_ih[k][d] && w->gru_b_hh[k][d], "null GRU weight");
    dp::Model *m = new dp::Model();
    m->in_ch = w->in_ch; m->width = w->width; m->hidden = w->hidden; m->n_class = w->n_class;
    m->seq_len = w->seq_len; m->n_layers = w->n_layers;
    m->row_elems = w->in_ch * w->width;
    m->kfold_pad = (m->row_elems + 63) / 64 * 64;
    // precision presets; the DP_FOLD_SPLIT / DP_GRU_SPLIT / DP_ACT switches exist for scripts/model_error_attribution.py,
    // which turns the accurate variant of one error source on at a time
    m->f16 = precision >> 1;
    m->fold_split = precision & 1;
    m->gru_split = (precision & 1) ? (1 << w->n_layers) - 1 : 0;
    m->fast_act = precision == 0 ? 1 : 0;           // MUFU.TANH only in the round-1 bf16 mode: its 2^-11 error is at the level of
                                                    // fp16's operand rounding (measured: fp16 logits 2.90e-3 -> 2.81e-3 max-abs)
    if (const char *e = getenv("DP_FOLD_SPLIT")) m->fold_split = atoi(e) ? 1 : 0;
    if (const char *e = getenv("DP_GRU_SPLIT")) m->gru_split = atoi(e) & ((1 << w->n_layers) - 1);
    if (const char *e = getenv("DP_ACT")) m->fast_act = !strcmp(e, "fast") ? 1 : !strcmp(e, "accurate") ? 0 : m->fast_act;
    m->alloc_lo = m->gru_split != 0;
    m->mcap = (max_rows + 127) / 128 * 128;
    // the arena row coordinate (slot * mcap + row) must fit the int32 TMA coordinate
    if ((long long)dp::NSLOT_BF * m->mcap >= (1ll << 31)) { delete m; dp::set_error("max_rows too large"); return DP_EINVAL; }
    if (cudaGetDevice(&m->device) != cudaSuccess) { delete m; dp::set_error("no CUDA device"); return DP_ECUDA; }
    const int rc = dp::build_model(m, w);
    if (rc) { dp::free_model(m); return rc; }
    if (cudaDeviceSynchronize() != cudaSuccess) { dp::free_model(m); dp::set_error("model upload failed"); return DP_ECUDA; }
    dp_model *h = new dp_model();
    h->m = m;
    *out = h;
    return DP_OK;
}

int dp_model_destroy(dp_model *h) {
    if (!h) return DP_OK;
    dp::free_model(h->m);
    delete h;
    return DP_OK;
}

int dp_model_forward_rows(dp_model *h, const uint8_t *img, int64_t n_rows, float *logits, uint8_t *pred, void *stream) {
    DP_REQUIRE(h && h->m, "null model");
    DP_REQUIRE(n_rows >= 0 && n_rows <= h->m->mcap, "n_rows exceeds the model's max_rows");
    if (n_rows == 0) return DP_OK;
    DP_REQUIRE(img, "null image");
    DP_REQUIRE(((uintptr_t)img & 15) == 0, "img must be 16-byte aligned");
    return dp::forward_rows(h->m, img, n_rows, logits, pred, (cudaStream_t)stream);
}

int dp_model_forward_windows(dp_model *h, const uint8_t *img, int64_t n_windows, int32_t H, float *logits, uint8_t *pred,
                             void *stream) {
    DP_REQUIRE(h && h->m, "null model");
    DP_REQUIRE(n_windows >= 0 && H >= h->m->seq_len, "H must be at least seq_len");
    DP_REQUIRE(n_windows * (int64_t)H <= h->m->mcap, "n_windows * H exceeds the model's max_rows");
    if (n_windows == 0) return DP_OK;
    DP_REQUIRE(img, "null image");
    DP_REQUIRE(((uintptr_t)img & 15) == 0, "img must be 16-byte aligned");
    return dp::forward_rows(h->m, img, n_windows * (int64_t)H, logits, pred, (cudaStream_t)stream, 0, H);
}

int dp_model_forward_calls(dp_model *h, const uint8_t *img, int64_t n_calls, float *logits, uint8_t *pred, void *stream) {
    DP_REQUIRE(h && h->m, "null model");
    DP_REQUIRE(n_calls >= 0 && n_calls * h->m->seq_len <= h->m->mcap, "n_calls * seq_len exceeds the model's max_rows");
    if (n_calls == 0) return DP_OK;
    DP_REQUIRE(img, "null image");
    DP_REQUIRE(((uintptr_t)img & 15) == 0, "img must be 16-byte aligned");
    return dp::forward_rows(h->m, img, n_calls, logits, pred, (cudaStream_t)stream, 1);
}

int dp_vote(const uint8_t *pred, int32_t n_windows, int32_t H, int32_t n_class, int8_t *calls, const uint8_t *labels,
            unsigned long long *confusion, void *stream) {
    DP_REQUIRE(n_windows >= 0 && H >= 0, "negative extent");
    if (n_windows == 0 || H == 0) return DP_OK;
    DP_REQUIRE(pred && calls, "null argument");
    const long long n = (long long)n_windows * H;
    dp::vote_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pred, n_windows, H, calls, labels,
                                                                                  confusion, n_class);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

int dp_cross_entropy(const float *logits, const uint8_t *labels, int64_t n, int32_t n_class, float *loss, double *sum,
                     int32_t *bad_labels, void *stream) {
    DP_REQUIRE(n >= 0 && n_class >= 1 && n_class <= 64, "bad extent");
    if (n == 0) return DP_OK;
    DP_REQUIRE(logits && labels && (loss || sum), "null argument");
    dp::cross_entropy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(logits, labels, n, n_class, loss, sum,
                                                                                           bad_labels);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

int dp_model_peek(dp_model *h, int32_t what, int64_t n_rows, float *host_out) {
    DP_REQUIRE(h && h->m && host_out, "null argument");
    dp::Model *m = h->m;
    DP_REQUIRE(n_rows >= 0 && n_rows <= m->mcap, "n_rows out of range");
    DP_CUDA_CHECK(cudaDeviceSynchronize());
    const size_t n = (size_t)n_rows * dp::HP;
    if (what >= 0 && what < 3) {
        std::vector<uint16_t> tmp(n);
        DP_CUDA_CHECK(cudaMemcpy(tmp.data(), m->arena + (size_t)dp::bf_slot_y(what) * m->mcap * dp::HP, n * 2, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; i++) host_out[i] = dp::h162f(tmp[i], m->f16 != 0);
        if (m->alloc_lo) {
            DP_CUDA_CHECK(cudaMemcpy(tmp.data(), m->arena_lo + (size_t)dp::bf_slot_y(what) * m->mcap * dp::HP, n * 2, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < n; i++) host_out[i] += dp::h162f(tmp[i], m->f16 != 0);
        }
        return DP_OK;
    }
    const int k = what / 100, d = (what / 10) % 10, t = what % 10;
    DP_REQUIRE(what >= 100 && k >= 1 && k <= m->n_layers && d < 2 && t < 3, "unknown peek id");
    if (k - 1 < m->bf_state) {
        std::vector<uint16_t> tmp(n);
        const int bslot = d == 0 ? dp::bf_slot_hf(k - 1, t) : dp::bf_slot_hb(k - 1, t);
        DP_CUDA_CHECK(cudaMemcpy(tmp.data(), m->arena + (size_t)bslot * m->mcap * dp::HP, n * 2, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; i++) host_out[i] = dp::h162f(tmp[i], m->f16 != 0);
        return DP_OK;
    }
    const int slot = d == 0 ? dp::f32_slot_hf(k - 1, t) : dp::f32_slot_hb(k - 1, t);
    const size_t nb = (size_t)((n_rows + 127) / 128 * 128) * dp::HP;
    std::vector<float> tmp(nb);
    DP_CUDA_CHECK(cudaMemcpy(tmp.data(), m->state + (size_t)slot * m->mcap * dp::HP, nb * 4, cudaMemcpyDeviceToHost));
    for (long long r = 0; r < n_rows; r++)
        for (int u = 0; u < dp::HP; u++) {
            const long long mb = r >> 7;
            const int rr = (int)(r & 127), ub = u >> 6, cb = (u >> 4) & 3, k = (u >> 2) & 3, e = u & 3;
            host_out[(size_t)r * dp::HP + u] = tmp[(size_t)(((((mb * 8 + ub) * 4 + cb) * 4 + k) * 128 + rr) * 4 + e)];
        }
    return DP_OK;
}

int dp_model_profile(dp_model *h, int32_t enable) {
    DP_REQUIRE(h && h->m, "null model");
    h->m->profile = enable ? 1 : 0;
    return DP_OK;
}

int dp_model_profile_read(dp_model *h, float ms[4], int32_t launches[4]) {
    DP_REQUIRE(h && h->m && ms && launches, "null argument");
    dp::Model *m = h->m;
    for (int i = 0; i < 4; i++) { ms[i] = 0.f; launches[i] = 0; }
    DP_CUDA_CHECK(cudaDeviceSynchronize());
    const int n_gru = m->n_layers * 3;
    for (size_t b = 0; b + dp::PROF_MARKS <= m->prof_events.size(); b += dp::PROF_MARKS) {
        for (int i = 0; i + 1 < dp::PROF_MARKS; i++) {
            float t = 0.f;
            if (cudaEventElapsedTime(&t, m->prof_events[b + i], m->prof_events[b + i + 1]) != cudaSuccess) { cudaGetLastError(); continue; }
            const int cat = i == 0 ? 0 : i == 1 ? 1 : i < 2 + n_gru ? 2 : i == 2 + n_gru ? 3 : -1;
            if (cat < 0) continue;
            ms[cat] += t;
            launches[cat] += 1;
        }
    }
    for (cudaEvent_t e : m->prof_events) cudaEventDestroy(e);
    m->prof_events.clear();
    return DP_OK;
}

int dp_gemm_bf16(const void *A, int64_t lda, const void *B, int64_t ldb, const float *bias, void *out, int64_t ldc,
                 int32_t out_fp32, int64_t M, int64_t N, int64_t K, int32_t block_n, void *stream) {
    DP_REQUIRE(A && B && out, "null argument");
    DP_REQUIRE(M >= 1 && N >= 16 && N % 16 == 0 && K >= 1, "need M >= 1, N a multiple of 16, K >= 1");
    DP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "leading dimensions must be multiples of 8 elements");
    DP_REQUIRE(block_n == 128 || block_n == 256, "block_n must be 128 or 256");
    CUtensorMap tmA[4], tmB;
    int rc = dp::make_tmap_bf16_2d(&tmA[0], A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 2, 128);
    if (rc) return rc;
    tmA[1] = tmA[2] = tmA[3] = tmA[0];
    rc = dp::make_tmap_bf16_2d(&tmB, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, (uint32_t)block_n);
    if (rc) return rc;
    const int nch = (int)((K + 63) / 64);
    std::vector<dp::KChunk> tab(nch);
    for (int i = 0; i < nch; i++) tab[i] = dp::KChunk{0, 64 * i, 0, 64 * i};
    dp::KChunk *dtab = nullptr;
    DP_CUDA_CHECK(cudaMalloc((void **)&dtab, sizeof(dp::KChunk) * nch));
    DP_CUDA_CHECK(cudaMemcpyAsync(dtab, tab.data(), sizeof(dp::KChunk) * nch, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    DP_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));      // tab is pageable host memory
    dp::GemmArgs g;
    memset(&g, 0, sizeof(g));
    g.chunks = dtab; g.n_chunks = nch; g.M = (int)M; g.N = (int)N;
    g.tiles_m = (int)((M + 127) / 128); g.tiles_n = (int)((N + block_n - 1) / block_n);
    g.epi.bias = bias; g.epi.out = out; g.epi.ldc = ldc; g.epi.out_fp32 = out_fp32;
    rc = dp::launch_gemm_bf16(tmA, &tmB, g, block_n, (cudaStream_t)stream);
    cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
    cudaFree(dtab);
    if (rc) return rc;
    if (e != cudaSuccess) { dp::set_error("dp_gemm_bf16: %s", cudaGetErrorString(e)); return DP_ECUDA; }
    return DP_OK;
}

}  // extern "C"
