// Tensor-core kernels of the deePore model forward -- hand-written tcgen05 / TMEM / TMA for sm_100a.
// See tc_gemm.cuh for what each kernel computes and which reference code it replaces.
//
// CTA anatomy (both kernels, 256 threads, one CTA per SM, persistent over a static tile list):
//   warp 0    TMA producer: one elected lane walks the K-chunk table and fills a ring of smem stages
//   warp 1    MMA issuer:   one elected lane issues tcgen05.mma per stage, commits stage -> empty barrier,
//                           commits the finished accumulator -> tmem_full barrier
//   warp 2    TMEM allocator (512 columns = two accumulator stages)
//   warps 4-7 epilogue:     tcgen05.ld their 32 accumulator rows, apply the epilogue, store to HBM, release
//                           the accumulator stage
#include <stdlib.h>

#include "tc_gemm.cuh"

namespace dp {

using namespace tc;

// ------------------------------------------------------------------------------------------------
// host: tensor maps through the runtime's driver entry point (no link-time dependency on libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

int make_tmap_bf16_2d(CUtensorMap *out, const void *base, uint64_t rows, uint64_t cols, uint64_t pitch_bytes,
                      uint32_t box_rows) {
    if (!g_encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        DP_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        if (!fn || qres != cudaDriverEntryPointSuccess) {
            set_error("cuTensorMapEncodeTiled is not available from this driver");
            return DP_ECUDA;
        }
        g_encode = (EncodeTiledFn)fn;
    }
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstr[1] = {pitch_bytes};
    cuuint32_t box[2] = {64, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), gdim, gstr, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d): rows=%llu cols=%llu pitch=%llu box_rows=%u", (int)r,
                  (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)pitch_bytes, box_rows);
        return DP_ECUDA;
    }
    return DP_OK;
}

int device_sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev < 64 && cached[dev]) return cached[dev];
    int n = 148;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (dev < 64) cached[dev] = n;
    return n;
}

// ------------------------------------------------------------------------------------------------
constexpr int TC_THREADS = 256;
constexpr int BM = 128;            // accumulator rows per tile = TMEM lanes
constexpr int BK = 64;             // bf16 elements per K chunk = one 128-byte swizzle row
constexpr int A_STAGE_BYTES = BM * BK * 2;   // 16 KB

struct PipeState {
    int stage;
    uint32_t phase;
    __device__ __forceinline__ void advance(int nstages) {
        if (++stage == nstages) { stage = 0; phase ^= 1u; }
    }
};


template <int BN>
struct GemmCfg {
    static constexpr int B_STAGE_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
    static constexpr int STAGES = (BN >= 256) ? 4 : 6;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

// ================================================================================================
// generic table-driven GEMM
template <int BN>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                 const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmA3,
                 const __grid_constant__ CUtensorMap tmB, const GemmArgs g) {
    using Cfg = GemmCfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + STAGES * Cfg::STAGE_BYTES;
    // barriers: full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2]; then the TMEM base address word
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    uint32_t *tmem_slot_ptr = reinterpret_cast<uint32_t *>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles = g.tiles_m * g.tiles_n;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA0); prefetch_tmap(&tmA1); prefetch_tmap(&tmA2); prefetch_tmap(&tmA3); prefetch_tmap(&tmB);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int s = 0; s < 2; s++) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 4); }
        fence_barrier_init();
    }
    if (warp == 2) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        if (lane == 0) {
            PipeState ps{0, 0};
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const int m0 = (tile / g.tiles_n) * BM, n0 = (tile % g.tiles_n) * BN;
                for (int kc = 0; kc < g.n_chunks; kc++) {
                    const KChunk c = g.chunks[kc];
                    const int src = c.a_src & 0xff;
                    const CUtensorMap *ma = src == 0 ? &tmA0 : src == 1 ? &tmA1 : src == 2 ? &tmA2 : &tmA3;
                    mbar_wait(empty_bar(ps.stage), ps.phase ^ 1u);
                    mbar_expect_tx(full_bar(ps.stage), Cfg::STAGE_BYTES);
                    const uint32_t sa = smem_base + ps.stage * Cfg::STAGE_BYTES;
                    tma_load_2d(sa, ma, full_bar(ps.stage), c.a_col, m0 + c.a_drow);
                    tma_load_2d(sa + A_STAGE_BYTES, &tmB, full_bar(ps.stage), c.b_col, n0);
                    ps.advance(STAGES);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc_16(BM, BN, g.f16 != 0);
            PipeState ps{0, 0};
            int as = 0; uint32_t aphase = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                mbar_wait(tempty_bar(as), aphase ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int kc = 0; kc < g.n_chunks; kc++) {
                    mbar_wait(full_bar(ps.stage), ps.phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + ps.stage * Cfg::STAGE_BYTES;
                    const uint64_t da = make_smem_desc_sw128(sa), db = make_smem_desc_sw128(sa + A_STAGE_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; k++)
                        umma_bf16(d_tmem, da + 2u * k, db + 2u * k, idesc, (kc | k) != 0);
                    umma_commit(empty_bar(ps.stage));
                    ps.advance(STAGES);
                }
                umma_commit(tfull_bar(as));
                if (++as == 2) { as = 0; aphase ^= 1u; }
            }
        }
    } else if (warp >= 4) {
        const int wq = warp - 4;
        const GemmEpilogue &e = g.epi;
        const bool f16 = g.f16 != 0;
        const float asc = e.acc_scale != 0.f ? e.acc_scale : 1.f;
        int as = 0; uint32_t aphase = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int m0 = (tile / g.tiles_n) * BM, n0 = (tile % g.tiles_n) * BN;
            const int m = m0 + wq * 32 + lane;
            mbar_wait(tfull_bar(as), aphase);
            tc_fence_after();
            bool row_ok = m < g.M;
            long long m_out = m;
            if (e.compact_h > 0) {
                const int w = m / e.compact_h, r = m - w * e.compact_h;
                row_ok = row_ok && r < e.compact_keep;
                m_out = (long long)m - (long long)(e.compact_h - e.compact_keep) * w;
            }
            bool row_zero = false;
            if (e.mask_period > 0) { const int r = m % e.mask_period; row_zero = r < e.mask_lo || r >= e.mask_hi; }
            const uint32_t taddr = tmem_base + as * BN + ((uint32_t)(wq * 32) << 16);
#pragma unroll 1
            for (int cb = 0; cb < BN / 16; cb++) {
                const int n = n0 + cb * 16;
                if (n >= g.N) break;                       // warp-uniform
                uint32_t v[16];
                tmem_ld16(taddr + cb * 16, v);
                tmem_ld_wait();
                if (!row_ok) continue;
                float f[16];
#pragma unroll
                for (int i = 0; i < 16; i++) f[i] = __uint_as_float(v[i]) * asc;
                if (e.bias) {
#pragma unroll
                    for (int i = 0; i < 16; i += 4) {
                        const float4 b = __ldg(reinterpret_cast<const float4 *>(e.bias + n + i));
                        f[i] += b.x; f[i + 1] += b.y; f[i + 2] += b.z; f[i + 3] += b.w;
                    }
                }
                if (e.residual) {
                    const uint4 *rp = reinterpret_cast<const uint4 *>(e.residual + (long long)m * e.res_ld + n);
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const uint4 r = __ldg(rp + h);
                        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            float lo, hi;
                            unpack16(w[i], lo, hi, f16);
                            f[h * 8 + 2 * i] += lo;
                            f[h * 8 + 2 * i + 1] += hi;
                        }
                    }
                }
                if (row_zero) {
#pragma unroll
                    for (int i = 0; i < 16; i++) f[i] = 0.f;
                }
                long long off = m_out * e.ldc;
                if (e.nblk > 0) off += (long long)(n / e.nblk) * e.nblk_stride + (n % e.nblk); else off += n;
                if (e.out_fp32) {
                    float4 *op = reinterpret_cast<float4 *>(reinterpret_cast<float *>(e.out) + off);
#pragma unroll
                    for (int i = 0; i < 4; i++) op[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
                } else {
                    uint32_t p[8];
#pragma unroll
                    for (int i = 0; i < 8; i++) p[i] = pack16(f[2 * i], f[2 * i + 1], f16);
                    uint4 *op = reinterpret_cast<uint4 *>(reinterpret_cast<__nv_bfloat16 *>(e.out) + off);
                    op[0] = make_uint4(p[0], p[1], p[2], p[3]);
                    op[1] = make_uint4(p[4], p[5], p[6], p[7]);
                    if (e.out_lo) {
#pragma unroll
                        for (int i = 0; i < 8; i++)
                            p[i] = pack16(f[2 * i] - round16(f[2 * i], f16), f[2 * i + 1] - round16(f[2 * i + 1], f16), f16);
                        uint4 *lp = reinterpret_cast<uint4 *>(e.out_lo + off);
                        lp[0] = make_uint4(p[0], p[1], p[2], p[3]);
                        lp[1] = make_uint4(p[4], p[5], p[6], p[7]);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(as));
            if (++as == 2) { as = 0; aphase ^= 1u; }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

template <int BN>
static int launch_gemm_t(const CUtensorMap *tmA, const CUtensorMap *tmB, const GemmArgs &g, cudaStream_t st) {
    using Cfg = GemmCfg<BN>;
    static bool attr_done[64] = {false};
    int dev = 0;
    DP_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_done[dev]) {
        DP_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        if (dev < 64) attr_done[dev] = true;
    }
    const int n_tiles = g.tiles_m * g.tiles_n;
    if (n_tiles <= 0) return DP_OK;
    const int grid = n_tiles < device_sm_count() ? n_tiles : device_sm_count();
    gemm_bf16_kernel<BN><<<grid, TC_THREADS, Cfg::SMEM_BYTES, st>>>(tmA[0], tmA[1], tmA[2], tmA[3], *tmB, g);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

int launch_gemm_bf16(const CUtensorMap *tmA, const CUtensorMap *tmB, const GemmArgs &g, int block_n, cudaStream_t st) {
    if (block_n == 256) return launch_gemm_t<256>(tmA, tmB, g, st);
    if (block_n == 128) return launch_gemm_t<128>(tmA, tmB, g, st);
    set_error("launch_gemm_bf16: block_n must be 128 or 256");
    return DP_EINVAL;
}

// ================================================================================================
// GRU time step (both directions).  Tile = (problem, 128 rows, block of 64 hidden units).  Weight tile rows are
// gate-major [r 64 | z 64 | n 64] = 192 rows; accumulator columns per stage: [0,64) r, [64,128) z, [128,192) W_in x,
// [192,256) W_hn h.  Input chunks issue one N=192 MMA; hidden chunks an N=128 MMA onto r,z plus an N=64 MMA onto
// the separate W_hn h columns (torch.nn.GRU: n = tanh(W_in x + b_in + r * (W_hn h + b_hn))).
constexpr int GRU_UB = 64;                          // hidden units per tile
constexpr int GRU_BROWS = 3 * GRU_UB;               // 192 weight rows per tile
constexpr int GRU_B_STAGE_BYTES = GRU_BROWS * BK * 2;   // 24 KB
constexpr int GRU_STAGE_BYTES = A_STAGE_BYTES + GRU_B_STAGE_BYTES;   // 40 KB
constexpr int GRU_STAGES = 5;
constexpr int GRU_SMEM_BYTES = GRU_STAGES * GRU_STAGE_BYTES + 1024 + 256;
constexpr int GRU_HP = 512;                         // padded hidden size (row pitch of every state buffer)
constexpr int GRU_NUB = GRU_HP / GRU_UB;            // 8 unit blocks

__device__ __forceinline__ float sigmoid_acc(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float tanh_acc(float x) {
    const float e = __expf(2.f * x);
    return 1.f - __fdividef(2.f, e + 1.f);
}

// MUFU.TANH: ~2^-11 relative error -- below the bf16 operand rounding of the one-pass mode, not used in split mode
__device__ __forceinline__ float tanh_fast(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
template <bool FAST>
__device__ __forceinline__ float act_sigmoid(float x) {
    if (FAST) return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f);
    return sigmoid_acc(x);
}
template <bool FAST>
__device__ __forceinline__ float act_tanh(float x) { return FAST ? tanh_fast(x) : tanh_acc(x); }


// ---- GRU epilogue, shared by the 1-CTA and the CTA-pair kernel ----------------------------------------------------
// fp32 state layout (private to these kernels, fc1_argmax_kernel and dp_model_peek): BLOCKED so that one warp
// instruction moves 512 contiguous bytes.  Row m, unit u -> mb = m/128, r = m%128, ub = u/64, cb = (u/16)%4,
// k = (u/4)%4, e = u%4:   offset = ((((mb*8 + ub)*4 + cb)*4 + k)*128 + r)*4 + e.
// (Thread = accumulator row holds 16 consecutive units = 4 float4; the k-th float4 of 32 consecutive rows is contiguous.)
__host__ __device__ __forceinline__ long long gru_state_off(long long m, int u) {
    const long long mb = m >> 7;
    const int r = (int)(m & 127), ub = u >> 6, cb = (u >> 4) & 3, k = (u >> 2) & 3, e = u & 3;
    return ((((mb * 8 + ub) * 4 + cb) * 4 + k) * 128 + r) * 4 + e;
}
__device__ __forceinline__ void state_load16(const float *base, long long m, int u0, float (&f)[16]) {
    const float4 *p = reinterpret_cast<const float4 *>(base + gru_state_off(m, u0));
#pragma unroll
    for (int k = 0; k < 4; k++) { const float4 t = p[k * 128]; f[4 * k] = t.x; f[4 * k + 1] = t.y; f[4 * k + 2] = t.z; f[4 * k + 3] = t.w; }
}
__device__ __forceinline__ void state_store16(float *base, long long m, int u0, const float (&f)[16]) {
    float4 *p = reinterpret_cast<float4 *>(base + gru_state_off(m, u0));
#pragma unroll
    for (int k = 0; k < 4; k++) p[k * 128] = make_float4(f[4 * k], f[4 * k + 1], f[4 * k + 2], f[4 * k + 3]);
}
// bf16 row-major [row][512] store of 16 units per row for the warp's 32 rows, full 32-byte sectors per instruction:
// lanes 2j / 2j+1 write the two halves of row j's 32 bytes (values fetched from lane j by shuffle).  All 32 lanes call.
__device__ __forceinline__ void rows_store_bf16(__nv_bfloat16 *base, long long m_warp0, int u0, const float (&f)[16], int M,
                                                int lane, bool lo_part, bool f16) {
    uint32_t p[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        float a = f[2 * i], b = f[2 * i + 1];
        if (lo_part) { a -= round16(a, f16); b -= round16(b, f16); }
        p[i] = pack16(a, b, f16);
    }
#pragma unroll
    for (int h2 = 0; h2 < 2; h2++) {
        const int src = h2 * 16 + (lane >> 1);
        uint32_t v[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t a = __shfl_sync(0xffffffffu, p[i], src), b = __shfl_sync(0xffffffffu, p[4 + i], src);
            v[i] = (lane & 1) ? b : a;
        }
        const long long row = m_warp0 + src;
        if (row < M) *reinterpret_cast<uint4 *>(base + row * GRU_HP + u0 + (lane & 1) * 8) = make_uint4(v[0], v[1], v[2], v[3]);
    }
}

// 16 consecutive units of row m from bf16 rows [row][512] (32 bytes, two 16-byte loads)
__device__ __forceinline__ void rows_load_bf16(const __nv_bfloat16 *base, long long m, int u0, float (&f)[16], bool f16) {
    const uint4 *p = reinterpret_cast<const uint4 *>(base + m * GRU_HP + u0);
    const uint4 a = p[0], b = p[1];
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; i++) unpack16(w[i], f[2 * i], f[2 * i + 1], f16);
}

// Gate math + state update + stores for one group of 16 hidden units of the warp's 32 rows.
template <bool FAST_ACT>
__device__ __forceinline__ void gru_finish16(const GruProblem &P, int M, long long m_warp0, int lane, int u0,
                                             const uint32_t (&vr)[16], const uint32_t (&vz)[16], const uint32_t (&vx)[16],
                                             const uint32_t (&vh)[16], bool has_h, const float (&hp)[16], bool f16, float sc,
                                             int dbg = 0) {
    const long long m = m_warp0 + lane;
    const bool row_ok = m < M;
    float hn[16];
    // biases: per 4 units one float4 per gate (same address in every lane -> broadcast load)
    float4 bq[4];
#pragma unroll
    for (int i = 0; i < 16; i++) {
        if ((i & 3) == 0) {
#pragma unroll
            for (int gt = 0; gt < 4; gt++)
                bq[gt] = dbg != 9 ? __ldg(reinterpret_cast<const float4 *>(P.bias + gt * GRU_HP + u0 + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const float br = reinterpret_cast<const float *>(&bq[0])[i & 3], bz = reinterpret_cast<const float *>(&bq[1])[i & 3];
        const float bin = reinterpret_cast<const float *>(&bq[2])[i & 3], bhn = reinterpret_cast<const float *>(&bq[3])[i & 3];
        if (dbg == 8) { hn[i] = __uint_as_float(vr[i]) + __uint_as_float(vz[i]) + __uint_as_float(vx[i]) + br + bz + bin + bhn + hp[i]; continue; }
        const float r = act_sigmoid<FAST_ACT>(fmaf(__uint_as_float(vr[i]), sc, br));
        const float z = act_sigmoid<FAST_ACT>(fmaf(__uint_as_float(vz[i]), sc, bz));
        const float hh = has_h ? fmaf(__uint_as_float(vh[i]), sc, bhn) : bhn;
        const float nn = act_tanh<FAST_ACT>(fmaf(__uint_as_float(vx[i]), sc, bin) + r * hh);
        hn[i] = fmaf(z, hp[i] - nn, nn);                  // (1-z)*n + z*h
    }
    if (dbg == 6) { if (hn[3] == 123.456f) P.h_out[0] = hn[1]; return; }
    if (row_ok && P.h_out) state_store16(P.h_out, m, u0, hn);
    if (dbg == 7) return;
    if (P.h_out_bf) rows_store_bf16(P.h_out_bf, m_warp0, u0, hn, M, lane, false, f16);
    if (P.h_out_lo) rows_store_bf16(P.h_out_lo, m_warp0, u0, hn, M, lane, true, f16);
    if (P.partner || P.partner_bf) {                       // sum of directions (model.py:131-133)
        float pt[16];
        if (row_ok && P.partner) state_load16(P.partner, m, u0, pt);
        else if (row_ok) rows_load_bf16(P.partner_bf, m, u0, pt, f16);
        else {
#pragma unroll
            for (int i = 0; i < 16; i++) pt[i] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 16; i++) pt[i] += hn[i];
        rows_store_bf16(P.xsum, m_warp0, u0, pt, M, lane, false, f16);
        if (P.xsum_lo) rows_store_bf16(P.xsum_lo, m_warp0, u0, pt, M, lane, true, f16);
    }
    if (P.partner2 || P.partner2_bf) {
        float pt[16];
        if (row_ok && P.partner2) state_load16(P.partner2, m, u0, pt);
        else if (row_ok) rows_load_bf16(P.partner2_bf, m, u0, pt, f16);
        else {
#pragma unroll
            for (int i = 0; i < 16; i++) pt[i] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 16; i++) pt[i] += hp[i];
        rows_store_bf16(P.xsum2, m_warp0, u0, pt, M, lane, false, f16);
        if (P.xsum2_lo) rows_store_bf16(P.xsum2_lo, m_warp0, u0, pt, M, lane, true, f16);
    }
}

constexpr int GRU_THREADS = 384;                    // warps 0-3 as in the GEMM, warps 4-11 epilogue (two per TMEM lane quarter)

template <bool FAST_ACT>
__global__ void __launch_bounds__(GRU_THREADS, 1)
gru_step_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmA3,
                const __grid_constant__ CUtensorMap tmB, const GruArgs g) {
    constexpr int STAGES = GRU_STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + STAGES * GRU_STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    uint32_t *tmem_slot_ptr = reinterpret_cast<uint32_t *>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_per_prob = g.tiles_m * GRU_NUB;
    const int n_tiles = tiles_per_prob * g.n_prob;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA0); prefetch_tmap(&tmA1); prefetch_tmap(&tmA2); prefetch_tmap(&tmA3); prefetch_tmap(&tmB);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int s = 0; s < 2; s++) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 8); }
        fence_barrier_init();
    }
    if (warp == 2) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    // tile -> (problem, m block, unit block): unit block fastest so the 8 CTAs sharing an A tile run together
    if (warp == 0) {
        if (lane == 0) {
            PipeState ps{0, 0};
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const int pi = tile / tiles_per_prob, rem = tile % tiles_per_prob;
                const GruProblem &P = g.prob[pi];
                const int m0 = (rem / GRU_NUB) * BM, ub = rem % GRU_NUB;
                const int brow = P.b_row_base + ub * GRU_BROWS;
                for (int kc = 0; kc < P.n_chunks; kc++) {
                    const KChunk c = P.chunks[kc];
                    const int src = c.a_src & 0xff;
                    const CUtensorMap *ma = src == 0 ? &tmA0 : src == 1 ? &tmA1 : src == 2 ? &tmA2 : &tmA3;
                    mbar_wait(empty_bar(ps.stage), ps.phase ^ 1u);
                    mbar_expect_tx(full_bar(ps.stage), GRU_STAGE_BYTES);
                    const uint32_t sa = smem_base + ps.stage * GRU_STAGE_BYTES;
                    tma_load_2d(sa, ma, full_bar(ps.stage), c.a_col, m0 + c.a_drow);
                    tma_load_2d(sa + A_STAGE_BYTES, &tmB, full_bar(ps.stage), c.b_col, brow);
                    ps.advance(STAGES);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const bool f16 = g.f16 != 0;
            const uint32_t idesc192 = make_idesc_16(BM, 192, f16);
            const uint32_t idesc128 = make_idesc_16(BM, 128, f16);
            const uint32_t idesc64 = make_idesc_16(BM, 64, f16);
            PipeState ps{0, 0};
            int as = 0; uint32_t aphase = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const GruProblem &P = g.prob[tile / tiles_per_prob];
                mbar_wait(tempty_bar(as), aphase ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * 256;
                uint32_t acc_rz = 0, acc_nx = 0, acc_nh = 0;      // has this column group been written yet?
                for (int kc = 0; kc < P.n_chunks; kc++) {
                    const int kind = (P.chunks[kc].a_src >> 8) & 0xff;
                    mbar_wait(full_bar(ps.stage), ps.phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + ps.stage * GRU_STAGE_BYTES;
                    const uint64_t da = make_smem_desc_sw128(sa), db = make_smem_desc_sw128(sa + A_STAGE_BYTES);
                    const uint64_t dbn = make_smem_desc_sw128(sa + A_STAGE_BYTES + 128 * 128);   // rows 128..191 (n gate)
                    if (kind == 0) {
                        // r,z and nx columns share one instruction, so they must share the accumulate flag:
                        // input chunks come first in the table (host guarantees it), hence acc_rz == acc_nx here.
#pragma unroll
                        for (int k = 0; k < BK / 16; k++)
                            umma_bf16(d_tmem, da + 2u * k, db + 2u * k, idesc192, acc_nx | (uint32_t)k);
                        acc_rz = 1; acc_nx = 1;
                    } else if (g.debug == 2) {
#pragma unroll
                        for (int k = 0; k < BK / 16; k++)
                            umma_bf16(d_tmem, da + 2u * k, db + 2u * k, idesc192, 1u);
                    } else {
#pragma unroll
                        for (int k = 0; k < BK / 16; k++) {
                            umma_bf16(d_tmem, da + 2u * k, db + 2u * k, idesc128, acc_rz | (uint32_t)k);
                            umma_bf16(d_tmem + 192, da + 2u * k, dbn + 2u * k, idesc64, acc_nh | (uint32_t)k);
                        }
                        acc_rz = 1; acc_nh = 1;
                    }
                    umma_commit(empty_bar(ps.stage));
                    ps.advance(STAGES);
                }
                umma_commit(tfull_bar(as));
                if (++as == 2) { as = 0; aphase ^= 1u; }
            }
        }
    } else if (warp >= 4) {
        // 8 epilogue warps: warp w reads TMEM lanes 32*(w%4).. (hardware lane-quarter rule) and handles unit groups
        // cb = 2*half, 2*half+1 of the tile's 64 hidden units, half = (w-4)/4.
        const int wq = warp & 3, half = (warp - 4) >> 2;
        int as = 0; uint32_t aphase = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int pi = tile / tiles_per_prob, rem = tile % tiles_per_prob;
            const GruProblem &P = g.prob[pi];
            const int m0 = (rem / GRU_NUB) * BM, ub = rem % GRU_NUB;
            const long long m_warp0 = m0 + wq * 32;
            const long long m = m_warp0 + lane;
            bool has_h = false;
            for (int kc = 0; kc < P.n_chunks; kc++) has_h |= ((P.chunks[kc].a_src >> 8) & 0xff) != 0;
            // everything that does not depend on the accumulator is fetched BEFORE waiting for the MMAs
            float hp[2][16];
#pragma unroll
            for (int c2 = 0; c2 < 2; c2++) {
                if (P.h_prev && m < g.M && (g.debug & 0xff) != 5 && (g.debug & 0xff) != 1) state_load16(P.h_prev, m, ub * GRU_UB + (half * 2 + c2) * 16, hp[c2]);
                else {
#pragma unroll
                    for (int i = 0; i < 16; i++) hp[c2][i] = 0.f;
                }
            }
            mbar_wait(tfull_bar(as), aphase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + as * 256 + ((uint32_t)(wq * 32) << 16);
#pragma unroll
            for (int c2 = 0; c2 < 2; c2++) {
                if ((g.debug & 0xff) == 1) break;
                const int cb = half * 2 + c2;
                uint32_t vr[16], vz[16], vx[16], vh[16];
                tmem_ld16(taddr + cb * 16, vr);
                tmem_ld16(taddr + 64 + cb * 16, vz);
                tmem_ld16(taddr + 128 + cb * 16, vx);
                if (has_h) tmem_ld16(taddr + 192 + cb * 16, vh);
                tmem_ld_wait();
                gru_finish16<FAST_ACT>(P, g.M, m_warp0, lane, ub * GRU_UB + cb * 16, vr, vz, vx, vh, has_h, hp[c2], g.f16 != 0,
                                       g.acc_scale != 0.f ? g.acc_scale : 1.f, g.debug & 0xff);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(as));
            if (++as == 2) { as = 0; aphase ^= 1u; }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

// ================================================================================================
// GRU time step on CTA PAIRS (cta_group::2).  The single-CTA kernel above is shared-memory-bandwidth bound: every
// 128x192x16 MMA reads 10 KB of operands while TMA writes the next stage into the same shared memory.  Here a cluster
// of two CTAs computes a 256-row x 64-unit tile with ONE M=256, N=192 instruction per 16-wide k step:
//   * each CTA stages its own 128 rows of A and only HALF of the weight rows (96 of 192), so shared-memory traffic per
//     FLOP halves;
//   * accumulator columns per CTA: [W_in x 64 | r 64 | z 64 | W_hn h 64] (W_hn h stays separate because r multiplies it).
//     The weight matrix holds the W_ih gates in the order (n, r, z) and the W_hh gates in the order (r, z, n)
//     (model.cu), so an input chunk is one MMA into columns 0..191 and a hidden chunk one MMA into columns 64..255:
//     no zero-padded operand rows, no second MMA (an N=128 + N=64 pair for the hidden chunk cost 6 % of the launch);
//   * the accumulate flag is per instruction, so the hidden chunks always accumulate and columns 192..255 must be zero
//     when a stage is handed to the MMA issuer: the epilogue warps zero them at kernel start and again right after
//     reading them (tcgen05.st, free next to the loads); the first input chunk overwrites columns 0..191.
// Barriers: `full` (leader only; both CTAs' TMA bytes are credited to it), `empty` / `tmem_full` (per CTA, released by
// the leader's multicast tcgen05.commit), `tmem_empty` (leader, 32 arrivals: 16 epilogue warps of each CTA).
constexpr int GRU2_B_ROWS = 96;
constexpr int GRU2_STAGE_BYTES = A_STAGE_BYTES + GRU2_B_ROWS * 128;   // 16 KB A + 12 KB B
constexpr int GRU2_STAGES = 7;
constexpr int GRU2_EPI_WARPS = 16;
constexpr int GRU2_THREADS = 128 + 32 * GRU2_EPI_WARPS;   // warps 0-3: TMA producer, MMA issuer, TMEM allocator, spare
constexpr int GRU2_MAX_CHUNKS = 48;                  // per problem: 16 K-chunks x 3 split-bf16 passes
constexpr int GRU2_SMEM_BYTES = GRU2_STAGES * GRU2_STAGE_BYTES + 1024 + 256 + 2 * GRU2_MAX_CHUNKS * 16;
constexpr uint32_t GRU2_TX_BYTES = 2 * A_STAGE_BYTES + 2 * GRU2_B_ROWS * 128;   // both CTAs' A tiles and weight rows

template <bool FAST_ACT, int NP>
__global__ void __launch_bounds__(GRU2_THREADS, 1)
gru_step2_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                 const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmA3,
                 const __grid_constant__ CUtensorMap tmB96, const GruArgs g) {
    constexpr int STAGES = GRU2_STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + STAGES * GRU2_STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    uint8_t *smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));          // generic pointer to the aligned base
    uint32_t *tmem_slot_ptr = reinterpret_cast<uint32_t *>(smem_gen + (tmem_slot - smem_base));
    // the K-chunk tables of both problems, copied to shared memory once: the producer and the MMA issuer are single
    // threads and a global load per chunk in their loops (hundreds of cycles, serialised) paced the whole pipeline
    KChunk *s_chunks = reinterpret_cast<KChunk *>(smem_gen + STAGES * GRU2_STAGE_BYTES + 256);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // cluster = NP CTA pairs working on the SAME 256 rows and NP neighbouring unit blocks: every 128-row A tile is
    // needed by one CTA of each pair, so each of those CTAs fetches 128/NP of its rows and multicasts them to the others
    const uint32_t crank = cluster_ctarank();
    const uint32_t rank = crank & 1u, prk = crank >> 1;                       // rank within the pair, pair within the cluster
    const int pair = blockIdx.x / (2 * NP), n_pairs = gridDim.x / (2 * NP);   // cluster index / count (tile loop stride)
    constexpr int UBG = GRU_NUB / NP;                                         // unit-block groups per row block
    const int tiles_per_prob = g.tiles_m * UBG;                               // tiles_m counts 256-row blocks here
    const int n_tiles = tiles_per_prob * g.n_prob;
    constexpr uint16_t ALL_CTAS = (uint16_t)((1u << (2 * NP)) - 1u);

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA0); prefetch_tmap(&tmA1); prefetch_tmap(&tmA2); prefetch_tmap(&tmA3);
        prefetch_tmap(&tmB96);
    }
    if (warp == 1 && lane == 0) {
        // empty: one commit per pair of the cluster (each pair's MMAs read what every CTA of the cluster multicast)
        for (int s = 0; s < STAGES; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), NP); }
        for (int s = 0; s < 2; s++) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 2 * GRU2_EPI_WARPS); }
        fence_barrier_init();
    }
    if (warp == 2) { tmem_alloc_2sm(tmem_slot, 512); tmem_relinquish_2sm(); }
    if (warp >= 4)
        for (int pi = 0; pi < g.n_prob; pi++)
            for (int i = threadIdx.x - 128; i < g.prob[pi].n_chunks && i < GRU2_MAX_CHUNKS; i += GRU2_THREADS - 128)
                s_chunks[pi * GRU2_MAX_CHUNKS + i] = g.prob[pi].chunks[i];
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    if (warp >= 4) {                                   // W_hn h columns of both accumulator stages start at zero
        const uint32_t t0 = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + 192 + ((warp - 4) >> 2) * 16;
        tmem_st16_zero(t0); tmem_st16_zero(t0 + 256);
        tmem_st_wait();
    }
    tc_fence_before();
    cluster_sync_all();                                // the leader's MMAs write both CTAs' tensor memory
    tc_fence_after();

    if (warp == 0) {
        if (lane == 0) {
            PipeState ps{0, 0};
            for (int tile = pair; tile < n_tiles; tile += n_pairs) {
                const int pi = tile / tiles_per_prob, rem = tile % tiles_per_prob;
                const GruProblem &P = g.prob[pi];
                const int m0 = (rem / UBG) * 256 + (int)rank * BM, ub = (rem % UBG) * NP + (int)prk;
                const int brow = P.b_row_base + ub * GRU_BROWS;
                uint16_t amask = 0;                           // the CTAs of this rank in every pair
#pragma unroll
                for (int q = 0; q < NP; q++) amask |= (uint16_t)(1u << (2 * q + (int)rank));
                for (int kc = 0; kc < P.n_chunks; kc++) {
                    const KChunk c = s_chunks[pi * GRU2_MAX_CHUNKS + kc];
                    const int src = c.a_src & 0xff;
                    const CUtensorMap *ma = src == 0 ? &tmA0 : src == 1 ? &tmA1 : src == 2 ? &tmA2 : &tmA3;
                    mbar_wait(empty_bar(ps.stage), ps.phase ^ 1u);
                    const int xm = g.debug >> 8;              // timing experiments: 1 no A loads, 4 no B loads
                    if (rank == 0) mbar_expect_tx(full_bar(ps.stage), ((xm & 1) ? 0u : 2u * A_STAGE_BYTES) + ((xm & 4) ? 0u : 2u * GRU2_B_ROWS * 128u));
                    const uint32_t sa = smem_base + ps.stage * GRU2_STAGE_BYTES;
                    const uint32_t sb = sa + A_STAGE_BYTES;
                    if (xm & 1) { }
                    else if (NP == 1) tma_load_2d_2sm(sa, ma, full_bar(ps.stage), c.a_col, m0 + c.a_drow);
                    else tma_load_2d_2sm_mc(sa + prk * (uint32_t)(BM / NP) * 128u, ma, full_bar(ps.stage), c.a_col,
                                            m0 + (int)prk * (BM / NP) + c.a_drow, amask);
                    if (!(xm & 4)) tma_load_2d_2sm(sb, &tmB96, full_bar(ps.stage), c.b_col, brow + (int)rank * 96);
                    ps.advance(STAGES);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            const uint32_t idesc192 = make_idesc_16(256, 192, g.f16 != 0);
            PipeState ps{0, 0};
            int as = 0; uint32_t aphase = 0;
            for (int tile = pair; tile < n_tiles; tile += n_pairs) {
                const int pi = tile / tiles_per_prob;
                const GruProblem &P = g.prob[pi];
                mbar_wait(tempty_bar(as), aphase ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * 256;
                for (int kc = 0; kc < P.n_chunks; kc++) {
                    const int kind = (s_chunks[pi * GRU2_MAX_CHUNKS + kc].a_src >> 8) & 0xff;
                    mbar_wait(full_bar(ps.stage), ps.phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + ps.stage * GRU2_STAGE_BYTES;
                    const uint64_t da = make_smem_desc_sw128(sa);
                    const uint64_t db = make_smem_desc_sw128(sa + A_STAGE_BYTES);
                    if ((g.debug >> 8) & 2) {                  // timing experiment: no MMAs
                    } else if (kind == 0) {                    // input chunks come first (model.cu chunk tables):
#pragma unroll                                                 // columns 0..191 = [W_in x | r | z], started by the first one
                        for (int k = 0; k < BK / 16; k++)
                            umma_bf16_2sm(d_tmem, da + 2u * k, db + 2u * k, idesc192, (uint32_t)(kc | k));
                    } else {                                   // hidden chunks: columns 64..255 = [r | z | W_hn h]; the
#pragma unroll                                                 // epilogue left columns 192..255 zeroed
                        for (int k = 0; k < BK / 16; k++)
                            umma_bf16_2sm(d_tmem + 64, da + 2u * k, db + 2u * k, idesc192, 1u);
                    }
                    umma_commit_mc(empty_bar(ps.stage), ALL_CTAS);
                    ps.advance(STAGES);
                }
                umma_commit_mc(tfull_bar(as), (uint16_t)(3u << (2 * prk)));
                if (++as == 2) { as = 0; aphase ^= 1u; }
            }
        }
    } else if (warp >= 4) {
        // 16 epilogue warps: warp w reads TMEM lanes 32*(w%4).. (hardware lane-quarter rule) and handles the 16 hidden
        // units cb = (w-4)/4 of the tile's 64.  The epilogue is latency-bound per warp (state / partner loads, stores), so
        // its throughput comes from the number of warps in flight.
        const int wq = warp & 3, cb = (warp - 4) >> 2;
        int as = 0; uint32_t aphase = 0;
        for (int tile = pair; tile < n_tiles; tile += n_pairs) {
            const int pi = tile / tiles_per_prob, rem = tile % tiles_per_prob;
            const GruProblem &P = g.prob[pi];
            const int m0 = (rem / UBG) * 256 + (int)rank * BM, ub = (rem % UBG) * NP + (int)prk;
            const long long m_warp0 = m0 + wq * 32;
            const long long m = m_warp0 + lane;
            const bool has_h = P.h_prev != nullptr || P.h_prev_bf != nullptr;   // else no hidden chunks: columns 192..255 are never written
            const int dbg = g.debug & 0xff;
            // everything that does not depend on the accumulator is fetched BEFORE waiting for the MMAs
            float hp[16];
            if (P.h_prev && m < g.M && dbg != 5 && dbg != 1) state_load16(P.h_prev, m, ub * GRU_UB + cb * 16, hp);
            else if (P.h_prev_bf && m < g.M && dbg != 5 && dbg != 1) rows_load_bf16(P.h_prev_bf, m, ub * GRU_UB + cb * 16, hp, g.f16 != 0);
            else {
#pragma unroll
                for (int i = 0; i < 16; i++) hp[i] = 0.f;
            }
            mbar_wait(tfull_bar(as), aphase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + as * 256 + ((uint32_t)(wq * 32) << 16);
            if (dbg != 1) {
                uint32_t vr[16], vz[16], vx[16], vh[16];
                tmem_ld16(taddr + cb * 16, vx);
                tmem_ld16(taddr + 64 + cb * 16, vr);
                tmem_ld16(taddr + 128 + cb * 16, vz);
                if (has_h) tmem_ld16(taddr + 192 + cb * 16, vh);
                tmem_ld_wait();
                // the accumulator stage is free as soon as its values are in registers and the W_hn h columns, which the
                // hidden chunks only ever accumulate into, are back to zero
                if (has_h) { tmem_st16_zero(taddr + 192 + cb * 16); tmem_st_wait(); }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(tempty_bar(as), crank & ~1u);
                gru_finish16<FAST_ACT>(P, g.M, m_warp0, lane, ub * GRU_UB + cb * 16, vr, vz, vx, vh, has_h, hp, g.f16 != 0,
                                       g.acc_scale != 0.f ? g.acc_scale : 1.f, dbg);
            } else {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(tempty_bar(as), crank & ~1u);
            }
            if (++as == 2) { as = 0; aphase ^= 1u; }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 2) { tc_fence_after(); tmem_dealloc_2sm(tmem_base, 512); }
}

// ================================================================================================
// WIDE variant of the CTA-pair kernel: tile = 256 rows x 128 hidden units (two unit blocks).  The pair kernel above pulls
// (128 + 96) operand rows per CTA and k-chunk out of L2 for a 64-unit tile and is bound by that stream; here one A tile
// feeds the MMAs of TWO unit blocks, (128 + 2 x 96) rows for twice the FLOPs = 29 % fewer operand bytes per FLOP.  Price:
// the two accumulator sets [W_in x | r | z | W_hn h] x 64 fill all 512 TMEM columns, so there is no second accumulator
// stage -- the MMAs of the next tile start when every epilogue warp has its SECOND group of 16 units in registers, i.e.
// the tensor pipe idles for about one group's gate math per tile.  5-stage ring of (16 KB A + 2 x 12 KB weights).
constexpr int GRUW_STAGE_BYTES = A_STAGE_BYTES + 2 * GRU2_B_ROWS * 128;
constexpr int GRUW_STAGES = 5;
constexpr int GRUW_SMEM_BYTES = GRUW_STAGES * GRUW_STAGE_BYTES + 1024 + 256 + 2 * GRU2_MAX_CHUNKS * 16;

template <bool FAST_ACT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GRU2_THREADS, 1)
gru_step2w_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                  const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmA3,
                  const __grid_constant__ CUtensorMap tmB96, const GruArgs g) {
    constexpr int STAGES = GRUW_STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + STAGES * GRUW_STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    const uint32_t tfull_bar = bar_base + 8u * (2 * STAGES), tempty_bar = bar_base + 8u * (2 * STAGES + 1);
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 2);
    uint8_t *smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    uint32_t *tmem_slot_ptr = reinterpret_cast<uint32_t *>(smem_gen + (tmem_slot - smem_base));
    KChunk *s_chunks = reinterpret_cast<KChunk *>(smem_gen + STAGES * GRUW_STAGE_BYTES + 256);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    constexpr int UBG = GRU_NUB / 2;                                          // 128-unit groups per row block
    const int tiles_per_prob = g.tiles_m * UBG;
    const int n_tiles = tiles_per_prob * g.n_prob;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA0); prefetch_tmap(&tmA1); prefetch_tmap(&tmA2); prefetch_tmap(&tmA3);
        prefetch_tmap(&tmB96);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(tfull_bar, 1); mbar_init(tempty_bar, 2 * GRU2_EPI_WARPS);
        fence_barrier_init();
    }
    if (warp == 2) { tmem_alloc_2sm(tmem_slot, 512); tmem_relinquish_2sm(); }
    if (warp >= 4)
        for (int pi = 0; pi < g.n_prob; pi++)
            for (int i = threadIdx.x - 128; i < g.prob[pi].n_chunks && i < GRU2_MAX_CHUNKS; i += GRU2_THREADS - 128)
                s_chunks[pi * GRU2_MAX_CHUNKS + i] = g.prob[pi].chunks[i];
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    if (warp >= 4) {                                   // W_hn h columns of both unit blocks start at zero
        const uint32_t t0 = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + 192 + ((warp - 4) >> 2) * 16;
        tmem_st16_zero(t0); tmem_st16_zero(t0 + 256);
        tmem_st_wait();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();

    if (warp == 0) {
        if (lane == 0) {
            PipeState ps{0, 0};
            for (int tile = pair; tile < n_tiles; tile += n_pairs) {
                const int pi = tile / tiles_per_prob, rem = tile % tiles_per_prob;
                const GruProblem &P = g.prob[pi];
                const int m0 = (rem / UBG) * 256 + (int)rank * BM, ub = (rem % UBG) * 2;
                const int brow = P.b_row_base + ub * GRU_BROWS;
                for (int kc = 0; kc < P.n_chunks; kc++) {
                    const KChunk c = s_chunks[pi * GRU2_MAX_CHUNKS + kc];
                    const int src = c.a_src & 0xff;
                    const CUtensorMap *ma = src == 0 ? &tmA0 : src == 1 ? &tmA1 : src == 2 ? &tmA2 : &tmA3;
                    mbar_wait(empty_bar(ps.stage), ps.phase ^ 1u);
                    if (rank == 0) mbar_expect_tx(full_bar(ps.stage), 2u * (uint32_t)GRUW_STAGE_BYTES);
                    const uint32_t sa = smem_base + ps.stage * GRUW_STAGE_BYTES;
                    const uint32_t sb = sa + A_STAGE_BYTES;
                    tma_load_2d_2sm(sa, ma, full_bar(ps.stage), c.a_col, m0 + c.a_drow);
                    tma_load_2d_2sm(sb, &tmB96, full_bar(ps.stage), c.b_col, brow + (int)rank * 96);
                    tma_load_2d_2sm(sb + GRU2_B_ROWS * 128u, &tmB96, full_bar(ps.stage), c.b_col, brow + GRU_BROWS + (int)rank * 96);
                    ps.advance(STAGES);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            const uint32_t idesc192 = make_idesc_16(256, 192, g.f16 != 0);
            PipeState ps{0, 0};
            uint32_t aphase = 0;
            for (int tile = pair; tile < n_tiles; tile += n_pairs) {
                const int pi = tile / tiles_per_prob;
                const GruProblem &P = g.prob[pi];
                mbar_wait(tempty_bar, aphase ^ 1u);
                tc_fence_after();
                for (int kc = 0; kc < P.n_chunks; kc++) {
                    const int kind = (s_chunks[pi * GRU2_MAX_CHUNKS + kc].a_src >> 8) & 0xff;
                    mbar_wait(full_bar(ps.stage), ps.phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + ps.stage * GRUW_STAGE_BYTES;
                    const uint64_t da = make_smem_desc_sw128(sa);
                    const uint64_t db0 = make_smem_desc_sw128(sa + A_STAGE_BYTES);
                    const uint64_t db1 = make_smem_desc_sw128(sa + A_STAGE_BYTES + GRU2_B_ROWS * 128u);
                    const uint32_t d0 = tmem_base + (kind == 0 ? 0u : 64u);   // input: columns 0..191, hidden: 64..255
                    if (!((g.debug >> 8) & 2)) {
#pragma unroll
                        for (int k = 0; k < BK / 16; k++) {
                            const uint32_t acc = kind == 0 ? (uint32_t)(kc | k) : 1u;
                            umma_bf16_2sm(d0, da + 2u * k, db0 + 2u * k, idesc192, acc);
                            umma_bf16_2sm(d0 + 256, da + 2u * k, db1 + 2u * k, idesc192, acc);
                        }
                    }
                    umma_commit_mc(empty_bar(ps.stage), 3);
                    ps.advance(STAGES);
                }
                umma_commit_mc(tfull_bar, 3);
                aphase ^= 1u;
            }
        }
    } else if (warp >= 4) {
        // warp w: TMEM lanes 32*(w%4).., units cb = (w-4)/4 of BOTH unit blocks, one after the other; the accumulators are
        // handed back to the MMA issuer when the second group is in registers
        const int wq = warp & 3, cb = (warp - 4) >> 2;
        uint32_t aphase = 0;
        for (int tile = pair; tile < n_tiles; tile += n_pairs) {
            const int pi = tile / tiles_per_prob, rem = tile % tiles_per_prob;
            const GruProblem &P = g.prob[pi];
            const int m0 = (rem / UBG) * 256 + (int)rank * BM, ub = (rem % UBG) * 2;
            const long long m_warp0 = m0 + wq * 32;
            const long long m = m_warp0 + lane;
            const bool has_h = P.h_prev != nullptr || P.h_prev_bf != nullptr;
            const int dbg = g.debug & 0xff;
            const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16);
            if (dbg == 1) {                              // timing experiment: accumulators handed straight back
                mbar_wait(tfull_bar, aphase);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(tempty_bar, 0);
                aphase ^= 1u;
                continue;
            }
#pragma unroll 1
            for (int hb = 0; hb < 2; hb++) {
                const int u0 = (ub + hb) * GRU_UB + cb * 16;
                float hp[16];
                if (P.h_prev && m < g.M) state_load16(P.h_prev, m, u0, hp);
                else if (P.h_prev_bf && m < g.M) rows_load_bf16(P.h_prev_bf, m, u0, hp, g.f16 != 0);
                else {
#pragma unroll
                    for (int i = 0; i < 16; i++) hp[i] = 0.f;
                }
                if (hb == 0) { mbar_wait(tfull_bar, aphase); tc_fence_after(); }
                uint32_t vr[16], vz[16], vx[16], vh[16];
                const uint32_t t = taddr + hb * 256 + cb * 16;
                tmem_ld16(t, vx);
                tmem_ld16(t + 64, vr);
                tmem_ld16(t + 128, vz);
                if (has_h) tmem_ld16(t + 192, vh);
                tmem_ld_wait();
                if (has_h) { tmem_st16_zero(t + 192); tmem_st_wait(); }
                if (hb == 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(tempty_bar, 0);
                }
                gru_finish16<FAST_ACT>(P, g.M, m_warp0, lane, u0, vr, vz, vx, vh, has_h, hp, g.f16 != 0,
                                       g.acc_scale != 0.f ? g.acc_scale : 1.f, dbg);
            }
            aphase ^= 1u;
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 2) { tc_fence_after(); tmem_dealloc_2sm(tmem_base, 512); }
}

int launch_gru_step2w(const CUtensorMap *tmA, const CUtensorMap *tmB96, const GruArgs &g_in, cudaStream_t st) {
    static bool attr_done[64] = {false};
    int dev = 0;
    DP_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_done[dev]) {
        DP_CUDA_CHECK(cudaFuncSetAttribute(gru_step2w_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRUW_SMEM_BYTES));
        DP_CUDA_CHECK(cudaFuncSetAttribute(gru_step2w_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRUW_SMEM_BYTES));
        if (dev < 64) attr_done[dev] = true;
    }
    static int dbgw = -1;
    if (dbgw < 0) {
        const char *e = getenv("DP_GRU_DEBUG"); dbgw = e ? atoi(e) : 0;
        const char *e2 = getenv("DP_GRU_DEBUG2"); if (e2) dbgw |= atoi(e2) << 8;
    }
    GruArgs g = g_in;
    g.debug = dbgw;
    g.tiles_m = (g.M + 255) / 256;
    const int n_tiles = g.tiles_m * (GRU_NUB / 2) * g.n_prob;
    if (n_tiles <= 0) return DP_OK;
    for (int i = 0; i < g.n_prob; i++)
        if (g.prob[i].n_chunks > GRU2_MAX_CHUNKS) { set_error("gru_step2w: too many K chunks"); return DP_EINVAL; }
    const int max_pairs = device_sm_count() / 2;
    const int pairs = n_tiles < max_pairs ? n_tiles : max_pairs;
    if (g.fast_act) gru_step2w_kernel<true><<<2 * pairs, GRU2_THREADS, GRUW_SMEM_BYTES, st>>>(tmA[0], tmA[1], tmA[2], tmA[3], *tmB96, g);
    else gru_step2w_kernel<false><<<2 * pairs, GRU2_THREADS, GRUW_SMEM_BYTES, st>>>(tmA[0], tmA[1], tmA[2], tmA[3], *tmB96, g);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

// ================================================================================================
// STAGGERED wide variant (DP_GRU=stag): tile = 256 rows x 128 hidden units like gru_step2w_kernel, but the two unit
// blocks ("halves", 256 accumulator columns each) run their K loops L chunks apart.  The accumulator of half 0 is
// complete L chunk-times before that of half 1, and while it drains to registers the tensor pipe finishes half 1; half 0
// of the NEXT tile then starts L chunks ahead of half 1 while half 1 drains.  The two halves therefore act as the two
// accumulator stages the wide tile had no TMEM columns for, and every A tile still feeds the MMAs of two unit blocks
// (29 % fewer operand bytes per FLOP out of L2 than the 64-unit tile).  Price: an A tile stays in shared memory for L
// more chunks, so A and the weight tiles live in separate rings (A: NA x 16 KB released by half 1; weights: NB x 12 KB
// consumed in issue order B0(c), B1(c - L)).
// L = chunks between the two halves, NA = A ring stages (>= L + prefetch depth), NB = weight ring stages
constexpr int GRUS_B_BYTES = GRU2_B_ROWS * 128;  // 12 KB per CTA and weight tile
constexpr int GRUS_BAR_BYTES = 512;
__host__ __device__ constexpr int grus_smem_bytes(int na, int nb) {
    return na * A_STAGE_BYTES + nb * GRUS_B_BYTES + 1024 + GRUS_BAR_BYTES + 2 * GRU2_MAX_CHUNKS * 16;
}

template <bool FAST_ACT, int L, int NA, int NB>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GRU2_THREADS, 1)
gru_step2s_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                  const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmA3,
                  const __grid_constant__ CUtensorMap tmB96, const GruArgs g) {
    static_assert(grus_smem_bytes(NA, NB) <= 232448 && NA > L, "ring configuration");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t b_base = smem_base + NA * A_STAGE_BYTES;
    const uint32_t bar_base = b_base + NB * GRUS_B_BYTES;
    auto a_full = [&](int s) { return bar_base + 8u * s; };
    auto a_empty = [&](int s) { return bar_base + 8u * (NA + s); };
    auto b_full = [&](int s) { return bar_base + 8u * (2 * NA + s); };
    auto b_empty = [&](int s) { return bar_base + 8u * (2 * NA + NB + s); };
    auto tfull_bar = [&](int h) { return bar_base + 8u * (2 * NA + 2 * NB + h); };
    auto tempty_bar = [&](int h) { return bar_base + 8u * (2 * NA + 2 * NB + 2 + h); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * NA + 2 * NB + 4);
    static_assert(8 * (2 * NA + 2 * NB + 5) <= GRUS_BAR_BYTES, "barrier area too small");
    uint8_t *smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    uint32_t *tmem_slot_ptr = reinterpret_cast<uint32_t *>(smem_gen + (tmem_slot - smem_base));
    KChunk *s_chunks = reinterpret_cast<KChunk *>(smem_gen + (bar_base - smem_base) + GRUS_BAR_BYTES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    constexpr int UBG = GRU_NUB / 2;                                          // 128-unit groups per row block
    const int tiles_per_prob = g.tiles_m * UBG;
    const int n_tiles = tiles_per_prob * g.n_prob;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA0); prefetch_tmap(&tmA1); prefetch_tmap(&tmA2); prefetch_tmap(&tmA3);
        prefetch_tmap(&tmB96);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < NA; s++) { mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1); }
        for (int s = 0; s < NB; s++) { mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1); }
        for (int h = 0; h < 2; h++) { mbar_init(tfull_bar(h), 1); mbar_init(tempty_bar(h), 2 * GRU2_EPI_WARPS); }
        fence_barrier_init();
    }
    if (warp == 2) { tmem_alloc_2sm(tmem_slot, 512); tmem_relinquish_2sm(); }
    if (warp >= 4)
        for (int pi = 0; pi < g.n_prob; pi++)
            for (int i = threadIdx.x - 128; i < g.prob[pi].n_chunks && i < GRU2_MAX_CHUNKS; i += GRU2_THREADS - 128)
                s_chunks[pi * GRU2_MAX_CHUNKS + i] = g.prob[pi].chunks[i];
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    if (warp >= 4) {                                   // W_hn h columns of both halves start at zero
        const uint32_t t0 = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + 192 + ((warp - 4) >> 2) * 16;
        tmem_st16_zero(t0); tmem_st16_zero(t0 + 256);
        tmem_st_wait();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();

    if (warp == 0) {
        if (lane == 0) {
            PipeState pa{0, 0}, pb{0, 0};
            for (int tile = pair; tile < n_tiles; tile += n_pairs) {
                const int pi = tile / tiles_per_prob, rem = tile % tiles_per_prob;
                const GruProblem &P = g.prob[pi];
                const int m0 = (rem / UBG) * 256 + (int)rank * BM, ub = (rem % UBG) * 2;
                const int brow = P.b_row_base + ub * GRU_BROWS + (int)rank * 96;
                const int n = P.n_chunks;
                for (int c = 0; c < n + L; c++) {
                    if (c < n) {
                        const KChunk ch = s_chunks[pi * GRU2_MAX_CHUNKS + c];
                        const int src = ch.a_src & 0xff;
                        const CUtensorMap *ma = src == 0 ? &tmA0 : src == 1 ? &tmA1 : src == 2 ? &tmA2 : &tmA3;
                        mbar_wait(a_empty(pa.stage), pa.phase ^ 1u);
                        if (rank == 0) mbar_expect_tx(a_full(pa.stage), 2u * (uint32_t)A_STAGE_BYTES);
                        tma_load_2d_2sm(smem_base + pa.stage * A_STAGE_BYTES, ma, a_full(pa.stage), ch.a_col, m0 + ch.a_drow);
                        pa.advance(NA);
                        mbar_wait(b_empty(pb.stage), pb.phase ^ 1u);
                        if (rank == 0) mbar_expect_tx(b_full(pb.stage), 2u * (uint32_t)GRUS_B_BYTES);
                        tma_load_2d_2sm(b_base + pb.stage * GRUS_B_BYTES, &tmB96, b_full(pb.stage), ch.b_col, brow);
                        pb.advance(NB);
                    }
                    if (c >= L) {
                        const KChunk ch = s_chunks[pi * GRU2_MAX_CHUNKS + c - L];
                        mbar_wait(b_empty(pb.stage), pb.phase ^ 1u);
                        if (rank == 0) mbar_expect_tx(b_full(pb.stage), 2u * (uint32_t)GRUS_B_BYTES);
                        tma_load_2d_2sm(b_base + pb.stage * GRUS_B_BYTES, &tmB96, b_full(pb.stage), ch.b_col, brow + GRU_BROWS);
                        pb.advance(NB);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            const uint32_t idesc192 = make_idesc_16(256, 192, g.f16 != 0);
            PipeState ph{0, 0}, pt{0, 0}, pb{0, 0};          // A ring head (half 0) / tail (half 1, releases), weight ring
            uint32_t tphase = 0;
            for (int tile = pair; tile < n_tiles; tile += n_pairs) {
                const int pi = tile / tiles_per_prob;
                const int n = g.prob[pi].n_chunks;
                for (int c = 0; c < n + L; c++) {
                    if (c < n) {
                        const int kind = (s_chunks[pi * GRU2_MAX_CHUNKS + c].a_src >> 8) & 0xff;
                        if (c == 0) { mbar_wait(tempty_bar(0), tphase ^ 1u); tc_fence_after(); }
                        mbar_wait(a_full(ph.stage), ph.phase);
                        mbar_wait(b_full(pb.stage), pb.phase);
                        tc_fence_after();
                        const uint64_t da = make_smem_desc_sw128(smem_base + ph.stage * A_STAGE_BYTES);
                        const uint64_t db = make_smem_desc_sw128(b_base + pb.stage * GRUS_B_BYTES);
                        const uint32_t d0 = tmem_base + (kind == 0 ? 0u : 64u);   // input: columns 0..191, hidden: 64..255
                        if (!((g.debug >> 8) & 2)) {
#pragma unroll
                            for (int k = 0; k < BK / 16; k++)
                                umma_bf16_2sm(d0, da + 2u * k, db + 2u * k, idesc192, kind == 0 ? (uint32_t)(c | k) : 1u);
                        }
                        umma_commit_mc(b_empty(pb.stage), 3);
                        ph.advance(NA); pb.advance(NB);
                        if (c == n - 1) umma_commit_mc(tfull_bar(0), 3);
                    }
                    if (c >= L) {
                        const int cc = c - L;
                        const int kind = (s_chunks[pi * GRU2_MAX_CHUNKS + cc].a_src >> 8) & 0xff;
                        if (cc == 0) { mbar_wait(tempty_bar(1), tphase ^ 1u); tc_fence_after(); }
                        mbar_wait(b_full(pb.stage), pb.phase);
                        tc_fence_after();
                        const uint64_t da = make_smem_desc_sw128(smem_base + pt.stage * A_STAGE_BYTES);   // full since half 0 used it
                        const uint64_t db = make_smem_desc_sw128(b_base + pb.stage * GRUS_B_BYTES);
                        const uint32_t d0 = tmem_base + 256u + (kind == 0 ? 0u : 64u);
                        if (!((g.debug >> 8) & 2)) {
#pragma unroll
                            for (int k = 0; k < BK / 16; k++)
                                umma_bf16_2sm(d0, da + 2u * k, db + 2u * k, idesc192, kind == 0 ? (uint32_t)(cc | k) : 1u);
                        }
                        umma_commit_mc(b_empty(pb.stage), 3);
                        umma_commit_mc(a_empty(pt.stage), 3);
                        pt.advance(NA); pb.advance(NB);
                        if (cc == n - 1) umma_commit_mc(tfull_bar(1), 3);
                    }
                }
                tphase ^= 1u;
            }
        }
    } else if (warp >= 4) {
        // warp w: TMEM lanes 32*(w%4).., units cb = (w-4)/4 of half 0, then of half 1; each half's accumulator is handed
        // back as soon as this warp's share of it is in registers
        const int wq = warp & 3, cb = (warp - 4) >> 2;
        uint32_t tphase = 0;
        for (int tile = pair; tile < n_tiles; tile += n_pairs) {
            const int pi = tile / tiles_per_prob, rem = tile % tiles_per_prob;
            const GruProblem &P = g.prob[pi];
            const int m0 = (rem / UBG) * 256 + (int)rank * BM, ub = (rem % UBG) * 2;
            const long long m_warp0 = m0 + wq * 32;
            const long long m = m_warp0 + lane;
            const bool has_h = P.h_prev != nullptr || P.h_prev_bf != nullptr;
            const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16);
#pragma unroll 1
            for (int hb = 0; hb < 2; hb++) {
                const int u0 = (ub + hb) * GRU_UB + cb * 16;
                float hp[16];
                const int dbg = g.debug & 0xff;
                if (dbg == 1) {                          // timing experiment: accumulators handed straight back
                    mbar_wait(tfull_bar(hb), tphase);
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(tempty_bar(hb), 0);
                    continue;
                }
                if (P.h_prev && m < g.M) state_load16(P.h_prev, m, u0, hp);
                else if (P.h_prev_bf && m < g.M) rows_load_bf16(P.h_prev_bf, m, u0, hp, g.f16 != 0);
                else {
#pragma unroll
                    for (int i = 0; i < 16; i++) hp[i] = 0.f;
                }
                mbar_wait(tfull_bar(hb), tphase);
                tc_fence_after();
                uint32_t vr[16], vz[16], vx[16], vh[16];
                const uint32_t t = taddr + hb * 256 + cb * 16;
                tmem_ld16(t, vx);
                tmem_ld16(t + 64, vr);
                tmem_ld16(t + 128, vz);
                if (has_h) tmem_ld16(t + 192, vh);
                tmem_ld_wait();
                if (has_h) { tmem_st16_zero(t + 192); tmem_st_wait(); }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(tempty_bar(hb), 0);
                gru_finish16<FAST_ACT>(P, g.M, m_warp0, lane, u0, vr, vz, vx, vh, has_h, hp, g.f16 != 0,
                                       g.acc_scale != 0.f ? g.acc_scale : 1.f, dbg);
            }
            tphase ^= 1u;
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 2) { tc_fence_after(); tmem_dealloc_2sm(tmem_base, 512); }
}

template <bool FAST_ACT, int L, int NA, int NB>
static int launch_gru_step2s_t(const CUtensorMap *tmA, const CUtensorMap *tmB96, const GruArgs &g, int pairs, cudaStream_t st) {
    constexpr int SMEM = grus_smem_bytes(NA, NB);
    static bool attr_done[64] = {false};
    int dev = 0;
    DP_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_done[dev]) {
        DP_CUDA_CHECK(cudaFuncSetAttribute(gru_step2s_kernel<FAST_ACT, L, NA, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        if (dev < 64) attr_done[dev] = true;
    }
    gru_step2s_kernel<FAST_ACT, L, NA, NB><<<2 * pairs, GRU2_THREADS, SMEM, st>>>(tmA[0], tmA[1], tmA[2], tmA[3], *tmB96, g);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

int launch_gru_step2s(const CUtensorMap *tmA, const CUtensorMap *tmB96, const GruArgs &g_in, cudaStream_t st) {
    static int dbg = -1, cfg = 0;
    if (dbg < 0) {
        const char *e = getenv("DP_GRU_DEBUG"); dbg = e ? atoi(e) : 0;
        const char *e2 = getenv("DP_GRU_DEBUG2"); if (e2) dbg |= atoi(e2) << 8;
        const char *e3 = getenv("DP_GRUS_CFG"); cfg = e3 ? atoi(e3) : 0;      // ring configuration (lag, A stages, B stages)
    }
    GruArgs g = g_in;
    g.debug = dbg;
    g.tiles_m = (g.M + 255) / 256;
    const int n_tiles = g.tiles_m * (GRU_NUB / 2) * g.n_prob;
    if (n_tiles <= 0) return DP_OK;
    for (int i = 0; i < g.n_prob; i++)
        if (g.prob[i].n_chunks > GRU2_MAX_CHUNKS || g.prob[i].n_chunks < 1) { set_error("gru_step2s: bad K chunk count"); return DP_EINVAL; }
    const int max_pairs = device_sm_count() / 2;
    const int pairs = n_tiles < max_pairs ? n_tiles : max_pairs;
#define DP_GRUS_CASE(id, L_, NA_, NB_)                                                                           \
    if (cfg == id) return g.fast_act ? launch_gru_step2s_t<true, L_, NA_, NB_>(tmA, tmB96, g, pairs, st)          \
                                     : launch_gru_step2s_t<false, L_, NA_, NB_>(tmA, tmB96, g, pairs, st);
    DP_GRUS_CASE(1, 6, 9, 6)
    DP_GRUS_CASE(2, 4, 8, 7)
    DP_GRUS_CASE(3, 2, 7, 9)
    DP_GRUS_CASE(4, 2, 6, 10)
    DP_GRUS_CASE(5, 1, 6, 10)
#undef DP_GRUS_CASE
    return g.fast_act ? launch_gru_step2s_t<true, 3, 7, 8>(tmA, tmB96, g, pairs, st)
                      : launch_gru_step2s_t<false, 3, 7, 8>(tmA, tmB96, g, pairs, st);
}

template <bool FAST_ACT, int NP>
static int launch_gru_step2_t(const CUtensorMap *tmA, const CUtensorMap *tmB96, const GruArgs &g, int n_ctiles, cudaStream_t st) {
    auto kern = gru_step2_kernel<FAST_ACT, NP>;
    static bool attr_done[64] = {false};
    static int max_clusters[64] = {0};
    int dev = 0;
    DP_CUDA_CHECK(cudaGetDevice(&dev));
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2 * NP; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.blockDim = dim3(GRU2_THREADS, 1, 1);
    cfg.dynamicSmemBytes = GRU2_SMEM_BYTES;
    cfg.stream = st;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (dev >= 64 || !attr_done[dev]) {
        DP_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GRU2_SMEM_BYTES));
        // clusters that fit at once (a GPC whose SM count is not a multiple of the cluster size leaves SMs idle)
        cfg.gridDim = dim3((unsigned)(2 * NP * (device_sm_count() / (2 * NP))), 1, 1);
        int nc = 0;
        DP_CUDA_CHECK(cudaOccupancyMaxActiveClusters(&nc, kern, &cfg));
        if (nc <= 0) { set_error("gru_step2: no cluster of %d CTAs fits on this device", 2 * NP); return DP_ECUDA; }
        if (dev < 64) { max_clusters[dev] = nc; attr_done[dev] = true; }
        else max_clusters[0] = nc;
        if (getenv("DP_GRU_VERBOSE")) fprintf(stderr, "gru_step2: cluster of %d CTAs, %d clusters resident\n", 2 * NP, nc);
    }
    const int cap = max_clusters[dev < 64 ? dev : 0];
    const int clusters = n_ctiles < cap ? n_ctiles : cap;
    cfg.gridDim = dim3((unsigned)(2 * NP * clusters), 1, 1);
    DP_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA[0], tmA[1], tmA[2], tmA[3], *tmB96, g));
    return DP_OK;
}

// np = CTA pairs per cluster (1, 2 or 4); tmA must have boxes of 128 / np rows
int launch_gru_step2(const CUtensorMap *tmA, const CUtensorMap *tmB96, const GruArgs &g_in, int np, cudaStream_t st) {
    static int dbg = -1;
    if (dbg < 0) {
        const char *e = getenv("DP_GRU_DEBUG"); dbg = e ? atoi(e) : 0;
        const char *e2 = getenv("DP_GRU_DEBUG2"); if (e2) dbg |= atoi(e2) << 8;
    }
    GruArgs g = g_in;
    g.debug = dbg;
    g.tiles_m = (g.M + 255) / 256;
    if (np != 1 && np != 2 && np != 4) { set_error("gru_step2: pairs per cluster must be 1, 2 or 4"); return DP_EINVAL; }
    const int n_ctiles = g.tiles_m * (GRU_NUB / np) * g.n_prob;
    if (n_ctiles <= 0) return DP_OK;
    for (int i = 0; i < g.n_prob; i++)
        if (g.prob[i].n_chunks > GRU2_MAX_CHUNKS) { set_error("gru_step2: too many K chunks"); return DP_EINVAL; }
    if (g.fast_act) {
        if (np == 1) return launch_gru_step2_t<true, 1>(tmA, tmB96, g, n_ctiles, st);
        if (np == 2) return launch_gru_step2_t<true, 2>(tmA, tmB96, g, n_ctiles, st);
        return launch_gru_step2_t<true, 4>(tmA, tmB96, g, n_ctiles, st);
    }
    if (np == 1) return launch_gru_step2_t<false, 1>(tmA, tmB96, g, n_ctiles, st);
    if (np == 2) return launch_gru_step2_t<false, 2>(tmA, tmB96, g, n_ctiles, st);
    return launch_gru_step2_t<false, 4>(tmA, tmB96, g, n_ctiles, st);
}

int launch_gru_step(const CUtensorMap *tmA, const CUtensorMap *tmB, const GruArgs &g_in, cudaStream_t st) {
    static bool attr_done[64] = {false};
    int dev = 0;
    DP_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_done[dev]) {
        DP_CUDA_CHECK(cudaFuncSetAttribute(gru_step_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRU_SMEM_BYTES));
        DP_CUDA_CHECK(cudaFuncSetAttribute(gru_step_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRU_SMEM_BYTES));
        if (dev < 64) attr_done[dev] = true;
    }
    static int dbg = -1;
    if (dbg < 0) { const char *e = getenv("DP_GRU_DEBUG"); dbg = e ? atoi(e) : 0; }
    GruArgs g = g_in;
    g.debug = dbg;
    const int n_tiles = g.tiles_m * GRU_NUB * g.n_prob;
    if (n_tiles <= 0) return DP_OK;
    const int grid = n_tiles < device_sm_count() ? n_tiles : device_sm_count();
    if (g.fast_act) gru_step_kernel<true><<<grid, GRU_THREADS, GRU_SMEM_BYTES, st>>>(tmA[0], tmA[1], tmA[2], tmA[3], *tmB, g);
    else gru_step_kernel<false><<<grid, GRU_THREADS, GRU_SMEM_BYTES, st>>>(tmA[0], tmA[1], tmA[2], tmA[3], *tmB, g);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

}  // namespace dp
