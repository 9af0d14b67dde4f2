// PNG files of pileup tensors, written on the device ("batched PNG encode of device tensors").
//
// Replaces the per-window host work of Pileup.savePileupRGB (/root/reference/src/utils/Pileup.py:336-361: Image.new("RGBA",
// (coverageCutoff, windowCutoff)), 30 000 pixel stores, img.save(..., "PNG")): every window's uint8[H][W][4] tensor becomes
// a complete PNG byte stream in HBM -- signature, IHDR, ONE IDAT chunk holding a zlib stream of STORED deflate blocks
// (RFC 1950 / 1951: filter byte 0 + the raw RGBA row per scanline, <= 65 535 bytes per block), Adler-32, the chunk CRC-32s,
// IEND -- and only finished files cross PCIe.  Any PNG reader (PIL in modules.dataset.PileupDataset, dataset.py:47-52)
// decodes them to the identical pixels; the files are larger than zlib-compressed ones (120 kB instead of a few kB).
//
// One CTA per image, 256 threads:
//   phase 1  Adler-32 of the scanline stream: per-thread partial sums over a contiguous slice, combined with the
//            position weights (adler = 1 + sum d_i, b = n + sum (n - i) d_i, both mod 65521)
//   phase 2  every thread writes a contiguous slice of the IDAT bytes (type, zlib header, block headers, scanlines,
//            Adler-32) and computes the CRC-32 of exactly those bytes (table driven); the slices' CRCs are chained with the
//            polynomial shift x^(8 * slice length) mod P (zlib's crc32_combine identity), which is the same operator for all
//            slices but the last
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/deepore_b200.h"
#include "common.cuh"

namespace {

constexpr int PNG_THREADS = 256;
constexpr uint32_t CRC_POLY = 0xedb88320u;          // reflected CRC-32 polynomial (PNG, zlib)

__device__ __forceinline__ uint32_t crc_multmodp(uint32_t a, uint32_t b)     // a(x) * b(x) mod P, reflected bit order
{
    uint32_t m = 1u << 31, p = 0;
    for (;;) {
        if (a & m) {
            p ^= b;
            if ((a & (m - 1)) == 0) break;
        }
        m >>= 1;
        b = (b & 1u) ? (b >> 1) ^ CRC_POLY : b >> 1;
    }
    return p;
}

__device__ uint32_t crc_x8n(uint32_t n_bytes)       // x^(8 n) mod P
{
    uint32_t p = 1u << 31;                          // x^0
    uint32_t sq = crc_multmodp(1u << 30, 1u << 30);  // x^2 ...
    sq = crc_multmodp(sq, sq);                      // x^4
    sq = crc_multmodp(sq, sq);                      // x^8 = shift by one byte
    while (n_bytes) {
        if (n_bytes & 1u) p = crc_multmodp(sq, p);
        sq = crc_multmodp(sq, sq);
        n_bytes >>= 1;
    }
    return p;
}

struct PngGeom {
    int H, W;
    uint32_t line;            // 1 + 4 W bytes per scanline (filter byte + RGBA)
    uint32_t n_data;          // H * line: the uncompressed zlib payload
    uint32_t n_blocks;        // stored deflate blocks of <= 65535 bytes
    uint32_t n_idat;          // 2 + 5 n_blocks + n_data + 4: IDAT chunk data
    uint32_t n_file;          // whole file
};

__host__ __device__ inline PngGeom png_geom(int H, int W)
{
    PngGeom g;
    g.H = H; g.W = W;
    g.line = 1u + 4u * (uint32_t)W;
    g.n_data = (uint32_t)H * g.line;
    g.n_blocks = g.n_data ? (g.n_data + 65534u) / 65535u : 1u;
    g.n_idat = 2u + 5u * g.n_blocks + g.n_data + 4u;
    g.n_file = 8u + 25u + 12u + g.n_idat + 12u;
    return g;
}

__device__ __forceinline__ void put_be32(uint8_t *p, uint32_t v)
{
    p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v;
}

// byte `i` of the CRC-covered IDAT stream S = "IDAT" + zlib header + blocks + adler
__device__ __forceinline__ uint8_t idat_byte(const PngGeom &g, const uint8_t *img, uint32_t adler, uint32_t i)
{
    if (i < 4) return (uint8_t)"IDAT"[i];
    i -= 4;
    if (i < 2) return i == 0 ? 0x78 : 0x01;                      // CM = 8, CINFO = 7, no dictionary, check bits
    i -= 2;
    const uint32_t body = 5u * g.n_blocks + g.n_data;
    if (i >= body) {                                             // Adler-32, big-endian
        const uint32_t k = i - body;
        return (uint8_t)(adler >> (24 - 8 * k));
    }
    const uint32_t blk = i / 65540u, o = i - blk * 65540u;       // 5 header bytes + 65535 payload bytes per full block
    if (o < 5) {
        const uint32_t left = g.n_data - blk * 65535u;
        const uint32_t len = left < 65535u ? left : 65535u;
        if (o == 0) return blk + 1 == g.n_blocks ? 1 : 0;        // BFINAL, BTYPE = 00
        if (o == 1) return (uint8_t)len;
        if (o == 2) return (uint8_t)(len >> 8);
        if (o == 3) return (uint8_t)~len;
        return (uint8_t)(~len >> 8);
    }
    const uint32_t d = blk * 65535u + (o - 5u);
    const uint32_t row = d / g.line, col = d - row * g.line;
    return col == 0 ? 0 : img[(size_t)row * 4u * g.W + (col - 1u)];
}

__global__ void __launch_bounds__(PNG_THREADS)
png_encode_kernel(const uint8_t *__restrict__ imgs, int H, int W, uint8_t *__restrict__ out, int64_t out_stride)
{
    __shared__ uint32_t s_tab[256];
    __shared__ unsigned long long s_a[PNG_THREADS], s_b[PNG_THREADS];
    __shared__ uint32_t s_crc[PNG_THREADS];
    __shared__ uint32_t s_adler, s_shift;
    const int tid = threadIdx.x;
    const PngGeom g = png_geom(H, W);
    const uint8_t *img = imgs + (size_t)blockIdx.x * H * W * 4;
    uint8_t *file = out + (size_t)blockIdx.x * out_stride;
    {
        uint32_t c = (uint32_t)tid;
        for (int k = 0; k < 8; k++) c = (c & 1u) ? CRC_POLY ^ (c >> 1) : c >> 1;
        s_tab[tid] = c;
    }
    // ---- phase 1: Adler-32 over the scanline stream
    const uint32_t per = (g.n_data + PNG_THREADS - 1) / PNG_THREADS;
    {
        const uint32_t d0 = min((uint32_t)tid * per, g.n_data), d1 = min(d0 + per, g.n_data);
        unsigned long long a = 0, b = 0;                         // sum of bytes, sum of (d1 - i) * byte
        uint32_t row = d0 / g.line, col = d0 - row * g.line;
        for (uint32_t d = d0; d < d1; d++) {
            const uint32_t v = col == 0 ? 0u : img[(size_t)row * 4u * g.W + (col - 1u)];
            a += v;
            b += (unsigned long long)(d1 - d) * v;
            if (++col == g.line) { col = 0; row++; }
        }
        // bytes after this slice add a_t once each to the running b of the whole stream
        s_a[tid] = a;
        s_b[tid] = b + a * (unsigned long long)(g.n_data - d1);
    }
    __syncthreads();
    if (tid == 0) {
        unsigned long long A = 1, B = g.n_data;                  // the initial 1 of `a` is added n_data times to `b`
        for (int t = 0; t < PNG_THREADS; t++) { A += s_a[t]; B += s_b[t]; }
        s_adler = (uint32_t)((B % 65521ull) << 16) | (uint32_t)(A % 65521ull);
    }
    __syncthreads();
    const uint32_t adler = s_adler;
    // ---- phase 2: IDAT bytes + their CRC-32
    const uint32_t n_s = 4u + g.n_idat;                          // CRC covers the chunk type and the data
    const uint32_t sper = (n_s + PNG_THREADS - 1) / PNG_THREADS;
    const uint32_t i0 = min((uint32_t)tid * sper, n_s), i1 = min(i0 + sper, n_s);
    uint8_t *idat = file + 8 + 25 + 4;                           // behind the IDAT length field
    uint32_t crc = 0xffffffffu;
    for (uint32_t i = i0; i < i1; i++) {
        const uint8_t v = idat_byte(g, img, adler, i);
        idat[i] = v;
        crc = s_tab[(crc ^ v) & 0xffu] ^ (crc >> 8);
    }
    s_crc[tid] = crc ^ 0xffffffffu;                              // standard CRC-32 of the slice (0 for an empty one)
    if (tid == 1) s_shift = crc_x8n(sper);
    __syncthreads();
    if (tid == 0) {
        uint32_t c = s_crc[0];
        for (int t = 1; t < PNG_THREADS; t++) {
            const uint32_t t0 = min((uint32_t)t * sper, n_s), t1 = min(t0 + sper, n_s);
            if (t1 == t0) break;
            const uint32_t op = (t1 - t0 == sper) ? s_shift : crc_x8n(t1 - t0);
            c = crc_multmodp(op, c) ^ s_crc[t];                  // crc(A || B) = crc(A) * x^(8 |B|) + crc(B)   (zlib crc32_combine)
        }
        // ---- fixed parts: signature, IHDR, IDAT length + CRC, IEND
        const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
        for (int k = 0; k < 8; k++) file[k] = sig[k];
        uint8_t *h = file + 8;
        put_be32(h, 13);
        h[4] = 'I'; h[5] = 'H'; h[6] = 'D'; h[7] = 'R';
        put_be32(h + 8, (uint32_t)W); put_be32(h + 12, (uint32_t)H);
        h[16] = 8; h[17] = 6; h[18] = 0; h[19] = 0; h[20] = 0;    // 8 bits, RGBA, deflate, adaptive filtering, no interlace
        uint32_t hc = 0xffffffffu;
        for (int k = 4; k < 21; k++) hc = s_tab[(hc ^ h[k]) & 0xffu] ^ (hc >> 8);
        put_be32(h + 21, hc ^ 0xffffffffu);
        put_be32(file + 8 + 25, g.n_idat);
        put_be32(idat + n_s, c);
        uint8_t *e = idat + n_s + 4;
        put_be32(e, 0);
        e[4] = 'I'; e[5] = 'E'; e[6] = 'N'; e[7] = 'D';
        e[8] = 0xae; e[9] = 0x42; e[10] = 0x60; e[11] = 0x82;
    }
}

}  // namespace

extern "C" {

int64_t dp_png_size(int32_t H, int32_t W)
{
    if (H < 1 || W < 1 || H > 65535 || W > 16384) return DP_EINVAL;
    return (int64_t)png_geom(H, W).n_file;
}

int dp_png_encode(const uint8_t *img, int32_t n_images, int32_t H, int32_t W, uint8_t *out, int64_t out_stride, void *stream)
{
    if (n_images < 0 || H < 1 || W < 1 || H > 65535 || W > 16384) { dp::set_error("dp_png_encode: bad extent"); return DP_EINVAL; }
    if (n_images == 0) return DP_OK;
    if (!img || !out || out_stride < (int64_t)png_geom(H, W).n_file) { dp::set_error("dp_png_encode: null argument or stride below dp_png_size"); return DP_EINVAL; }
    png_encode_kernel<<<(unsigned)n_images, PNG_THREADS, 0, (cudaStream_t)stream>>>(img, H, W, out, out_stride);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

}  // extern "C"
