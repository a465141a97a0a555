// k_resize.cuh -- K4: bilinear downscale, the device form of
// tools/ImageTools.java:7-26 resizeImage (Graphics2D.drawImage with
// VALUE_INTERPOLATION_BILINEAR).  Used by the PNG path
// (core/ImageCompressionPng.java:66) and by the JPG scale loop / cached scale
// (core/ImageCompressionJpg.java:99, :222).
//
// Arithmetic = Java2D's software loop as published in OpenJDK's libawt
// (TransformHelper.c, LoopMacros.h bilinear helper, AlphaMath.c) [JDK-mem]:
// 32.32 fixed-point walk from the destination pixel centre minus one half, 2x2
// taps clamped to the image, 8-bit fractions, ((..) + (1<<15)) >> 16; pixels with
// alpha go through IntArgbPre (mul8) and back through SrcOver (div8).  The host
// computes the four fixed-point constants in double exactly like the JDK does.
//
// HBM-bound gather, laid out the way the north-star asks: one CTA owns a tile of TW x 8
// destination pixels.  The source rows its taps touch (two per destination row) are staged in
// shared memory as whole 16-byte vectors (128-bit coalesced loads of the contiguous row segment
// the tile's x range covers), the taps are read from shared memory as whole pixels, two colour
// channels are interpolated per 32-bit multiply (16-bit lanes: every partial sum is < 2^16), and
// the tile leaves as full-sector stores (32 consecutive 4-byte pixels per warp store; 3- and 1-byte pixels
// through a shared-memory row buffer and 128-bit stores).
#pragma once

#include "common.cuh"

namespace icb {

struct ResizeParams {
    const unsigned char *src;
    unsigned char *dst;
    int w, h, pitch;
    int dw, dh, dpitch;
    long long dxl, dyl, xbase, ybase;      // 32.32 fixed point
    int tw_log2;                           // destination pixels per CTA row = 1 << tw_log2 (4 .. 128)
    int seg_bytes;                         // shared-memory bytes per staged source row (multiple of 16)
    int tiles_x, ntiles;                   // destination tiles per row of tiles, and in all
};

constexpr int RS_ROWS = 8;                 // destination rows per CTA
constexpr int RS_THREADS = 256;
constexpr int RS_MAX_SEG = 2560;           // staged bytes per source row: 2 buffers x 16 rows x 2560 B = 80 KB at most
constexpr int RS_MAX_TW = 128;

__host__ __device__ inline int resize_smem_bytes(int seg_bytes, int tw, int bpp)
{
    return 2 * (2 * RS_ROWS * seg_bytes) + (bpp == 4 ? 0 : RS_ROWS * ((tw * bpp + 15) & ~15)) + 1024;     // two staging buffers
}

__device__ __forceinline__ int mul8(int a, int b)       // AlphaMath.c mul8table[a][b]
{
    return (int)(((unsigned)b * ((unsigned)a * 0x10101u) + (1u << 23)) >> 24);
}

__device__ __forceinline__ unsigned div8_inc(int a)     // AlphaMath.c: the per-row increment of div8table[a][.], a >= 1
{
    return ((0xffu << 24) + (unsigned)a / 2u) / (unsigned)a;
}

// the two taps and the 8-bit fraction of one axis (TransformHelper.c: whole part clamped to the image)
__device__ __forceinline__ void rs_taps(long long l, int n, int &i0, int &i1, int &f)
{
    const int iw = (int)(l >> 32);
    f = (int)(((unsigned)l) >> 24);
    i0 = iw < 0 ? 0 : iw;
    i1 = (iw + 1 < n) ? (iw < 0 ? 0 : iw + 1) : i0;
    if (i0 > n - 1) i0 = i1 = n - 1;
}

// One pixel (BPP bytes, in the low bytes of the result) at byte offset `off` of a staged row.
template <int BPP>
__device__ __forceinline__ uint32_t rs_px(const unsigned char *row, int off)
{
    if (BPP == 4) return *reinterpret_cast<const uint32_t *>(row + off);
    if (BPP == 1) return row[off];
    const uint32_t *w = reinterpret_cast<const uint32_t *>(row + (off & ~3));
    return __funnelshift_r(w[0], w[1], 8 * (off & 3)) & 0xffffffu;
}

// Premultiply one tap by its alpha (IntArgb -> IntArgbPre), two channels per multiply: e / o hold channels
// 0, 2 / 1, 3 in 16-bit lanes.  AlphaMath.c mul8table[a][c] = (c * a * 0x10101 + 2^23) >> 24 equals
// (t + (t >> 8)) >> 8 with t = a * c + 128 for every a, c in 0..255 (checked exhaustively), and neither t nor
// t + (t >> 8) leaves 16 bits, so both lanes go through one 32-bit multiply.  The alpha lane keeps alpha itself.
template <int AOFF>
__device__ __forceinline__ void rs_premul2(uint32_t p, uint32_t &e, uint32_t &o)
{
    const uint32_t a = (p >> (8 * AOFF)) & 0xffu;
    uint32_t te = e * a + 0x00800080u, to = o * a + 0x00800080u;
    te = ((te + ((te >> 8) & 0x00ff00ffu)) >> 8) & 0x00ff00ffu;
    to = ((to + ((to >> 8) & 0x00ff00ffu)) >> 8) & 0x00ff00ffu;
    if (AOFF == 0) te = (te & 0xffff0000u) | a;
    else if (AOFF == 2) te = (te & 0x0000ffffu) | (a << 16);
    else if (AOFF == 1) to = (to & 0xffff0000u) | a;
    else to = (to & 0x0000ffffu) | (a << 16);
    e = te; o = to;
}

__device__ __forceinline__ void rs_cp_async16(void *smem_dst, const void *gmem_src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src)
                 : "memory");
}

// geometry of destination tile t: its first pixel, its size, and the staged byte range of the source rows
struct RsTile {
    int dx_first, dy_first, ncols, nrows, byte0, nvec;
};

template <int BPP>
__device__ __forceinline__ RsTile rs_tile(const ResizeParams &p, int t)
{
    const long long half = 1ll << 31;
    const int tw = 1 << p.tw_log2;
    RsTile g;
    const int ty = t / p.tiles_x, tx = t - ty * p.tiles_x;
    g.dx_first = tx * tw; g.dy_first = ty * RS_ROWS;
    g.ncols = min(tw, p.dw - g.dx_first); g.nrows = min(RS_ROWS, p.dh - g.dy_first);
    int xa, xb, f_unused;
    rs_taps(p.xbase + (long long)g.dx_first * p.dxl - half, p.w, xa, xb, f_unused);
    const int xs0 = xa;
    rs_taps(p.xbase + (long long)(g.dx_first + g.ncols - 1) * p.dxl - half, p.w, xa, xb, f_unused);
    const int xs1 = max(xa, xb);
    g.byte0 = (xs0 * BPP) & ~15;
    g.nvec = min(p.seg_bytes, (((xs1 + 1) * BPP + 15) & ~15) - g.byte0) >> 4;     // (the host sized seg_bytes for it)
    return g;
}

// Persistent CTAs, two staging buffers: while a tile is interpolated out of one buffer, the source rows of the
// CTA's next tile stream into the other (cp.async, 16 bytes per request, no registers in between).
template <int BPP, int AOFF>
__global__ void __launch_bounds__(RS_THREADS) k_resize(const ResizeParams p)
{
    extern __shared__ __align__(16) unsigned char rs_smem[];
    const int tid = threadIdx.x;
    const int tw = 1 << p.tw_log2;
    const int stage_bytes = 2 * RS_ROWS * p.seg_bytes;                      // one buffer: [2 * RS_ROWS][seg_bytes]
    unsigned char *s_out = rs_smem + 2 * stage_bytes;                       // [RS_ROWS][out_stride]   (BPP != 4)
    const int out_stride = BPP == 4 ? 0 : (tw * BPP + 15) & ~15;
    unsigned *s_inc = reinterpret_cast<unsigned *>(s_out + RS_ROWS * out_stride);   // div8 increments
    const long long half = 1ll << 31;

    if (AOFF >= 0) s_inc[tid] = tid ? div8_inc(tid) : 0u;

    // rows y0 / y1 of every destination row of tile t, bytes [byte0, byte0 + 16 nvec): warp w stages the two
    // source rows of destination row w, lanes walk the row's vectors.  The tile's geometry and the warp's row
    // fraction are returned: the same tile is interpolated one iteration later, nothing is computed twice.
    auto stage = [&](int t, int buf, RsTile &g, int &yf) {
        g = rs_tile<BPP>(p, t);
        const int lane = tid & 31, rr = tid >> 5;
        yf = 0;
        if (rr < g.nrows) {
            int y0, y1;
            rs_taps(p.ybase + (long long)(g.dy_first + rr) * p.dyl - half, p.h, y0, y1, yf);
            const unsigned char *g0 = p.src + (size_t)y0 * p.pitch + g.byte0, *g1 = p.src + (size_t)y1 * p.pitch + g.byte0;
            unsigned char *s0 = rs_smem + buf * stage_bytes + (2 * rr) * p.seg_bytes, *s1 = s0 + p.seg_bytes;
            for (int v = lane; v < g.nvec; v += 32) {
                rs_cp_async16(s0 + 16 * v, g0 + 16 * v);
                rs_cp_async16(s1 + 16 * v, g1 + 16 * v);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    RsTile tg_next = {};
    int yf_next = 0;
    if ((int)blockIdx.x < p.ntiles) stage(blockIdx.x, 0, tg_next, yf_next);
    int k = 0;
    for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x, k++) {
    const int buf = k & 1;
    const RsTile tg = tg_next;
    const int yf = yf_next;
    if (t + (int)gridDim.x < p.ntiles) {
        stage(t + gridDim.x, buf ^ 1, tg_next, yf_next);
        asm volatile("cp.async.wait_group 1;" ::: "memory");               // everything but the prefetch has landed
    } else {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const int dx_first = tg.dx_first, dy_first = tg.dy_first, ncols = tg.ncols, nrows = tg.nrows, byte0 = tg.byte0;
    const unsigned char *s_in = rs_smem + buf * stage_bytes;

    // ---- interpolate: warp rr = destination row rr of the tile, lane l = pixels l, l + 32, ... of that row
    // (neighbouring lanes read neighbouring taps: shared-memory reads are conflict-free or two-way, and every
    // store instruction of the warp covers one contiguous stretch of the row)
    {
        const int lane = tid & 31, rr = tid >> 5;
        if (rr < nrows) {
            const int dy = dy_first + rr;
            const unsigned char *r0 = s_in + (2 * rr) * p.seg_bytes - byte0, *r1 = r0 + p.seg_bytes;
            const uint32_t wy1 = (uint32_t)yf, wy0 = 256u - wy1;
            const long long dx32 = p.dxl * 32;
            long long xl = p.xbase + (long long)(dx_first + lane) * p.dxl - half;
            for (int j = lane; j < ncols; j += 32, xl += dx32) {
                const int xw = (int)(xl >> 32);
                const uint32_t wx1 = ((uint32_t)xl) >> 24, wx0 = 256u - wx1;
                const int x0 = min(max(xw, 0), p.w - 1), x1 = min(max(xw + 1, 0), p.w - 1);     // == rs_taps, branch-free
                const uint32_t p00 = rs_px<BPP>(r0, x0 * BPP), p01 = rs_px<BPP>(r0, x1 * BPP);
                const uint32_t p10 = rs_px<BPP>(r1, x0 * BPP), p11 = rs_px<BPP>(r1, x1 * BPP);
                // channels 0 / 2 and 1 / 3 of each tap in the 16-bit lanes of one word each
                uint32_t e00 = p00 & 0x00ff00ffu, e01 = p01 & 0x00ff00ffu, e10 = p10 & 0x00ff00ffu, e11 = p11 & 0x00ff00ffu;
                uint32_t o00 = 0, o01 = 0, o10 = 0, o11 = 0;
                if (BPP >= 3) {
                    o00 = (p00 >> 8) & 0x00ff00ffu; o01 = (p01 >> 8) & 0x00ff00ffu;
                    o10 = (p10 >> 8) & 0x00ff00ffu; o11 = (p11 >> 8) & 0x00ff00ffu;
                }
                bool opaque = true;
                if (AOFF >= 0) {
                    constexpr int A = AOFF < 0 ? 0 : AOFF;
                    opaque = (((p00 & p01 & p10 & p11) >> (8 * A)) & 0xffu) == 0xffu;
                    if (!opaque) {                                  // IntArgb -> IntArgbPre, every tap by its own alpha
                        rs_premul2<A>(p00, e00, o00); rs_premul2<A>(p01, e01, o01);
                        rs_premul2<A>(p10, e10, o10); rs_premul2<A>(p11, e11, o11);
                    }
                }
                // horizontal pass, two channels per multiply: c00 * (256 - xf) + c01 * xf == (c00 << 8) + (c01 - c00) * xf
                const uint32_t te = e00 * wx0 + e01 * wx1, be = e10 * wx0 + e11 * wx1;
                uint32_t v0 = ((te & 0xffffu) * wy0 + (be & 0xffffu) * wy1 + (1u << 15)) >> 16;
                uint32_t v1 = 0, v2 = 0, v3 = 0;
                if (BPP >= 3) {
                    const uint32_t to = o00 * wx0 + o01 * wx1, bo = o10 * wx0 + o11 * wx1;
                    v1 = ((to & 0xffffu) * wy0 + (bo & 0xffffu) * wy1 + (1u << 15)) >> 16;
                    v2 = ((te >> 16) * wy0 + (be >> 16) * wy1 + (1u << 15)) >> 16;
                    if (BPP == 4) v3 = ((to >> 16) * wy0 + (bo >> 16) * wy1 + (1u << 15)) >> 16;
                }
                if (AOFF >= 0 && !opaque) {                         // back from premultiplied (SrcOver into a cleared destination)
                    uint32_t vv[4] = {v0, v1, v2, v3};
                    const uint32_t a = vv[AOFF < 0 ? 0 : AOFF];
                    const uint32_t inc = s_inc[a];
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        if (c == AOFF) continue;
                        vv[c] = a == 0u ? 0u : (a == 255u ? vv[c] : (vv[c] >= a ? 255u : (vv[c] * inc + (1u << 23)) >> 24));
                    }
                    v0 = vv[0]; v1 = vv[1]; v2 = vv[2]; v3 = vv[3];
                }
                if (BPP == 4) {
                    *reinterpret_cast<uint32_t *>(p.dst + (size_t)dy * p.dpitch + (size_t)(dx_first + j) * 4) =
                        v0 | (v1 << 8) | (v2 << 16) | (v3 << 24);
                } else if (BPP == 3) {
                    unsigned char *o = s_out + rr * out_stride + 3 * j;
                    o[0] = (unsigned char)v0; o[1] = (unsigned char)v1; o[2] = (unsigned char)v2;
                } else {
                    s_out[rr * out_stride + j] = (unsigned char)v0;
                }
            }
        }
    }
    if (BPP != 4) {
        __syncthreads();
        // the tile's rows leave as 16-byte vectors (tile x offset and row pitch are multiples of 16 bytes);
        // the ragged end of the image's last tile goes out in bytes
        const int row_bytes = ncols * BPP, nv = row_bytes >> 4;
        const size_t xoff = (size_t)dx_first * BPP;
        for (int i = tid; i < nrows * nv; i += RS_THREADS) {
            const int rr = i / nv, v = i - rr * nv;
            reinterpret_cast<uint4 *>(p.dst + (size_t)(dy_first + rr) * p.dpitch + xoff)[v] =
                reinterpret_cast<const uint4 *>(s_out + rr * out_stride)[v];
        }
        const int tail = row_bytes & 15;
        for (int i = tid; i < nrows * tail; i += RS_THREADS) {
            const int rr = i / tail, b = (nv << 4) + (i - rr * tail);
            p.dst[(size_t)(dy_first + rr) * p.dpitch + xoff + b] = s_out[rr * out_stride + b];
        }
    }
    __syncthreads();                            // the buffer just read is the next prefetch target
    }
}

// Fallback for extreme downscales (more source pixels per destination pixel than a staged row segment holds):
// one thread per destination pixel, taps read straight from global memory.  Same arithmetic.
template <int BPP, int AOFF>
__global__ void __launch_bounds__(256) k_resize_gather(const ResizeParams p)
{
    const int dx = blockIdx.x * 32 + (threadIdx.x & 31);
    const int dy = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (dx >= p.dw || dy >= p.dh) return;
    const long long half = 1ll << 31;
    int y0, y1, yf, x0, x1, xf;
    rs_taps(p.ybase + (long long)dy * p.dyl - half, p.h, y0, y1, yf);
    rs_taps(p.xbase + (long long)dx * p.dxl - half, p.w, x0, x1, xf);
    const unsigned char *r0 = p.src + (size_t)y0 * p.pitch, *r1 = p.src + (size_t)y1 * p.pitch;
    const unsigned char *p00 = r0 + (size_t)x0 * BPP, *p01 = r0 + (size_t)x1 * BPP;
    const unsigned char *p10 = r1 + (size_t)x0 * BPP, *p11 = r1 + (size_t)x1 * BPP;
    int a00 = 255, a01 = 255, a10 = 255, a11 = 255;
    if (AOFF >= 0) {
        a00 = p00[AOFF < 0 ? 0 : AOFF]; a01 = p01[AOFF < 0 ? 0 : AOFF];
        a10 = p10[AOFF < 0 ? 0 : AOFF]; a11 = p11[AOFF < 0 ? 0 : AOFF];
    }
    int res[BPP];
#pragma unroll
    for (int c = 0; c < BPP; c++) {
        int c00 = p00[c], c01 = p01[c], c10 = p10[c], c11 = p11[c];
        if (AOFF >= 0 && c != AOFF) {
            c00 = mul8(a00, c00); c01 = mul8(a01, c01); c10 = mul8(a10, c10); c11 = mul8(a11, c11);
        }
        const int top = (c00 << 8) + (c01 - c00) * xf;
        const int bot = (c10 << 8) + (c11 - c10) * xf;
        res[c] = ((top << 8) + (bot - top) * yf + (1 << 15)) >> 16;
    }
    unsigned char *o = p.dst + (size_t)dy * p.dpitch + (size_t)dx * BPP;
#pragma unroll
    for (int c = 0; c < BPP; c++) {
        int v = res[c];
        if (AOFF >= 0 && c != AOFF) {
            const int a = res[AOFF < 0 ? 0 : AOFF];
            if (a == 0) v = 0;
            else if (a != 255) v = v >= a ? 255 : (int)(((unsigned)v * div8_inc(a) + (1u << 23)) >> 24);
        }
        o[c] = (unsigned char)v;
    }
}

// f2 (input staging): ImageReadParam.setSourceSubsampling(s, s, 0, 0) as the reference's decode rule
// applies it (core/ImageCompression.java:140-153) -- source pixel (x*s, y*s) for every destination
// pixel, point sampling.  One thread = four destination pixels (one 16-byte store for 4-byte pixels,
// three words for 3-byte pixels); 4-byte source pixels are read as words.  HBM-bound gather: a
// 32-byte sector is fetched for every BPP bytes kept when s*BPP >= 32.
template <int BPP>
__global__ void __launch_bounds__(256) k_subsample(const unsigned char *__restrict__ src, int pitch, int s,
                                                   unsigned char *__restrict__ dst, int dw, int dh, int dpitch)
{
    const int dx0 = (blockIdx.x * 32 + (threadIdx.x & 31)) * 4;
    const int dy = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (dx0 >= dw || dy >= dh) return;
    const unsigned char *row = src + (size_t)dy * s * pitch;
    uint32_t px[4] = {0, 0, 0, 0};
    const int n = min(4, dw - dx0);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (i >= n) break;
        const unsigned char *q = row + (size_t)(dx0 + i) * s * BPP;
        if (BPP == 4) {
            px[i] = __ldg(reinterpret_cast<const uint32_t *>(q));
        } else {
#pragma unroll
            for (int c = 0; c < BPP; c++) px[i] |= (uint32_t)__ldg(q + c) << (8 * c);
        }
    }
    unsigned char *o = dst + (size_t)dy * dpitch + (size_t)dx0 * BPP;
    if (n == 4 && BPP == 4) {
        *reinterpret_cast<uint4 *>(o) = make_uint4(px[0], px[1], px[2], px[3]);
    } else if (n == 4 && BPP == 3) {
        uint32_t *o32 = reinterpret_cast<uint32_t *>(o);
        o32[0] = px[0] | (px[1] << 24);
        o32[1] = (px[1] >> 8) | (px[2] << 16);
        o32[2] = (px[2] >> 16) | (px[3] << 8);
    } else {
        for (int i = 0; i < n; i++)
            for (int c = 0; c < BPP; c++) o[i * BPP + c] = (unsigned char)(px[i] >> (8 * c));
    }
}

}  // namespace icb
