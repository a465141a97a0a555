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
// HBM-bound gather: each source byte is touched by at most two output pixels.
#pragma once

#include "common.cuh"

namespace icb {

struct ResizeParams {
    const unsigned char *src;
    unsigned char *dst;
    int w, h, pitch;
    int dw, dh, dpitch;
    long long dxl, dyl, xbase, ybase;      // 32.32 fixed point
};

__device__ __forceinline__ int mul8(int a, int b)       // AlphaMath.c mul8table[a][b]
{
    return (int)(((unsigned)b * ((unsigned)a * 0x10101u) + (1u << 23)) >> 24);
}

__device__ __forceinline__ int div8(int a, int b)       // AlphaMath.c div8table[a][b], a >= 1
{
    if (b >= a) return 255;
    const unsigned inc = ((0xffu << 24) + (unsigned)a / 2u) / (unsigned)a;
    return (int)(((unsigned)b * inc + (1u << 23)) >> 24);
}

template <int BPP, int AOFF>
__global__ void __launch_bounds__(256) k_resize(const ResizeParams p)
{
    const int dx = blockIdx.x * 32 + (threadIdx.x & 31);
    const int dy = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (dx >= p.dw || dy >= p.dh) return;
    const long long half = 1ll << 31;
    const long long yl = p.ybase + (long long)dy * p.dyl - half;
    const long long xl = p.xbase + (long long)dx * p.dxl - half;
    const int yw = (int)(yl >> 32), xw = (int)(xl >> 32);
    const int yf = (int)(((unsigned)yl) >> 24), xf = (int)(((unsigned)xl) >> 24);
    int y0 = yw < 0 ? 0 : yw;
    int y1 = (yw + 1 < p.h) ? (yw < 0 ? 0 : yw + 1) : y0;
    if (y0 > p.h - 1) y0 = y1 = p.h - 1;
    int x0 = xw < 0 ? 0 : xw;
    int x1 = (xw + 1 < p.w) ? (xw < 0 ? 0 : xw + 1) : x0;
    if (x0 > p.w - 1) x0 = x1 = p.w - 1;
    const unsigned char *r0 = p.src + (size_t)y0 * p.pitch, *r1 = p.src + (size_t)y1 * p.pitch;
    const unsigned char *p00 = r0 + (size_t)x0 * BPP, *p01 = r0 + (size_t)x1 * BPP;
    const unsigned char *p10 = r1 + (size_t)x0 * BPP, *p11 = r1 + (size_t)x1 * BPP;
    int a00 = 255, a01 = 255, a10 = 255, a11 = 255;
    if (AOFF >= 0) {
        a00 = __ldg(p00 + AOFF); a01 = __ldg(p01 + AOFF);
        a10 = __ldg(p10 + AOFF); a11 = __ldg(p11 + AOFF);
    }
    int res[BPP];
#pragma unroll
    for (int c = 0; c < BPP; c++) {
        int c00 = __ldg(p00 + c), c01 = __ldg(p01 + c), c10 = __ldg(p10 + c), c11 = __ldg(p11 + c);
        if (AOFF >= 0 && c != AOFF) {
            c00 = mul8(a00, c00); c01 = mul8(a01, c01); c10 = mul8(a10, c10); c11 = mul8(a11, c11);
        }
        const int top = (c00 << 8) + (c01 - c00) * xf;
        const int bot = (c10 << 8) + (c11 - c10) * xf;
        const int v = (top << 8) + (bot - top) * yf;
        res[c] = (v + (1 << 15)) >> 16;
    }
    unsigned char *o = p.dst + (size_t)dy * p.dpitch + (size_t)dx * BPP;
    if (AOFF >= 0) {
        const int a = res[AOFF];
#pragma unroll
        for (int c = 0; c < BPP; c++) {
            int v;
            if (c == AOFF) v = a;
            else if (a == 0) v = 0;
            else if (a == 255) v = res[c];
            else v = div8(a, res[c]);
            o[c] = (unsigned char)v;
        }
    } else {
#pragma unroll
        for (int c = 0; c < BPP; c++) o[c] = (unsigned char)res[c];
    }
}

// f2 (input staging): ImageReadParam.setSourceSubsampling(s, s, 0, 0) as the reference's decode rule
// applies it (core/ImageCompression.java:140-153) -- source pixel (x*s, y*s) for every destination
// pixel, point sampling.  One thread per destination pixel; HBM-bound gather (a 32-byte sector is
// fetched for every BPP bytes kept when s*BPP >= 32).
template <int BPP>
__global__ void __launch_bounds__(256) k_subsample(const unsigned char *__restrict__ src, int pitch, int s,
                                                   unsigned char *__restrict__ dst, int dw, int dh, int dpitch)
{
    const int dx = blockIdx.x * 32 + (threadIdx.x & 31);
    const int dy = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (dx >= dw || dy >= dh) return;
    const unsigned char *p = src + (size_t)dy * s * pitch + (size_t)dx * s * BPP;
    unsigned char *o = dst + (size_t)dy * dpitch + (size_t)dx * BPP;
#pragma unroll
    for (int c = 0; c < BPP; c++) o[c] = __ldg(p + c);
}

}  // namespace icb
