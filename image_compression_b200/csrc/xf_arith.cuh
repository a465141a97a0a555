// xf_arith.cuh -- the exact-integer arithmetic of the transform kernel (colour conversion by dp2a,
// row pass on sample bytes, column pass), kept apart from the kernel so that the same code compiles
// for the host: tests/test_xf_arith.py builds tests/xf_arith_host.cu with nvcc and checks these
// forms against the CPU oracle without a GPU (dp2a / dp4a / prmt are emulated on the host).
#pragma once

#include <stdint.h>

#ifndef __CUDACC__
#define __host__
#define __device__
#define __forceinline__ inline
struct alignas(8) uint2 { unsigned int x, y; };
#endif
#define ICB_HD __host__ __device__ __forceinline__

namespace icb {

// jfdctint.c constants (CONST_BITS 13)
#define ICB_FIX_0_298631336 2446
#define ICB_FIX_0_390180644 3196
#define ICB_FIX_0_541196100 4433
#define ICB_FIX_0_765366865 6270
#define ICB_FIX_0_899976223 7373
#define ICB_FIX_1_175875602 9633
#define ICB_FIX_1_501321110 12299
#define ICB_FIX_1_847759065 15137
#define ICB_FIX_1_961570560 16069
#define ICB_FIX_2_053119869 16819
#define ICB_FIX_2_562915447 20995
#define ICB_FIX_3_072711026 25172

// ---- integer dot products (sm_61+): a = 16-bit pairs / signed bytes (coefficients), b = bytes ----
template <bool HI, bool SIGNED_A>
ICB_HD int dp2a(uint32_t a, uint32_t b, int c)
{
#ifdef __CUDA_ARCH__
    int d;
    if (HI) {
        if (SIGNED_A) asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
        else asm("dp2a.hi.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    } else {
        if (SIGNED_A) asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
        else asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    }
    return d;
#else       // PTX ISA dp2a: d = c + a.h0 * b.byte[2 HI] + a.h1 * b.byte[2 HI + 1]
    const int64_t a0 = SIGNED_A ? (int64_t)(int16_t)(a & 0xffffu) : (int64_t)(a & 0xffffu);
    const int64_t a1 = SIGNED_A ? (int64_t)(int16_t)(a >> 16) : (int64_t)(a >> 16);
    const int64_t b0 = (b >> (HI ? 16 : 0)) & 0xffu, b1 = (b >> (HI ? 24 : 8)) & 0xffu;
    return (int)(uint32_t)((int64_t)c + a0 * b0 + a1 * b1);
#endif
}
ICB_HD int dp4a_su(uint32_t a, uint32_t b, int c)
{
#ifdef __CUDA_ARCH__
    int d;
    asm("dp4a.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#else
    int64_t acc = c;
    for (int i = 0; i < 4; i++) acc += (int64_t)(int8_t)((a >> (8 * i)) & 0xffu) * (int64_t)((b >> (8 * i)) & 0xffu);
    return (int)(uint32_t)acc;
#endif
}
// PRMT, default mode: result byte i = byte (selector nibble i) of the eight bytes {y, x}
ICB_HD uint32_t prmt(uint32_t x, uint32_t y, uint32_t sel)
{
#ifdef __CUDA_ARCH__
    return __byte_perm(x, y, sel);
#else
    const uint64_t v = ((uint64_t)y << 32) | x;
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7))) & 0xffu) << (8 * i);
    return r;
#endif
}
ICB_HD constexpr uint32_t pair16(int lo, int hi)
{
    return ((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16);
}

// Row pass (jfdctint.c pass 1) of one 8-sample row held as bytes: wa = x0..x3, wb = x4..x7.
// Keeps PASS1_BITS = 2 extra bits and removes the +128 level shift (only output 0 sees it, by
// exactly 8 * 128 << 2).  Each output before its descale is sum_j M[k][j] * x_j with
//   M[k][7 - j] = M[k][j] (k even), -M[k][j] (k odd), all |M| <= 11363.
template <int K>
ICB_HD int fdct_row_k(uint32_t wa, uint32_t wb)        // output K of the row pass
{
    constexpr int RND = 1 << 10;
    // even half: d2 = t13 * (c541 + c765) + t12 * c541,  d6 = t13 * c541 + t12 * (c541 - c1847)
    constexpr int A2 = ICB_FIX_0_541196100 + ICB_FIX_0_765366865, B2 = ICB_FIX_0_541196100;
    constexpr int A6 = ICB_FIX_0_541196100, B6 = ICB_FIX_0_541196100 - ICB_FIX_1_847759065;
    // odd half over (t7, t6, t5, t4) = (x0 - x7, x1 - x6, x2 - x5, x3 - x4)
    constexpr int Z = ICB_FIX_1_175875602;
    constexpr int O1[4] = {ICB_FIX_1_501321110 - ICB_FIX_0_899976223 + Z - ICB_FIX_0_390180644, Z,
                           Z - ICB_FIX_0_390180644, Z - ICB_FIX_0_899976223};
    constexpr int O3[4] = {Z, ICB_FIX_3_072711026 - ICB_FIX_2_562915447 + Z - ICB_FIX_1_961570560,
                           Z - ICB_FIX_2_562915447, Z - ICB_FIX_1_961570560};
    constexpr int O5[4] = {Z - ICB_FIX_0_390180644, Z - ICB_FIX_2_562915447,
                           ICB_FIX_2_053119869 - ICB_FIX_2_562915447 + Z - ICB_FIX_0_390180644, Z};
    constexpr int O7[4] = {Z - ICB_FIX_0_899976223, Z - ICB_FIX_1_961570560, Z,
                           ICB_FIX_0_298631336 - ICB_FIX_0_899976223 + Z - ICB_FIX_1_961570560};
#define ICB_ROW4(c0, c1, c2, c3, c4, c5, c6, c7)                                                       \
    dp2a<true, true>(pair16(c6, c7), wb,                                                               \
                     dp2a<false, true>(pair16(c4, c5), wb,                                             \
                                       dp2a<true, true>(pair16(c2, c3), wa, dp2a<false, true>(pair16(c0, c1), wa, RND))))
    if (K == 0) return dp4a_su(0x04040404u, wa, dp4a_su(0x04040404u, wb, -4096));
    if (K == 4) return dp4a_su(0x04fcfc04u, wa, dp4a_su(0x04fcfc04u, wb, 0));
    if (K == 2) return ICB_ROW4(A2, B2, -B2, -A2, -A2, -B2, B2, A2) >> 11;
    if (K == 6) return ICB_ROW4(A6, B6, -B6, -A6, -A6, -B6, B6, A6) >> 11;
    if (K == 1) return ICB_ROW4(O1[0], O1[1], O1[2], O1[3], -O1[3], -O1[2], -O1[1], -O1[0]) >> 11;
    if (K == 3) return ICB_ROW4(O3[0], O3[1], O3[2], O3[3], -O3[3], -O3[2], -O3[1], -O3[0]) >> 11;
    if (K == 5) return ICB_ROW4(O5[0], O5[1], O5[2], O5[3], -O5[3], -O5[2], -O5[1], -O5[0]) >> 11;
    return ICB_ROW4(O7[0], O7[1], O7[2], O7[3], -O7[3], -O7[2], -O7[1], -O7[0]) >> 11;
#undef ICB_ROW4
}

ICB_HD void fdct_row_bytes(uint32_t wa, uint32_t wb, int *d)
{
    d[0] = fdct_row_k<0>(wa, wb); d[1] = fdct_row_k<1>(wa, wb); d[2] = fdct_row_k<2>(wa, wb); d[3] = fdct_row_k<3>(wa, wb);
    d[4] = fdct_row_k<4>(wa, wb); d[5] = fdct_row_k<5>(wa, wb); d[6] = fdct_row_k<6>(wa, wb); d[7] = fdct_row_k<7>(wa, wb);
}

// Column pass (jfdctint.c pass 2): descales by CONST_BITS + PASS1_BITS = 15, IJG's butterflies.
// Every output is returned SHIFTED LEFT BY 16 (plus junk below): all constants are doubled, so the
// bits IJG's ">> 15" would keep land in the upper half-word and the int16 pack that follows picks
// them with the same PRMT it needs anyway -- no shift instructions.  (Wrap-around of the doubled
// sum is harmless: only bits 16..31 of it are used and the coefficients fit 16 bits.)
ICB_HD void fdct_col_hi16(int &d0, int &d1, int &d2, int &d3, int &d4, int &d5, int &d6, int &d7)
{
    constexpr int RND2 = 1 << 15;       // 2 * (1 << 14)
    constexpr int Z = 2 * ICB_FIX_1_175875602;
    constexpr int C541 = 2 * ICB_FIX_0_541196100, C765 = 2 * ICB_FIX_0_765366865, C1847 = 2 * ICB_FIX_1_847759065;
    constexpr int C298 = 2 * ICB_FIX_0_298631336, C390 = 2 * ICB_FIX_0_390180644, C899 = 2 * ICB_FIX_0_899976223;
    constexpr int C1501 = 2 * ICB_FIX_1_501321110, C1961 = 2 * ICB_FIX_1_961570560, C2053 = 2 * ICB_FIX_2_053119869;
    constexpr int C2562 = 2 * ICB_FIX_2_562915447, C3072 = 2 * ICB_FIX_3_072711026;
    const int tmp0 = d0 + d7, tmp7 = d0 - d7;
    const int tmp1 = d1 + d6, tmp6 = d1 - d6;
    const int tmp2 = d2 + d5, tmp5 = d2 - d5;
    const int tmp3 = d3 + d4, tmp4 = d3 - d4;
    const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3;
    const int tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    d0 = (tmp10 + tmp11) * 16384 + 32768;       // ((x + 2) >> 2) << 16
    d4 = (tmp10 - tmp11) * 16384 + 32768;
#ifndef ICB_XF_COLEXPANDED  // IJG's own factoring: 12 multiplies, more additions.  Measured 2-4 % faster than the expanded
                            // form below (19 multiplies, fewer additions): the fma-heavy pipe is the kernel's bound
    const int z1e = (tmp12 + tmp13) * C541 + RND2;
    d2 = z1e + tmp13 * C765;
    d6 = z1e - tmp12 * C1847;
    const int z3s = tmp4 + tmp6, z4s = tmp5 + tmp7;
    const int z5 = (z3s + z4s) * Z + RND2;
    const int z3 = z5 - z3s * C1961, z4 = z5 - z4s * C390;
    const int z1 = (tmp4 + tmp7) * (-C899), z2 = (tmp5 + tmp6) * (-C2562);
    d7 = tmp4 * C298 + z1 + z3;
    d5 = tmp5 * C2053 + z2 + z4;
    d3 = tmp6 * C3072 + z2 + z3;
    d1 = tmp7 * C1501 + z1 + z4;
#else
    d2 = tmp13 * (C541 + C765) + tmp12 * C541 + RND2;
    d6 = tmp13 * C541 + tmp12 * (C541 - C1847) + RND2;
    const int z5 = (tmp4 + tmp5 + tmp6 + tmp7) * Z + RND2;
    d7 = z5 + tmp4 * (C298 - C899 - C1961) - tmp7 * C899 - tmp6 * C1961;
    d5 = z5 + tmp5 * (C2053 - C2562 - C390) - tmp6 * C2562 - tmp7 * C390;
    d3 = z5 + tmp6 * (C3072 - C2562 - C1961) - tmp5 * C2562 - tmp4 * C1961;
    d1 = z5 + tmp7 * (C1501 - C899 - C390) - tmp4 * C899 - tmp5 * C390;
#endif
}

// Whole block, one column of the intermediate at a time: output K of all eight rows straight from the
// sample bytes (sw[2r], sw[2r+1] = row r), then that column's pass, then its eight results packed in
// pairs of rows -- P[4 K + j] = (element (2j, K), element (2j+1, K)) as int16 halves.  At no point
// are more than one column of 32-bit intermediates alive (the row-major form keeps all 64).
template <int K>
ICB_HD void fdct_block_col(const uint32_t (&sw)[16], uint32_t (&P)[32])
{
    int v0 = fdct_row_k<K>(sw[0], sw[1]), v1 = fdct_row_k<K>(sw[2], sw[3]);
    int v2 = fdct_row_k<K>(sw[4], sw[5]), v3 = fdct_row_k<K>(sw[6], sw[7]);
    int v4 = fdct_row_k<K>(sw[8], sw[9]), v5 = fdct_row_k<K>(sw[10], sw[11]);
    int v6 = fdct_row_k<K>(sw[12], sw[13]), v7 = fdct_row_k<K>(sw[14], sw[15]);
    fdct_col_hi16(v0, v1, v2, v3, v4, v5, v6, v7);
    P[4 * K + 0] = prmt((uint32_t)v0, (uint32_t)v1, 0x7632);
    P[4 * K + 1] = prmt((uint32_t)v2, (uint32_t)v3, 0x7632);
    P[4 * K + 2] = prmt((uint32_t)v4, (uint32_t)v5, 0x7632);
    P[4 * K + 3] = prmt((uint32_t)v6, (uint32_t)v7, 0x7632);
}
ICB_HD void fdct_block_bytes(const uint32_t (&sw)[16], uint32_t (&P)[32])
{
    fdct_block_col<0>(sw, P); fdct_block_col<1>(sw, P); fdct_block_col<2>(sw, P); fdct_block_col<3>(sw, P);
    fdct_block_col<4>(sw, P); fdct_block_col<5>(sw, P); fdct_block_col<6>(sw, P); fdct_block_col<7>(sw, P);
}
// zigzag index -> natural index (for constant expressions only)
constexpr unsigned char kZZc[64] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5,
    12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
    58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
// int16 pair (natural index NA in the low half, NB in the high half) out of the packed columns
template <int NA, int NB>
ICB_HD uint32_t fdct_pick2(const uint32_t (&P)[32])
{
    constexpr int ra = NA >> 3, ka = NA & 7, rb = NB >> 3, kb = NB & 7;
    constexpr uint32_t sel = (uint32_t)(2 * (ra & 1)) | ((uint32_t)(2 * (ra & 1) + 1) << 4) |
                             ((uint32_t)(4 + 2 * (rb & 1)) << 8) | ((uint32_t)(5 + 2 * (rb & 1)) << 12);
    return prmt(P[4 * ka + (ra >> 1)], P[4 * kb + (rb >> 1)], sel);
}
// 16-byte chunk KK (zigzag positions 8 KK .. 8 KK + 7) of the block's output
template <int KK>
ICB_HD void fdct_zz_chunk(const uint32_t (&P)[32], uint32_t &x, uint32_t &y, uint32_t &z, uint32_t &w)
{
    x = fdct_pick2<kZZc[8 * KK + 0], kZZc[8 * KK + 1]>(P);
    y = fdct_pick2<kZZc[8 * KK + 2], kZZc[8 * KK + 3]>(P);
    z = fdct_pick2<kZZc[8 * KK + 4], kZZc[8 * KK + 5]>(P);
    w = fdct_pick2<kZZc[8 * KK + 6], kZZc[8 * KK + 7]>(P);
}

struct Ycc { int y, cb, cr; };

// jccolor.c rgb_ycc_convert, SCALEBITS 16 (a2.2) -- scalar form, image-edge strips only
ICB_HD Ycc rgb2ycc(int r, int g, int b)
{
    Ycc o;
    o.y = (19595 * r + 38470 * g + 7471 * b + 32768) >> 16;
    o.cb = (-11059 * r - 21709 * g + 32768 * b + 8421375) >> 16;
    o.cr = (32768 * r - 27439 * g - 5329 * b + 8421375) >> 16;
    return o;
}

// One colour component of four packed 24-bit pixels (three words) by dp2a: coefficients (c0, c1, c2)
// in MEMORY byte order, K the rounding constant.  Results keep SCALEBITS = 16 fraction bits.
//   w0 = p0.0 p0.1 p0.2 p1.0   w1 = p1.1 p1.2 p2.0 p2.1   w2 = p2.2 p3.0 p3.1 p3.2
template <bool SGN, int C0, int C1, int C2, int K>
ICB_HD void comp4(uint32_t w0, uint32_t w1, uint32_t w2, int (&t)[4])
{
    t[0] = dp2a<true, SGN>(pair16(C2, 0), w0, dp2a<false, SGN>(pair16(C0, C1), w0, K));
    t[1] = dp2a<false, SGN>(pair16(C1, C2), w1, dp2a<true, SGN>(pair16(0, C0), w0, K));
    t[2] = dp2a<false, SGN>(pair16(C2, 0), w2, dp2a<true, SGN>(pair16(C0, C1), w1, K));
    t[3] = dp2a<true, SGN>(pair16(C1, C2), w2, dp2a<false, SGN>(pair16(0, C0), w2, K));
}

// Negated, offset chroma: n = 256 - c, from X = -(c's linear form):  c = (8421375 - X) >> 16
//   <=>  n = (X + 65535 - 8421375 + (256 << 16)) >> 16, 1 <= n <= 256 (so the shift is a plain one)
constexpr int ICB_NEG_CHROMA_K = 65535 - 8421375 + (256 << 16);

// Eight pixels of one row (24 bytes): luma packed as bytes into y_lo / y_hi, and the four
// horizontal pair sums of the negated chroma added into ncb[] / ncr[].
template <bool BGR>
ICB_HD void convert_row8(const unsigned char *p, uint32_t &y_lo, uint32_t &y_hi, int (&ncb)[4],
                                             int (&ncr)[4])
{
    const uint2 *q = reinterpret_cast<const uint2 *>(p);
    const uint2 a = q[0], b = q[1], c = q[2];
    const uint32_t w[6] = {a.x, a.y, b.x, b.y, c.x, c.y};
#pragma unroll
    for (int h = 0; h < 2; h++) {
        int ty[4], tb[4], tr[4];
        if (BGR) {
            comp4<false, 7471, 38470, 19595, 32768>(w[3 * h], w[3 * h + 1], w[3 * h + 2], ty);
            comp4<true, -32768, 21709, 11059, ICB_NEG_CHROMA_K>(w[3 * h], w[3 * h + 1], w[3 * h + 2], tb);
            comp4<true, 5329, 27439, -32768, ICB_NEG_CHROMA_K>(w[3 * h], w[3 * h + 1], w[3 * h + 2], tr);
        } else {
            comp4<false, 19595, 38470, 7471, 32768>(w[3 * h], w[3 * h + 1], w[3 * h + 2], ty);
            comp4<true, 11059, 21709, -32768, ICB_NEG_CHROMA_K>(w[3 * h], w[3 * h + 1], w[3 * h + 2], tb);
            comp4<true, -32768, 27439, 5329, ICB_NEG_CHROMA_K>(w[3 * h], w[3 * h + 1], w[3 * h + 2], tr);
        }
        // luma 0..255 sits in byte 2 of each t (byte 3 is zero)
        const uint32_t y4 = prmt(prmt(ty[0], ty[1], 0x0062), prmt(ty[2], ty[3], 0x0062), 0x5410);
        if (h == 0) y_lo = y4; else y_hi = y4;
        ncb[2 * h] += ((uint32_t)tb[0] >> 16) + ((uint32_t)tb[1] >> 16);
        ncb[2 * h + 1] += ((uint32_t)tb[2] >> 16) + ((uint32_t)tb[3] >> 16);
        ncr[2 * h] += ((uint32_t)tr[0] >> 16) + ((uint32_t)tr[1] >> 16);
        ncr[2 * h + 1] += ((uint32_t)tr[2] >> 16) + ((uint32_t)tr[3] >> 16);
    }
}

}  // namespace icb
