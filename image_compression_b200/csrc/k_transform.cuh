// k_transform.cuh -- K1: RGB8 -> YCbCr -> 4:2:0 -> islow FDCT, once per image.
//
// Replaces, for every candidate quality at once, the quality-independent front
// half of each javax.imageio encode the reference issues per search step
// (core/ImageCompressionJpg.java:174-175 -> :136-147): IJG jccolor.c
// rgb_ycc_convert (a2.2), jcsample.c h2v2_downsample + edge expansion (a2.3),
// jfdctint.c jpeg_fdct_islow (a2.4) and jccoefct.c's MCU order and dummy blocks
// (a2.6).  Output: unquantised coefficients, int16, MCU-interleaved block order
// (Y00 Y01 Y10 Y11 Cb Cr), 64 per block in zigzag order = 128 B per block.
//
// HBM-bound by design: 3 B/px in + 3 B/px out.  One persistent CTA streams
// 16-MCU tiles: a 2-D TMA box (768 B x 16 rows, zero-filled out of bounds) lands
// in shared memory under an mbarrier; 8x2-pixel strips are colour-converted and
// downsampled into a swizzled int32 sample buffer; one thread then runs a whole
// 8x8 block's 16 butterflies in registers and the tile leaves as coalesced
// 128-bit stores.  No tensor cores: the DCT here is 8-point integer butterflies
// with IJG's exact rounding, not a dense contraction.
#pragma once

#include <cuda.h>
#include "common.cuh"

namespace icb {

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!done);
}

__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int x, int y, uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar))
        : "memory");
}

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

// One 8-point islow butterfly.  PASS1: rows, keeps PASS1_BITS=2 extra bits and
// removes the +128 level shift (raw samples come in: only output 0 sees it,
// by exactly 8*128<<2).  PASS2: columns, descales by CONST_BITS+PASS1_BITS.
template <bool PASS1>
__device__ __forceinline__ void fdct8(int &d0, int &d1, int &d2, int &d3, int &d4, int &d5, int &d6, int &d7)
{
    constexpr int SH = PASS1 ? 11 : 15;
    constexpr int RND = 1 << (SH - 1);
    int tmp0 = d0 + d7, tmp7 = d0 - d7;
    int tmp1 = d1 + d6, tmp6 = d1 - d6;
    int tmp2 = d2 + d5, tmp5 = d2 - d5;
    int tmp3 = d3 + d4, tmp4 = d3 - d4;
    int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3;
    int tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    if (PASS1) {
        d0 = ((tmp10 + tmp11) << 2) - 4096;
        d4 = (tmp10 - tmp11) << 2;
    } else {
        d0 = (tmp10 + tmp11 + 2) >> 2;
        d4 = (tmp10 - tmp11 + 2) >> 2;
    }
    int z1 = (tmp12 + tmp13) * ICB_FIX_0_541196100 + RND;
    d2 = (z1 + tmp13 * ICB_FIX_0_765366865) >> SH;
    d6 = (z1 - tmp12 * ICB_FIX_1_847759065) >> SH;
    z1 = tmp4 + tmp7;
    int z2 = tmp5 + tmp6;
    int z3 = tmp4 + tmp6;
    int z4 = tmp5 + tmp7;
    int z5 = (z3 + z4) * ICB_FIX_1_175875602 + RND;
    z3 = z5 - z3 * ICB_FIX_1_961570560;
    z4 = z5 - z4 * ICB_FIX_0_390180644;
    z1 = z1 * (-ICB_FIX_0_899976223);
    z2 = z2 * (-ICB_FIX_2_562915447);
    d7 = (tmp4 * ICB_FIX_0_298631336 + z1 + z3) >> SH;
    d5 = (tmp5 * ICB_FIX_2_053119869 + z2 + z4) >> SH;
    d3 = (tmp6 * ICB_FIX_3_072711026 + z2 + z3) >> SH;
    d1 = (tmp7 * ICB_FIX_1_501321110 + z1 + z4) >> SH;
}

// zigzag index -> natural index, compile-time copy for register indexing
__device__ constexpr unsigned char kZZ[64] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5,
    12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
    58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct Ycc { int y, cb, cr; };

// jccolor.c rgb_ycc_convert, SCALEBITS 16 (a2.2)
__device__ __forceinline__ Ycc rgb2ycc(int r, int g, int b)
{
    Ycc o;
    o.y = (19595 * r + 38470 * g + 7471 * b + 32768) >> 16;
    o.cb = (-11059 * r - 21709 * g + 32768 * b + 8421375) >> 16;
    o.cr = (32768 * r - 27439 * g - 5329 * b + 8421375) >> 16;
    return o;
}

template <int K>
__device__ __forceinline__ int byte_of(const uint32_t (&w)[6])
{
    return (int)((w[K >> 2] >> (8 * (K & 3))) & 0xffu);
}

// swizzled 16-byte chunk address inside the sample / staging buffers
__device__ __forceinline__ int samp_off(int blk, int chunk)      // in int32 units, 64 per block
{
    return blk * 64 + ((chunk ^ (blk & 7)) << 2);
}

template <bool BGR>
__device__ __forceinline__ void convert_row8(const unsigned char *p, int (&yv)[8], int (&cb)[8], int (&cr)[8])
{
    uint32_t w[6];
    const uint2 *q = reinterpret_cast<const uint2 *>(p);
    uint2 a = q[0], b = q[1], c = q[2];
    w[0] = a.x; w[1] = a.y; w[2] = b.x; w[3] = b.y; w[4] = c.x; w[5] = c.y;
#define ICB_PX(i)                                                                          \
    {                                                                                      \
        int c0 = byte_of<3 * i>(w), c1 = byte_of<3 * i + 1>(w), c2 = byte_of<3 * i + 2>(w); \
        Ycc o = BGR ? rgb2ycc(c2, c1, c0) : rgb2ycc(c0, c1, c2);                           \
        yv[i] = o.y; cb[i] = o.cb; cr[i] = o.cr;                                           \
    }
    ICB_PX(0) ICB_PX(1) ICB_PX(2) ICB_PX(3) ICB_PX(4) ICB_PX(5) ICB_PX(6) ICB_PX(7)
#undef ICB_PX
}

template <bool BGR>
__device__ __forceinline__ Ycc load_px(const unsigned char *tile, int ly, int lx)
{
    const unsigned char *p = tile + ly * TILE_ROW_BYTES + lx * 3;
    int c0 = p[0], c1 = p[1], c2 = p[2];
    return BGR ? rgb2ycc(c2, c1, c0) : rgb2ycc(c0, c1, c2);
}

template <bool BGR>
__global__ void __launch_bounds__(XF_THREADS, 4)
k_transform(const __grid_constant__ CUtensorMap tmap, const Geom g, int16_t *__restrict__ coef,
            int tiles_per_row, int ntiles)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char *in_buf = smem_raw;                                             // XF_STAGES x 12 KB
    int *samp = reinterpret_cast<int *>(smem_raw + XF_STAGES * TILE_IN_BYTES);    // 96 x 64 int32
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + XF_STAGES * TILE_IN_BYTES + TILE_BLOCKS * 256);

    const int tid = threadIdx.x;
    pdl_launch_dependents();        // k_entropy_prepare may set up while this grid drains
    if (tid == 0) {
        for (int s = 0; s < XF_STAGES; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // prologue: fill the pipeline
    if (tid == 0) {
        for (int s = 0; s < XF_STAGES; s++) {
            int t = blockIdx.x + s * gridDim.x;
            if (t < ntiles) {
                int my = t / tiles_per_row, tx = t - my * tiles_per_row;
                mbar_expect_tx(&full[s], TILE_IN_BYTES);
                tma_load_2d(in_buf + s * TILE_IN_BYTES, &tmap, tx * (TILE_ROW_BYTES / 4), my * 16, &full[s]);
            }
        }
    }

    int k = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, k++) {
        const int s = k % XF_STAGES;
        const uint32_t parity = (k / XF_STAGES) & 1;
        const int my = t / tiles_per_row, tx = t - my * tiles_per_row;
        const int mcu0 = tx * TILE_MCUS;
        unsigned char *tile = in_buf + s * TILE_IN_BYTES;
        mbar_wait(&full[s], parity);

        // ---- stage A: colour conversion + 4:2:0 downsample, 8x2-pixel strips ----
#pragma unroll
        for (int round = 0; round < 2; round++) {
            const int sid = tid + round * XF_THREADS;
            const int sc = sid & 31, sr = sid >> 5;
            const int m = sc >> 1, xi = sc & 1;
            if (mcu0 + m >= g.mcu_cols) continue;
            const int x0 = mcu0 * 16 + sc * 8, y0 = my * 16 + 2 * sr;
            const int yblk = m * 6 + (sr >> 2) * 2 + xi;
            const int yrow = (2 * sr) & 7;
            int4 ya0, ya1, yb0, yb1, cbv, crv;
            if (x0 + 7 < g.w && y0 + 1 < g.h) {
                int y_a[8], cb_a[8], cr_a[8], y_b[8], cb_b[8], cr_b[8];
                convert_row8<BGR>(tile + (2 * sr) * TILE_ROW_BYTES + sc * 24, y_a, cb_a, cr_a);
                convert_row8<BGR>(tile + (2 * sr + 1) * TILE_ROW_BYTES + sc * 24, y_b, cb_b, cr_b);
                ya0 = make_int4(y_a[0], y_a[1], y_a[2], y_a[3]);
                ya1 = make_int4(y_a[4], y_a[5], y_a[6], y_a[7]);
                yb0 = make_int4(y_b[0], y_b[1], y_b[2], y_b[3]);
                yb1 = make_int4(y_b[4], y_b[5], y_b[6], y_b[7]);
                // jcsample.c h2v2_downsample: bias 1,2,1,2 by output column (a2.3)
                cbv = make_int4((cb_a[0] + cb_a[1] + cb_b[0] + cb_b[1] + 1) >> 2,
                                (cb_a[2] + cb_a[3] + cb_b[2] + cb_b[3] + 2) >> 2,
                                (cb_a[4] + cb_a[5] + cb_b[4] + cb_b[5] + 1) >> 2,
                                (cb_a[6] + cb_a[7] + cb_b[6] + cb_b[7] + 2) >> 2);
                crv = make_int4((cr_a[0] + cr_a[1] + cr_b[0] + cr_b[1] + 1) >> 2,
                                (cr_a[2] + cr_a[3] + cr_b[2] + cr_b[3] + 2) >> 2,
                                (cr_a[4] + cr_a[5] + cr_b[4] + cr_b[5] + 1) >> 2,
                                (cr_a[6] + cr_a[7] + cr_b[6] + cr_b[7] + 2) >> 2);
            } else {
                // image edge: expand_right_edge / expand_bottom_edge by clamping; rows past
                // the last downsampled chroma row replicate THAT row (jcprepct.c)
                const int lx0 = sc * 8;
                int yy[2][8];
#pragma unroll
                for (int r = 0; r < 2; r++) {
                    int ly = min(y0 + r, g.h - 1) - my * 16;
#pragma unroll
                    for (int i = 0; i < 8; i++) {
                        int lx = min(x0 + i, g.w - 1) - mcu0 * 16;
                        yy[r][i] = load_px<BGR>(tile, ly, lx).y;
                    }
                }
                (void)lx0;
                const int cy = min(my * 8 + sr, g.crows_real - 1);
                const int la = 2 * cy - my * 16, lb = min(2 * cy + 1, g.h - 1) - my * 16;
                int cbo[4], cro[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    int lxa = min(x0 + 2 * j, g.w - 1) - mcu0 * 16;
                    int lxb = min(x0 + 2 * j + 1, g.w - 1) - mcu0 * 16;
                    Ycc p00 = load_px<BGR>(tile, la, lxa), p01 = load_px<BGR>(tile, la, lxb);
                    Ycc p10 = load_px<BGR>(tile, lb, lxa), p11 = load_px<BGR>(tile, lb, lxb);
                    int bias = 1 + (j & 1);
                    cbo[j] = (p00.cb + p01.cb + p10.cb + p11.cb + bias) >> 2;
                    cro[j] = (p00.cr + p01.cr + p10.cr + p11.cr + bias) >> 2;
                }
                ya0 = make_int4(yy[0][0], yy[0][1], yy[0][2], yy[0][3]);
                ya1 = make_int4(yy[0][4], yy[0][5], yy[0][6], yy[0][7]);
                yb0 = make_int4(yy[1][0], yy[1][1], yy[1][2], yy[1][3]);
                yb1 = make_int4(yy[1][4], yy[1][5], yy[1][6], yy[1][7]);
                cbv = make_int4(cbo[0], cbo[1], cbo[2], cbo[3]);
                crv = make_int4(cro[0], cro[1], cro[2], cro[3]);
            }
            *reinterpret_cast<int4 *>(samp + samp_off(yblk, 2 * yrow)) = ya0;
            *reinterpret_cast<int4 *>(samp + samp_off(yblk, 2 * yrow + 1)) = ya1;
            *reinterpret_cast<int4 *>(samp + samp_off(yblk, 2 * yrow + 2)) = yb0;
            *reinterpret_cast<int4 *>(samp + samp_off(yblk, 2 * yrow + 3)) = yb1;
            *reinterpret_cast<int4 *>(samp + samp_off(m * 6 + 4, 2 * sr + xi)) = cbv;
            *reinterpret_cast<int4 *>(samp + samp_off(m * 6 + 5, 2 * sr + xi)) = crv;
        }
        __syncthreads();            // samples complete; the RGB tile is dead

        // ---- stage B: one thread = one 8x8 block, 16 butterflies in registers ----
        if (tid < TILE_BLOCKS) {
            const int blk = tid;
            const int m = blk / 6, c = blk - m * 6;
            const int gx = mcu0 + m;
            if (gx < g.mcu_cols) {
                int src = blk;
                bool dummy = false;
                if (c < 4) {        // jccoefct.c dummy blocks (a2.6): luma only
                    const int yi = c >> 1, xi = c & 1;
                    const bool row_real = (2 * my + yi) < g.yblk_h;
                    const bool col_real = (2 * gx + xi) < g.yblk_w;
                    if (!(row_real && col_real)) {
                        dummy = true;
                        if (row_real) src = m * 6 + yi * 2;                       // block to its left
                        else src = m * 6 + ((2 * gx + 1) < g.yblk_w ? 1 : 0);   // MCU_buffer[blkn-1]
                    }
                }
                int d[64];
#pragma unroll
                for (int ch = 0; ch < 16; ch++) {
                    int4 v = *reinterpret_cast<const int4 *>(samp + samp_off(src, ch));
                    d[4 * ch] = v.x; d[4 * ch + 1] = v.y; d[4 * ch + 2] = v.z; d[4 * ch + 3] = v.w;
                }
#pragma unroll
                for (int r = 0; r < 8; r++)
                    fdct8<true>(d[8 * r], d[8 * r + 1], d[8 * r + 2], d[8 * r + 3], d[8 * r + 4], d[8 * r + 5],
                                d[8 * r + 6], d[8 * r + 7]);
#pragma unroll
                for (int cc = 0; cc < 8; cc++)
                    fdct8<false>(d[cc], d[8 + cc], d[16 + cc], d[24 + cc], d[32 + cc], d[40 + cc], d[48 + cc],
                                 d[56 + cc]);
                if (dummy) {
#pragma unroll
                    for (int i = 1; i < 64; i++) d[i] = 0;
                }
                // zigzag + int16 pack into the (now free) RGB stage buffer, swizzled chunks
                uint4 *outp = reinterpret_cast<uint4 *>(tile + blk * 128);
#pragma unroll
                for (int kk = 0; kk < 8; kk++) {
                    uint4 o;
                    o.x = __byte_perm(d[kZZ[8 * kk + 0]], d[kZZ[8 * kk + 1]], 0x5410);
                    o.y = __byte_perm(d[kZZ[8 * kk + 2]], d[kZZ[8 * kk + 3]], 0x5410);
                    o.z = __byte_perm(d[kZZ[8 * kk + 4]], d[kZZ[8 * kk + 5]], 0x5410);
                    o.w = __byte_perm(d[kZZ[8 * kk + 6]], d[kZZ[8 * kk + 7]], 0x5410);
                    outp[kk ^ (blk & 7)] = o;
                }
            }
        }
        __syncthreads();            // coefficients staged

        // ---- copy-out: 96 blocks x 128 B, consecutive lanes -> consecutive 16 B ----
        {
            const size_t base = ((size_t)my * g.mcu_cols + mcu0) * 6;
            uint4 *gout = reinterpret_cast<uint4 *>(coef) + base * 8;
#pragma unroll
            for (int i = 0; i < (TILE_BLOCKS * 8) / XF_THREADS; i++) {
                const int idx = tid + i * XF_THREADS;
                const int blk = idx >> 3, kk = idx & 7;
                if (mcu0 + blk / 6 < g.mcu_cols) {
                    uint4 v = reinterpret_cast<const uint4 *>(tile + blk * 128)[kk ^ (blk & 7)];
                    gout[idx] = v;
                }
            }
        }
        __syncthreads();            // stage buffer free again

        if (tid == 0) {
            const int tn = t + XF_STAGES * gridDim.x;
            if (tn < ntiles) {
                const int my2 = tn / tiles_per_row, tx2 = tn - my2 * tiles_per_row;
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(&full[s], TILE_IN_BYTES);
                tma_load_2d(tile, &tmap, tx2 * (TILE_ROW_BYTES / 4), my2 * 16, &full[s]);
            }
        }
    }
}

constexpr int XF_SMEM_BYTES = XF_STAGES * TILE_IN_BYTES + TILE_BLOCKS * 256 + 64;

}  // namespace icb
