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
// downsampled into a swizzled BYTE sample buffer (64 B per block); one thread then
// runs a whole 8x8 block in registers and the tile leaves as coalesced 128-bit
// stores.  No tensor cores: exact IJG rounding sits between the two DCT passes.
//
// The kernel is bound by integer issue (alu and fma pipes both take two cycles per
// warp instruction), so the arithmetic is arranged for the fewest instructions,
// every form algebraically identical to IJG's in exact integer arithmetic:
//   * colour conversion by dp2a (16-bit coefficient pair x two pixel bytes) straight
//     on the packed 24-bit pixels -- the four pixels of three words each take two
//     dp2a per component with their own lo/hi coefficient pairing, no byte extraction;
//     Cb/Cr are computed negated and offset (256 - c) so that the +32768 coefficient
//     becomes -32768 and fits the signed 16-bit operand;
//   * the row pass of the DCT runs on the sample BYTES: before its descale every
//     output is a fixed linear form of the eight samples (coefficients <= 11363), i.e.
//     four dp2a per output, two dp4a for outputs 0 and 4;
//   * the column pass keeps the butterflies for the even half and uses the expanded
//     three-multiply form for each odd output.
#pragma once

#include <cuda.h>
#include "common.cuh"
#include "xf_arith.cuh"

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

// zigzag index -> natural index, compile-time copy for register indexing
__device__ constexpr unsigned char kZZ[64] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5,
    12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
    58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

template <bool BGR>
__device__ __forceinline__ Ycc load_px(const unsigned char *tile, int ly, int lx)
{
    const unsigned char *p = tile + ly * TILE_ROW_BYTES + lx * 3;
    int c0 = p[0], c1 = p[1], c2 = p[2];
    return BGR ? rgb2ycc(c2, c1, c0) : rgb2ycc(c0, c1, c2);
}

// sample buffer: 64 bytes per block (row r at byte 8 r), 16-byte chunks swizzled so that both the
// strip writers and the one-thread-per-block readers are conflict free
__device__ __forceinline__ int samp_chunk(int blk, int chunk)     // in 16-byte units
{
    return blk * 4 + (chunk ^ ((blk >> 1) & 3));
}

template <bool BGR>
__global__ void __launch_bounds__(XF_THREADS, XF_CTAS)
k_transform(const __grid_constant__ CUtensorMap tmap, const Geom g, int16_t *__restrict__ coef,
            int tiles_per_row, int ntiles)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char *in_buf = smem_raw;                                             // XF_STAGES x 12 KB
    uint4 *samp = reinterpret_cast<uint4 *>(smem_raw + XF_STAGES * TILE_IN_BYTES);    // 96 blocks x 64 sample bytes
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + XF_STAGES * TILE_IN_BYTES + TILE_BLOCKS * 64);

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

    const int cta_slot = blockIdx.x / max(1, (int)(gridDim.x + XF_CTAS - 1) / XF_CTAS);    // CTAs of one SM are gridDim / XF_CTAS apart
    int k = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, k++) {
        const int s = k % XF_STAGES;
        const uint32_t parity = (k / XF_STAGES) & 1;
        const int my = t / tiles_per_row, tx = t - my * tiles_per_row;
        const int mcu0 = tx * TILE_MCUS;
        unsigned char *tile = in_buf + s * TILE_IN_BYTES;
        mbar_wait(&full[s], parity);

#ifndef ICB_XF_DBG
#define ICB_XF_DBG 0        // experiments only: 1 = no arithmetic (TMA in, barriers, copy-out), 2 = no global stores
#endif
        // ---- stage A: colour conversion + 4:2:0 downsample, 8x2-pixel strips ----
#pragma unroll
        for (int round = 0; round < ((ICB_XF_DBG & 1) ? 0 : 2); round++) {
            const int sid = tid + round * XF_THREADS;
            const int sc = sid & 31, sr = sid >> 5;
            const int m = sc >> 1, xi = sc & 1;
            if (mcu0 + m >= g.mcu_cols) continue;
            const int x0 = mcu0 * 16 + sc * 8, y0 = my * 16 + 2 * sr;
            const int yblk = m * 6 + (sr >> 2) * 2 + xi;
            const int yrow = (2 * sr) & 7;
            uint4 yv;               // rows yrow, yrow + 1 of the luma block = one 16-byte chunk
            uint32_t cbv, crv;      // four chroma samples of row sr, columns 4 xi .. 4 xi + 3
            if (x0 + 7 < g.w && y0 + 1 < g.h) {
                int ncb[4] = {0, 0, 0, 0}, ncr[4] = {0, 0, 0, 0};
                convert_row8<BGR>(tile + (2 * sr) * TILE_ROW_BYTES + sc * 24, yv.x, yv.y, ncb, ncr);
                convert_row8<BGR>(tile + (2 * sr + 1) * TILE_ROW_BYTES + sc * 24, yv.z, yv.w, ncb, ncr);
                // jcsample.c h2v2_downsample: bias 1,2,1,2 by output column (a2.3); the sums are of
                // n = 256 - c, so (c00 + c01 + c10 + c11 + bias) >> 2 = (1024 + bias - sum n) >> 2
                cbv = (uint32_t)((1025 - ncb[0]) >> 2) | ((uint32_t)((1026 - ncb[1]) >> 2) << 8) |
                      ((uint32_t)((1025 - ncb[2]) >> 2) << 16) | ((uint32_t)((1026 - ncb[3]) >> 2) << 24);
                crv = (uint32_t)((1025 - ncr[0]) >> 2) | ((uint32_t)((1026 - ncr[1]) >> 2) << 8) |
                      ((uint32_t)((1025 - ncr[2]) >> 2) << 16) | ((uint32_t)((1026 - ncr[3]) >> 2) << 24);
            } else {
                // image edge: expand_right_edge / expand_bottom_edge by clamping; rows past
                // the last downsampled chroma row replicate THAT row (jcprepct.c)
                uint32_t yy[4] = {0, 0, 0, 0};
#pragma unroll
                for (int r = 0; r < 2; r++) {
                    int ly = min(y0 + r, g.h - 1) - my * 16;
#pragma unroll
                    for (int i = 0; i < 8; i++) {
                        int lx = min(x0 + i, g.w - 1) - mcu0 * 16;
                        yy[2 * r + (i >> 2)] |= (uint32_t)load_px<BGR>(tile, ly, lx).y << (8 * (i & 3));
                    }
                }
                yv = make_uint4(yy[0], yy[1], yy[2], yy[3]);
                const int cy = min(my * 8 + sr, g.crows_real - 1);
                const int la = 2 * cy - my * 16, lb = min(2 * cy + 1, g.h - 1) - my * 16;
                cbv = 0; crv = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    int lxa = min(x0 + 2 * j, g.w - 1) - mcu0 * 16;
                    int lxb = min(x0 + 2 * j + 1, g.w - 1) - mcu0 * 16;
                    Ycc p00 = load_px<BGR>(tile, la, lxa), p01 = load_px<BGR>(tile, la, lxb);
                    Ycc p10 = load_px<BGR>(tile, lb, lxa), p11 = load_px<BGR>(tile, lb, lxb);
                    int bias = 1 + (j & 1);
                    cbv |= (uint32_t)((p00.cb + p01.cb + p10.cb + p11.cb + bias) >> 2) << (8 * j);
                    crv |= (uint32_t)((p00.cr + p01.cr + p10.cr + p11.cr + bias) >> 2) << (8 * j);
                }
            }
            samp[samp_chunk(yblk, yrow >> 1)] = yv;
            {
                uint32_t *s32 = reinterpret_cast<uint32_t *>(samp);
                const int w = ((sr & 1) << 1) + xi;                    // word inside the chunk
                s32[samp_chunk(m * 6 + 4, sr >> 1) * 4 + w] = cbv;
                s32[samp_chunk(m * 6 + 5, sr >> 1) * 4 + w] = crv;
            }
        }
        __syncthreads();            // samples complete; the RGB tile is dead

        // ---- stage B: one thread = one 8x8 block, both DCT passes in registers ----
        // 96 blocks on 4 warps: one warp sits a tile out.  Which one rotates with the tile count
        // and with the CTA's position among the CTAs sharing the SM, so that every SM
        // sub-partition (warp w of each CTA lands on sub-partition w) carries 3/4 of the DCT work
        // instead of three of them carrying all of it.
        const int vt = (tid + 32 * ((k + cta_slot) & 3)) & (XF_THREADS - 1);
        if (!(ICB_XF_DBG & 1) && vt < TILE_BLOCKS) {
            const int blk = vt;
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
                uint32_t sw[16];
#pragma unroll
                for (int ch = 0; ch < 4; ch++) {
                    const uint4 v = samp[samp_chunk(src, ch)];
                    sw[4 * ch] = v.x; sw[4 * ch + 1] = v.y; sw[4 * ch + 2] = v.z; sw[4 * ch + 3] = v.w;
                }
#ifdef ICB_XF_ROWMAJOR      // kept for A/B: all 64 intermediates alive between the passes (122 registers)
                int d[64];
#pragma unroll
                for (int r = 0; r < 8; r++) fdct_row_bytes(sw[2 * r], sw[2 * r + 1], &d[8 * r]);
#pragma unroll
                for (int cc = 0; cc < 8; cc++)
                    fdct_col_hi16(d[cc], d[8 + cc], d[16 + cc], d[24 + cc], d[32 + cc], d[40 + cc], d[48 + cc],
                                  d[56 + cc]);
                if (dummy) {
#pragma unroll
                    for (int i = 1; i < 64; i++) d[i] = 0;
                }
                // zigzag + int16 pack (upper half-words, see fdct_col_hi16) into the (now free) RGB stage buffer, swizzled chunks
                uint4 *outp = reinterpret_cast<uint4 *>(tile + blk * 128);
#pragma unroll
                for (int kk = 0; kk < 8; kk++) {
                    uint4 o;
                    o.x = __byte_perm(d[kZZ[8 * kk + 0]], d[kZZ[8 * kk + 1]], 0x7632);
                    o.y = __byte_perm(d[kZZ[8 * kk + 2]], d[kZZ[8 * kk + 3]], 0x7632);
                    o.z = __byte_perm(d[kZZ[8 * kk + 4]], d[kZZ[8 * kk + 5]], 0x7632);
                    o.w = __byte_perm(d[kZZ[8 * kk + 6]], d[kZZ[8 * kk + 7]], 0x7632);
                    outp[kk ^ (blk & 7)] = o;
                }
#else
                uint32_t P[32];
                fdct_block_bytes(sw, P);
                if (dummy) {
                    P[0] &= 0xffffu;
#pragma unroll
                    for (int i = 1; i < 32; i++) P[i] = 0;
                }
                // zigzag + int16 pack into the (now free) RGB stage buffer, swizzled chunks
                uint4 *outp = reinterpret_cast<uint4 *>(tile + blk * 128);
                uint4 o;
#define ICB_CHUNK(KK) fdct_zz_chunk<KK>(P, o.x, o.y, o.z, o.w); outp[KK ^ (blk & 7)] = o;
                ICB_CHUNK(0) ICB_CHUNK(1) ICB_CHUNK(2) ICB_CHUNK(3) ICB_CHUNK(4) ICB_CHUNK(5) ICB_CHUNK(6) ICB_CHUNK(7)
#undef ICB_CHUNK
#endif
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
                if (!(ICB_XF_DBG & 2) && mcu0 + blk / 6 < g.mcu_cols) {
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

constexpr int XF_SMEM_BYTES = XF_STAGES * TILE_IN_BYTES + TILE_BLOCKS * 64 + 64;

}  // namespace icb
