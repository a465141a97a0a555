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
#include "k_control.cuh"
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

// natural index -> zigzag position
__device__ constexpr unsigned char kZZinv[64] = {
    0, 1, 5, 6, 14, 15, 27, 28, 2, 4, 7, 13, 16, 26, 29, 42,
    3, 8, 12, 17, 25, 30, 41, 43, 9, 11, 18, 24, 31, 40, 44, 53,
    10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38, 46, 51, 55, 60,
    21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48, 49, 57, 58, 62, 63};

// Survivor lists written by the transform kernel (k_entropy.cuh has the consumer side): the unit of
// tile t gets its entries while the tile's coefficients are still in shared memory.  ent == nullptr
// switches the lists off (coefficient-only callers).
struct XfLists {
    uint32_t *ent;                  // [nunits][UNIT_CAP]
    unsigned int *ent_cnt;          // [nunits]
    unsigned short *unit_info;      // [nunits][16] what k_job_begin links the units with: [0], [1] entry index of the first
                                    //   MCU's Cb / Cr DC entry; [2..4] raw DC of the last MCU's Y11 / Cb / Cr; [6..11] the
                                    //   first MCU's Y00 / Cb / Cr DC entries themselves (32 bits each, predictor field 0)
    float list_q;                   // quality whose tables bound the lists (k_entropy_prepare has the rule)
    int table_mode;
    FitState *st;                   // kernel clock of the job: block 0 stamps FitState.kt[0][0]; the END of this kernel is
                                    // stamped by the job-start kernel when its griddepcontrol.wait returns.  nullptr: none
};

// survivor test on a column-pass result d = x * 65536 + junk (0 <= junk < 65536): |x| >= t  <=>  d lies
// outside [(1 - t) << 16, t << 16)  <=>  (unsigned)(d - A) >= B with A = (1 - t) << 16, B = (2 t - 1) << 16
template <int NAT>
__device__ __forceinline__ void xf_test(int d, uint32_t A, uint32_t B, uint32_t &mlo, uint32_t &mhi)
{
    constexpr int z = kZZinv[NAT];
    constexpr uint32_t blo = z < 32 ? 1u << (z & 31) : 0u, bhi = z < 32 ? 0u : 1u << (z & 31);
    if ((uint32_t)d - A >= B) {
        if (z < 32) mlo |= blo; else mhi |= bhi;
    }
}

// fdct_block_col (xf_arith.cuh) plus the survivor mask of column K: thr = this block's table,
// [4 K + j] = (A, B) of rows 2 j and 2 j + 1
template <int K>
__device__ __forceinline__ void fdct_block_col_m(const uint32_t (&sw)[16], uint32_t (&P)[32], const uint4 *thr,
                                                 uint32_t &mlo, uint32_t &mhi)
{
    int v0 = fdct_row_k<K>(sw[0], sw[1]), v1 = fdct_row_k<K>(sw[2], sw[3]);
    int v2 = fdct_row_k<K>(sw[4], sw[5]), v3 = fdct_row_k<K>(sw[6], sw[7]);
    int v4 = fdct_row_k<K>(sw[8], sw[9]), v5 = fdct_row_k<K>(sw[10], sw[11]);
    int v6 = fdct_row_k<K>(sw[12], sw[13]), v7 = fdct_row_k<K>(sw[14], sw[15]);
    fdct_col_hi16(v0, v1, v2, v3, v4, v5, v6, v7);
    uint4 t = thr[4 * K];
    xf_test<K>(v0, t.x, t.y, mlo, mhi); xf_test<8 + K>(v1, t.z, t.w, mlo, mhi);
    t = thr[4 * K + 1];
    xf_test<16 + K>(v2, t.x, t.y, mlo, mhi); xf_test<24 + K>(v3, t.z, t.w, mlo, mhi);
    t = thr[4 * K + 2];
    xf_test<32 + K>(v4, t.x, t.y, mlo, mhi); xf_test<40 + K>(v5, t.z, t.w, mlo, mhi);
    t = thr[4 * K + 3];
    xf_test<48 + K>(v6, t.x, t.y, mlo, mhi); xf_test<56 + K>(v7, t.z, t.w, mlo, mhi);
    P[4 * K + 0] = prmt((uint32_t)v0, (uint32_t)v1, 0x7632);
    P[4 * K + 1] = prmt((uint32_t)v2, (uint32_t)v3, 0x7632);
    P[4 * K + 2] = prmt((uint32_t)v4, (uint32_t)v5, 0x7632);
    P[4 * K + 3] = prmt((uint32_t)v6, (uint32_t)v7, 0x7632);
}

constexpr int XF_OFF_SAMP = XF_STAGES * TILE_IN_BYTES;
constexpr int XF_OFF_BAR = XF_OFF_SAMP + TILE_BLOCKS * 64;
constexpr int XF_OFF_THR = XF_OFF_BAR + 64;             // 2 tables x 64 x (A, B)
constexpr int XF_OFF_WTOT = XF_OFF_THR + 1024;
constexpr int XF_SMEM_BYTES = XF_OFF_WTOT + 64;
static_assert(XF_OFF_BAR % 16 == 0, "shared-memory carve-up");

template <bool BGR, bool LISTS>
__global__ void __launch_bounds__(XF_THREADS, XF_CTAS)
k_transform(const __grid_constant__ CUtensorMap tmap, const Geom g, int16_t *__restrict__ coef,
            int tiles_per_row, int ntiles, const XfLists xl)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char *in_buf = smem_raw;                                             // XF_STAGES TMA boxes
    uint4 *samp = reinterpret_cast<uint4 *>(smem_raw + XF_OFF_SAMP);              // TILE_BLOCKS x 64 sample bytes
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + XF_OFF_BAR);
    uint4 *s_thr = reinterpret_cast<uint4 *>(smem_raw + XF_OFF_THR);              // [table][4 K + j]
    uint32_t *s_wtot = reinterpret_cast<uint32_t *>(smem_raw + XF_OFF_WTOT);      // survivor entries per warp of blocks

    const int tid = threadIdx.x, lane = tid & 31;
    constexpr bool lists = LISTS;           // (the coefficient-only instance carries no survivor test at all)
    pdl_launch_dependents();        // the job-start kernel may set up while this grid drains
    if (tid == 0) {
        if (blockIdx.x == 0 && xl.st) xl.st->kt[0][0] = gtime();
        for (int s = 0; s < XF_STAGES; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // survivor thresholds of the quality the lists are built for (a coefficient is non-zero at quality q
    // iff |x| >= 4 * table(q), and table(q) never grows with q); without lists nothing survives the test
    {
        const int tb = tid >> 6, nat = tid & 63, r = nat >> 3, K = nat & 7;
        uint32_t A = 0x80000000u, B = 0xffffffffu;
        if (lists) {
            const int t = scaled_qval(tb ? c_base_chr[nat] : c_base_lum[nat], xl.list_q, xl.table_mode) << 2;
            A = (uint32_t)((1 - t) * 65536);
            B = (uint32_t)((2 * t - 1) * 65536);
        }
        reinterpret_cast<uint2 *>(s_thr)[tb * 64 + K * 8 + r] = make_uint2(A, B);
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

        // ---- stage A: colour conversion + 4:2:0 downsample, 8x2-pixel strips ----
#pragma unroll
        for (int round = 0; round < (TILE_MCUS * 16 + XF_THREADS - 1) / XF_THREADS; round++) {
            const int sid = tid + round * XF_THREADS;
            const int sr = sid / (2 * TILE_MCUS), sc = sid - sr * (2 * TILE_MCUS);
            const int m = sc >> 1, xi = sc & 1;
            if ((TILE_MCUS * 16) % XF_THREADS != 0 && sid >= TILE_MCUS * 16) continue;
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
        // With fewer blocks than threads one warp sits a tile out.  Which one rotates with the tile
        // count and with the CTA's position among the CTAs sharing the SM, so that every SM
        // sub-partition (warp w of each CTA lands on sub-partition w) carries its share of the DCT work.
        const int vt = TILE_BLOCKS <= XF_THREADS - 32 ? (tid + 32 * ((k + cta_slot) & 3)) & (XF_THREADS - 1) : tid;
        const int nm = min(TILE_MCUS, g.mcu_cols - mcu0);       // MCUs of this tile
        uint32_t mlo = 0, mhi = 0;                              // survivor mask of this thread's block, zigzag positions
        const int blk = vt;
        const int bm = blk / 6, bc = blk - bm * 6;
        const bool have = blk < TILE_BLOCKS && bm < nm;
        if (have) {
            const int gx = mcu0 + bm;
            int src = blk;
            bool dummy = false;
            if (bc < 4) {        // jccoefct.c dummy blocks (a2.6): luma only
                const int yi = bc >> 1, xi = bc & 1;
                const bool row_real = (2 * my + yi) < g.yblk_h;
                const bool col_real = (2 * gx + xi) < g.yblk_w;
                if (!(row_real && col_real)) {
                    dummy = true;
                    if (row_real) src = bm * 6 + yi * 2;                       // block to its left
                    else src = bm * 6 + ((2 * gx + 1) < g.yblk_w ? 1 : 0);    // MCU_buffer[blkn-1]
                }
            }
            uint32_t sw[16];
#pragma unroll
            for (int ch = 0; ch < 4; ch++) {
                const uint4 v = samp[samp_chunk(src, ch)];
                sw[4 * ch] = v.x; sw[4 * ch + 1] = v.y; sw[4 * ch + 2] = v.z; sw[4 * ch + 3] = v.w;
            }
            uint32_t P[32];
            const uint4 *thr = s_thr + (bc >= 4 ? 32 : 0);
            if (lists) {
                fdct_block_col_m<0>(sw, P, thr, mlo, mhi); fdct_block_col_m<1>(sw, P, thr, mlo, mhi);
                fdct_block_col_m<2>(sw, P, thr, mlo, mhi); fdct_block_col_m<3>(sw, P, thr, mlo, mhi);
                fdct_block_col_m<4>(sw, P, thr, mlo, mhi); fdct_block_col_m<5>(sw, P, thr, mlo, mhi);
                fdct_block_col_m<6>(sw, P, thr, mlo, mhi); fdct_block_col_m<7>(sw, P, thr, mlo, mhi);
            } else {
                fdct_block_bytes(sw, P);
            }
            if (dummy) {
                P[0] &= 0xffffu;
#pragma unroll
                for (int i = 1; i < 32; i++) P[i] = 0;
                mlo = 0; mhi = 0;
            }
            mlo |= 1u;                                          // DC always codes a symbol
            // zigzag + int16 pack into the (now free) RGB stage buffer, swizzled chunks
            uint4 *outp = reinterpret_cast<uint4 *>(tile + blk * 128);
            uint4 o;
#define ICB_CHUNK(KK) fdct_zz_chunk<KK>(P, o.x, o.y, o.z, o.w); outp[KK ^ (blk & 7)] = o;
            ICB_CHUNK(0) ICB_CHUNK(1) ICB_CHUNK(2) ICB_CHUNK(3) ICB_CHUNK(4) ICB_CHUNK(5) ICB_CHUNK(6) ICB_CHUNK(7)
#undef ICB_CHUNK
        }
        // entries in front of this block inside the unit: warp scan now, the warps' totals across the barrier
        uint32_t cnt = 0, incl = 0;
        if (lists) {
            cnt = __popc(mlo) + __popc(mhi);
            incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += u;
            }
            if (lane == 31) s_wtot[vt >> 5] = incl;
        }
        __syncthreads();            // coefficients staged

        // ---- survivor list of the unit (k_entropy.cuh entry format), from the staged coefficients ----
        if (lists) {
            uint32_t *out = xl.ent + (size_t)t * UNIT_CAP;
            const uint32_t w0 = s_wtot[0], w1 = s_wtot[1], w2 = s_wtot[2], w3 = s_wtot[3];
            const int vw = vt >> 5;
            uint32_t rank = incl - cnt + (vw > 0 ? w0 : 0u) + (vw > 1 ? w1 : 0u) + (vw > 2 ? w2 : 0u);
            const uint32_t total = w0 + w1 + w2 + w3;
            if (have) {
                const unsigned char *mine = tile + blk * 128;
                const int sw7 = blk & 7;
                const uint32_t tag = bc >= 4 ? 1u << 17 : 0u;
                // DC predictor = previous block of the same component in scan order; the first MCU's Y00, Cb
                // and Cr look into the previous unit: k_job_begin fills those three in
                int pb = blk - 1;
                if (bc == 0) pb = blk - 3; else if (bc >= 4) pb = blk - 6;
                int xp = 0;
                if (pb >= 0) xp = *reinterpret_cast<const short *>(tile + pb * 128 + (pb & 7) * 16);
                const uint32_t dc = *reinterpret_cast<const unsigned short *>(mine + sw7 * 16);
                const uint32_t dc_entry = ENTRY_DC | tag | ((uint32_t)xp << 18) | dc;
                if (bm == 0 && bc >= 4) xl.unit_info[16 * (size_t)t + (bc - 4)] = (unsigned short)rank;
                if (bm == nm - 1 && bc >= 3) xl.unit_info[16 * (size_t)t + 2 + (bc - 3)] = (unsigned short)dc;
                if (bm == 0 && (bc == 0 || bc >= 4))
                    reinterpret_cast<uint32_t *>(xl.unit_info + 16 * (size_t)t + 6)[bc == 0 ? 0 : bc - 3] = dc_entry;
                out[rank++] = dc_entry;
                uint32_t m = mlo & ~1u;
                while (m) {
                    const uint32_t z = __ffs(m) - 1;
                    m &= m - 1;
                    out[rank++] = (uint32_t)*reinterpret_cast<const unsigned short *>(mine + (((z >> 3) ^ sw7) << 4) + ((z & 7) << 1)) |
                                  tag | (z << 18);
                }
                m = mhi;
                while (m) {
                    const uint32_t z = __ffs(m) + 31;
                    m &= m - 1;
                    out[rank++] = (uint32_t)*reinterpret_cast<const unsigned short *>(mine + (((z >> 3) ^ sw7) << 4) + ((z & 7) << 1)) |
                                  tag | (z << 18);
                }
            }
            // dead (zero) entries up to a whole 64-entry round: the pass reads rounds without masking tails
            for (uint32_t i = total + tid; i < ((total + 63u) & ~63u); i += XF_THREADS) out[i] = 0u;
            if (tid == 0) xl.ent_cnt[t] = total;
        }

        // ---- copy-out: TILE_BLOCKS x 128 B, consecutive lanes -> consecutive 16 B ----
        {
            const size_t base = ((size_t)my * g.mcu_cols + mcu0) * 6;
            uint4 *gout = reinterpret_cast<uint4 *>(coef) + base * 8;
#pragma unroll
            for (int i = 0; i < (TILE_BLOCKS * 8 + XF_THREADS - 1) / XF_THREADS; i++) {
                const int idx = tid + i * XF_THREADS;
                const int b = idx >> 3, kk = idx & 7;
                if (((TILE_BLOCKS * 8) % XF_THREADS == 0 || idx < TILE_BLOCKS * 8) && b < nm * 6) {
                    uint4 v = reinterpret_cast<const uint4 *>(tile + b * 128)[kk ^ (b & 7)];
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

}  // namespace icb
