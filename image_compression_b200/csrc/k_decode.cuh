// k_decode.cuh -- SURVEY section 8 row f2 (input staging): the JPEG reader in front of the JPG seam, on the GPU.
//
// Replaces reader.read(0, param) of core/ImageCompression.java:119-155 for the files that make up the reference's
// workload (baseline SOF0, 8 bit, YCbCr 4:2:0 -- also 4:2:2 and 4:4:4 --, one interleaved scan, restart intervals or not): the caller uploads the
// FILE (a few MB) instead of the decoded raster (36 MB for 12 MP), which is what bounds the end-to-end rate on a
// PCIe link.  Anything else is refused on the host (ICB_ERR_UNSUPPORTED: the caller keeps the stock ImageIO decode).
// Arithmetic = the JDK reader's libjavajpeg = IJG libjpeg 6b with its defaults [JDK-mem], restated in
// oracle/jpeg_decode_oracle.c and pinned there against libjpeg-turbo: jdhuff.c decode_mcu, jidctint.c
// jpeg_idct_islow with jddctmgr.c's dequantisation and the & RANGE_MASK range limit, jdsample.c
// h2v2_fancy_upsample with jdmainct.c's context rows, jdcolor.c ycc_rgb_convert.
//
// A Huffman stream has no entry points, so it is decoded the self-synchronising way (Weissenberger & Schmidt,
// ICPP 2018, restated from the paper): the unstuffed bit stream is cut into subsequences of DEC_SUB_BITS bits, one
// thread each.  Round 0 decodes every subsequence from a guessed state (symbol boundary, first block of an MCU) and
// each thread carries on into the following subsequences until it arrives in a state somebody already decoded
// from; round r > 0 repeats that for every subsequence whose left neighbour's exit state is not the one it was
// decoded from.  Thread 0 starts exact and correct records are never overwritten by wrong ones, so the exact
// prefix grows every round -- by as much as the code needs to re-synchronise (bit position, zigzag index and the
// block's place in the 4:2:0 MCU: a few thousand bits on photographs).  The fixed point is the sequential decode;
// a job that has not reached it after DEC_ROUNDS rounds reports ICB_ERR_UNSUPPORTED (host decode).  Then: block counts -> prefix -> a last
// decode that writes the coefficients, DC prediction as a scan over MCUs, IDCT into planes, upsample + colour.
// Kernels, all on the job's stream:
//   k_dec_scan     per 4 KB of file: stuffed zeros and RSTn markers counted, position of the first other marker
//   k_dec_offsets  prefix of those counts (one CTA)
//   k_dec_unstuff  the entropy-coded bytes without stuffed zeros and restart markers, zero padded; where each restart interval starts
//   k_dec_sync     one synchronisation round (DEC_ROUNDS launches, no-ops once nothing changes)
//   k_dec_blocks   blocks in front of every subsequence (one CTA); clears the blocks two neighbouring subsequences share
//   k_dec_write    coefficients (quantised, natural order, DC as differences) into the block array: whole blocks
//                  through 128 bytes of shared memory per thread as 16-byte vectors, shared blocks by 2-byte stores
//   k_dec_dc_local / k_dec_dc_cta   per-component prefix of the DC differences over the MCUs
//   k_dec_idct     DC prediction undone, dequantise + islow inverse DCT, one thread per block -> Y / Cb / Cr planes
//   k_dec_color    fancy h2v2 / h2v1 upsampling (or none, 4:4:4) + YCbCr -> RGB, eight pixels per thread -> the decoded raster
// The chain is captured once per slot as a CUDA graph (api.cu enqueue_decode): the kernels read their parameters from
// a block in device memory, grids follow the slot's capacities.
#pragma once

#include "common.cuh"

namespace icb {

constexpr int DEC_SUB_BITS = 4096;          // bits per subsequence (a multiple of 32); ICB_DEC_SUB overrides per process
constexpr int DEC_ROUNDS = 10;              // synchronisation rounds queued per job (<= 32: five bits of the record)
constexpr int DEC_CARRY = 4;                // subsequences a thread carries on through per round
constexpr int DEC_CHUNK = 4096;             // file bytes per CTA of the scan / unstuff kernels
constexpr int DEC_LUT_BITS = 10;
#ifndef ICB_DEC_THREADS
#define ICB_DEC_THREADS 256
#endif
constexpr int DEC_THREADS = ICB_DEC_THREADS;            // threads (= subsequences) per CTA of the decode kernels: they share 34 KB of tables

constexpr int DEC_LUT2 = 16;                // second-level tables per Huffman table (distinct 10-bit prefixes of long codes)

// One look-up entry says everything the decode loop needs about the symbol at the head of the window:
//   [5:0]  bits to skip = code length + value bits      [12:6] how far the zigzag index moves: 1 for a DC symbol,
//   [16:13] value bits (size)   [21:17] code length      run + 1 for a coefficient, 16 for ZRL, 64 for end of block
// 0 = not in the table; 0x80000000 | k = a code longer than DEC_LUT_BITS whose next six bits index second-level table k.
__host__ __device__ inline unsigned int dec_entry(int cls, unsigned int len, unsigned int sym)
{
    const unsigned int sz = sym & 15u, r = cls ? sym >> 4 : 0u;
    const unsigned int zadv = cls == 0 ? 1u : (sz ? r + 1u : (r == 15u ? 16u : 64u));
    return (len + sz) | (zadv << 6) | (sz << 13) | (len << 17);
}

// Huffman decode tables of one file, built on the host (jpeg_parse.h jpeg_build_tables): per table t = class * 2 + id
// (0, 1 DC; 2, 3 AC) a look-up by the next DEC_LUT_BITS bits, second-level tables for the longer codes, and jdhuff.c's
// maxcode / valoffset / huffval for whatever is in neither (more than DEC_LUT2 long prefixes, or no such code).
struct alignas(16) DecTables {
    unsigned int lut[4][1 << DEC_LUT_BITS];
    unsigned int lut2[4][DEC_LUT2][64];
    int maxcode[4][18];
    int valoff[4][18];
    unsigned char huffval[4][256];
    unsigned short qt[3][64];               // per component, natural order
    unsigned char tab_dc[6], tab_ac[6];     // table index (0..3) of block k of an MCU: Y Y Y Y Cb Cr
    unsigned char pad[4];
};

struct DecState {
    unsigned int scan_end;                  // file offset of the first marker after the scan header (atomicMin)
    unsigned int total_bytes;               // unstuffed entropy-coded bytes
    unsigned int active[DEC_ROUNDS];        // subsequences re-decoded in round r
    unsigned int end_bit;                   // bit position at which the last block of the image ended
    unsigned int blocks_found;              // blocks the whole stream holds
    unsigned int n_rst;                     // RSTn markers in the scan
    int error;                              // != 0: not decodable here
    int pad;
};

struct DecParams {
    const unsigned char *file;              // the file in HBM
    unsigned int file_len, scan_off;
    unsigned int *chunk_drop;               // [nchunks + 1] stuffed zeros per chunk, then their exclusive prefix
    unsigned char *bits;                    // unstuffed entropy-coded bytes (+ 64 zero bytes)
    unsigned long long *rec;                // [nsub_max] one record per subsequence (dec_rec)
    unsigned int sub_bits;                  // bits per subsequence
    unsigned int carry;                     // subsequences a thread carries on through per round
    int wbuf;                               // write pass: whole blocks through shared memory (1) or every coefficient straight to HBM (0, A/B)
    unsigned int *blk_base;                 // [nsub_max] blocks in front of each subsequence
    int *mcu_pref;                          // [mcus][3] DC differences in front of each MCU inside its CTA of k_dec_dc_local
    int *cta_pref;                          // [ceil(mcus / DEC_DC_THREADS)][3] ... and in front of that CTA
    int16_t *coef;                          // [nblocks][64] natural order
    unsigned char *planes;                  // Y [mcu_rows * 8 vs][mcu_cols * 8 hs], then Cb, Cr [mcu_rows * 8][mcu_cols * 8]
    unsigned char *rgb;                     // decoded raster
    int rgb_pitch, bgr;
    int w, h, mcu_cols, mcu_rows;
    int hs, vs, nb;                         // luma sampling (2x2 4:2:0, 2x1 4:2:2, 1x1 4:4:4) and blocks per MCU = hs * vs + 2
    unsigned int nblocks;
    unsigned int nsub_max;                  // subsequences the file could hold at most
    int nchunks;                            // DEC_CHUNK-byte pieces of the file from scan_off on
    int ri;                                 // restart interval in MCUs (DRI), 0 = none
    unsigned int *chunk_rst;                // [nchunks + 1] RSTn markers per chunk, then their exclusive prefix
    unsigned int *rst_pos;                  // [rst_cap] byte offset in `bits` at which the interval after marker k starts
    unsigned int rst_cap;
    const DecTables *tab;
    DecState *st;
};

// The coefficient array starts at zero (the decoder only stores non-zero values).  A kernel, not a memset node: how a
// memset of 36 MB is carried out is the driver's choice, and every job of every slot needs one.
__global__ void __launch_bounds__(256) k_dec_zero(uint4 *dst, size_t nvec)
{
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < nvec; i += (size_t)gridDim.x * 256) dst[i] = z;
}

// ---- entropy-coded segment: stuffed zeros and restart markers out, first other marker found -----------------
// What happens to byte i of the scan (prev / next = its neighbours): a zero after 0xFF is stuffing and goes; with a
// restart interval, 0xFF 0xD0..0xD7 is an RSTn marker and both bytes go; 0xFF followed by anything else ends the scan.
__device__ __forceinline__ bool dec_is_rst(unsigned int b, int ri) { return ri > 0 && b >= 0xD0u && b <= 0xD7u; }
__device__ __forceinline__ bool dec_keep(unsigned int prev, unsigned int b, unsigned int next, int ri)
{
    if (prev == 0xFFu && (b == 0u || dec_is_rst(b, ri))) return false;
    if (b == 0xFFu && dec_is_rst(next, ri)) return false;
    return true;
}

// A thread takes the 16 file bytes of one ALIGNED group (one 16-byte load; the file buffer is 256-byte aligned and the
// groups start at scan_off rounded down), plus the byte in front and the byte behind; bytes outside [scan_off, end) do
// not exist for it.  Chunk c = the 256 groups from group 256 c on.
struct DecGroup {
    unsigned int w[4];          // the 16 bytes, little-endian words
    unsigned int prev, next;    // neighbours (0 outside the scan)
    unsigned int first, last;   // byte indices (0..16) of the part of the group inside [scan_off, end)
    unsigned int addr;          // file offset of byte 0
};

__device__ __forceinline__ DecGroup dec_load_group(const DecParams &p, unsigned int chunk, unsigned int end)
{
    DecGroup g;
    g.addr = (p.scan_off & ~15u) + (chunk * 256u + threadIdx.x) * 16u;
    g.first = g.addr < p.scan_off ? p.scan_off - g.addr : 0u;
    g.last = g.addr >= end ? 0u : min(16u, end - g.addr);
    if (g.first > g.last) g.first = g.last;
    g.w[0] = g.w[1] = g.w[2] = g.w[3] = 0u;
    g.prev = g.next = 0u;
    if (g.last > g.first) {
        const uint4 v = *reinterpret_cast<const uint4 *>(p.file + g.addr);       // (the buffer is padded to whole groups)
        g.w[0] = v.x; g.w[1] = v.y; g.w[2] = v.z; g.w[3] = v.w;
        if (g.addr + g.first > p.scan_off) g.prev = p.file[g.addr + g.first - 1];
        if (g.addr + 16u < end) g.next = p.file[g.addr + 16u];
    }
    return g;
}
__device__ __forceinline__ unsigned int dec_group_byte(const DecGroup &g, unsigned int i) { return (g.w[i >> 2] >> (8u * (i & 3u))) & 0xffu; }

__global__ void __launch_bounds__(256) k_dec_scan(const DecParams *pp)
{
    const DecParams p = *pp;
    __shared__ unsigned int s_cnt[2];
    if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const DecGroup g = dec_load_group(p, blockIdx.x, p.file_len);
    unsigned int drops = 0, rsts = 0;
    {
        unsigned int prev = g.prev;
#pragma unroll
        for (unsigned int i = 0; i < 16; i++) {
            if (i < g.first || i >= g.last) continue;
            const unsigned int b = dec_group_byte(g, i);
            const unsigned int next = i + 1 < 16 ? (i + 1 < g.last ? dec_group_byte(g, i + 1) : 0u) : g.next;
            if (!dec_keep(prev, b, next, p.ri)) drops++;
            if (prev == 0xFFu && g.addr + i > p.scan_off) {
                if (dec_is_rst(b, p.ri)) rsts++;
                else if (b != 0u && b != 0xFFu) atomicMin(&p.st->scan_end, g.addr + i - 1);      // a marker: the scan ends here
            }
            prev = b;
        }
    }
    drops = __reduce_add_sync(0xffffffffu, drops);
    rsts = __reduce_add_sync(0xffffffffu, rsts);
    if ((threadIdx.x & 31) == 0) { if (drops) atomicAdd(&s_cnt[0], drops); if (rsts) atomicAdd(&s_cnt[1], rsts); }
    __syncthreads();
    if (threadIdx.x == 0) { p.chunk_drop[blockIdx.x] = s_cnt[0]; p.chunk_rst[blockIdx.x] = s_cnt[1]; }
}

__global__ void __launch_bounds__(1024) k_dec_offsets(const DecParams *pp)
{
    const DecParams p = *pp;
    const int nchunks = p.nchunks;
    __shared__ unsigned int s_part[2][1024];
    const int tid = threadIdx.x;
    const int per = (nchunks + 1023) / 1024;
    const int i0 = min(nchunks, tid * per), i1 = min(nchunks, i0 + per);
    unsigned int mine[2] = {0, 0};
    for (int i = i0; i < i1; i++) { mine[0] += p.chunk_drop[i]; mine[1] += p.chunk_rst[i]; }
    s_part[0][tid] = mine[0]; s_part[1][tid] = mine[1];
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        unsigned int t0 = 0, t1 = 0;
        if (tid >= o) { t0 = s_part[0][tid - o]; t1 = s_part[1][tid - o]; }
        __syncthreads();
        s_part[0][tid] += t0; s_part[1][tid] += t1;
        __syncthreads();
    }
    unsigned int run0 = s_part[0][tid] - mine[0], run1 = s_part[1][tid] - mine[1];
    for (int i = i0; i < i1; i++) {
        const unsigned int c0 = p.chunk_drop[i], c1 = p.chunk_rst[i];
        p.chunk_drop[i] = run0; p.chunk_rst[i] = run1;
        run0 += c0; run1 += c1;
    }
}

// The bytes that stay are first packed into the CTA's shared memory (each thread at its offset inside the chunk's
// output), then leave as 16-byte vectors: one store per 16 output bytes instead of sixteen.
__global__ void __launch_bounds__(256) k_dec_unstuff(const DecParams *pp)
{
    const DecParams p = *pp;
    __shared__ unsigned int s_warp[2][8];
    __shared__ __align__(16) unsigned char s_out[DEC_CHUNK + 16];
    const unsigned int end = min(p.st->scan_end, p.file_len);
    const unsigned int chunk_addr = (p.scan_off & ~15u) + blockIdx.x * DEC_CHUNK;
    if (chunk_addr >= end) return;
    const DecGroup g = dec_load_group(p, blockIdx.x, end);
    unsigned int keep = 0, rsts = 0;
    unsigned int kept_mask = 0, rst_mask = 0;
    {
        unsigned int prev = g.prev;
#pragma unroll
        for (unsigned int i = 0; i < 16; i++) {
            if (i < g.first || i >= g.last) continue;
            const unsigned int b = dec_group_byte(g, i);
            const unsigned int next = i + 1 < 16 ? (i + 1 < g.last ? dec_group_byte(g, i + 1) : 0u) : g.next;
            if (dec_keep(prev, b, next, p.ri)) { keep++; kept_mask |= 1u << i; }
            if (prev == 0xFFu && dec_is_rst(b, p.ri)) { rsts++; rst_mask |= 1u << i; }
            prev = b;
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned int ik = keep, ir = rsts;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, ik, o), u = __shfl_up_sync(0xffffffffu, ir, o);
        if (lane >= o) { ik += t; ir += u; }
    }
    if (lane == 31) { s_warp[0][warp] = ik; s_warp[1][warp] = ir; }
    __syncthreads();
    unsigned int off = ik - keep, roff = ir - rsts, cta_keep = 0;
    for (int w = 0; w < 8; w++) { if (w < warp) { off += s_warp[0][w]; roff += s_warp[1][w]; } cta_keep += s_warp[0][w]; }
    // output offset of the chunk = bytes of the scan in front of it minus what was dropped there
    const unsigned int in_front = chunk_addr > p.scan_off ? chunk_addr - p.scan_off : 0u;
    const unsigned int out0 = in_front - p.chunk_drop[blockIdx.x];
    const unsigned int shift = out0 & 15u;                      // the chunk's output starts `shift` bytes into a 16-byte vector
    unsigned int o = off, rk = p.chunk_rst[blockIdx.x] + roff;
#pragma unroll
    for (unsigned int i = 0; i < 16; i++) {
        if (kept_mask & (1u << i)) s_out[shift + o++] = (unsigned char)dec_group_byte(g, i);
        if (rst_mask & (1u << i)) { if (rk < p.rst_cap) p.rst_pos[rk] = out0 + o; rk++; }
    }
    __syncthreads();
    {
        // whole vectors from shared memory; the ragged first and last vector of the chunk in bytes (the neighbours own the rest)
        unsigned char *dst = p.bits + (out0 - shift);
        const unsigned int nbytes = shift + cta_keep, nvec = (nbytes + 15u) >> 4;
        for (unsigned int v = threadIdx.x; v < nvec; v += 256) {
            const unsigned int lo = v * 16u, hi = min(lo + 16u, nbytes);
            if (lo >= shift && hi == lo + 16u) {
                reinterpret_cast<uint4 *>(dst)[v] = reinterpret_cast<const uint4 *>(s_out)[v];
            } else {
                for (unsigned int k = max(lo, shift); k < hi; k++) dst[k] = s_out[k];
            }
        }
    }
    if (g.last > g.first && g.addr + g.last == end) {           // the thread that holds the last byte of the scan
        const unsigned int total = out0 + o;
        p.st->total_bytes = total;
        p.st->n_rst = rk;
        for (unsigned int i = 0; i < 64; i++) p.bits[total + i] = 0;
    }
}

// ---- Huffman decode of one subsequence ----------------------------------------------------------------
// Per subsequence ONE 64-bit record, written whole (atomicMin), so that entry state, exit state and block count
// always belong to the same decode:
//   [63:59] DEC_ROUNDS - 1 - round   [58:42] owner = the thread that wrote it     (priority: later rounds win,
//   [41:27] entry state              [26:12] exit state      [11:0] blocks completed    inside a round the leftmost)
// A state is 15 bits: bit offset past the subsequence's nominal start / end (6 bits; a symbol overshoots by < 32)
// and bz = block-of-MCU * 64 + zigzag index (9 bits).  12 bits of block count: a block takes at least 4 bits, so a
// subsequence of up to 8192 bits; 17 bits of owner: up to 131 071 subsequences per scan.
constexpr unsigned long long DEC_REC_NONE = ~0ull;
constexpr unsigned int DEC_STATE_NONE = 0x7fffu;

__device__ __forceinline__ unsigned long long dec_rec(int round, unsigned int owner, unsigned int entry, unsigned int exit_,
                                                      unsigned int nblk)
{
    return ((unsigned long long)(unsigned int)(DEC_ROUNDS - 1 - round) << 59) | ((unsigned long long)(owner & 0x1ffffu) << 42) |
           ((unsigned long long)entry << 27) | ((unsigned long long)exit_ << 12) | (unsigned long long)(nblk & 0xfffu);
}
__device__ __forceinline__ unsigned int dec_rec_entry(unsigned long long r) { return (unsigned int)(r >> 27) & 0x7fffu; }
__device__ __forceinline__ unsigned int dec_rec_exit(unsigned long long r) { return (unsigned int)(r >> 12) & 0x7fffu; }
__device__ __forceinline__ unsigned int dec_rec_nblk(unsigned long long r) { return (unsigned int)r & 0xfffu; }

// Bit reader: a 64-bit window whose top bit is stream bit `pos`; `have` valid bits.  One 32-bit word (big-endian
// byte order in memory) is pulled in through the read-only path whenever fewer than 32 bits are left, so a symbol
// (code <= 16 bits + value <= 15 bits) always sits in the top 32 bits.
struct DecBits {
    const unsigned int *words;
    unsigned long long win;
    unsigned int pos, have, next;          // next = index of the word `pre` holds
    unsigned int pre;                      // the word that enters the window next, already loaded: with 200 KB of an SM's
                                           // L1 taken as shared memory by six of these CTAs a refill is an L2 round trip;
                                           // asked for one refill ahead, it is back before it is needed

    __device__ __forceinline__ void open(const unsigned char *bits, unsigned int pbit)
    {
        words = reinterpret_cast<const unsigned int *>(bits);
        pos = pbit;
        next = (pbit >> 5) + 2;
        const unsigned int a = __byte_perm(__ldg(words + next - 2), 0, 0x0123), b = __byte_perm(__ldg(words + next - 1), 0, 0x0123);
        pre = __ldg(words + next);
        win = (((unsigned long long)a << 32) | b) << (pbit & 31u);
        have = 64u - (pbit & 31u);
    }
    __device__ __forceinline__ unsigned int peek() const { return (unsigned int)(win >> 32); }
    __device__ __forceinline__ void skip(unsigned int n)
    {
        win <<= n;
        pos += n;
        have -= n;
        if (have < 32u) {
            win |= (unsigned long long)__byte_perm(pre, 0, 0x0123) << (32u - have);
            have += 32u;
            next++;
            pre = __ldg(words + next);
        }
    }
};

__constant__ unsigned char c_dec_zz[64] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct alignas(16) DecSmem {
    unsigned int lut[4][1 << DEC_LUT_BITS];
    unsigned int lut2[4][DEC_LUT2][64];
    int maxcode[4][18];
    int valoff[4][18];
    unsigned char huffval[4][256];
    unsigned char tab_dc[8], tab_ac[8];
    unsigned char zz[64];
};

__device__ __forceinline__ void dec_load_tables(DecSmem &s, const DecTables *t)
{
    const uint4 *src = reinterpret_cast<const uint4 *>(t->lut);             // lut and lut2 are contiguous in both structs
    uint4 *dst = reinterpret_cast<uint4 *>(s.lut);
    constexpr int nvec = (4 * (1 << DEC_LUT_BITS) + 4 * DEC_LUT2 * 64) / 4;
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = src[i];
    for (int i = threadIdx.x; i < 4 * 18; i += blockDim.x) {
        (&s.maxcode[0][0])[i] = (&t->maxcode[0][0])[i];
        (&s.valoff[0][0])[i] = (&t->valoff[0][0])[i];
    }
    for (int i = threadIdx.x; i < 4 * 256; i += blockDim.x) (&s.huffval[0][0])[i] = (&t->huffval[0][0])[i];
    if (threadIdx.x < 6) { s.tab_dc[threadIdx.x] = t->tab_dc[threadIdx.x]; s.tab_ac[threadIdx.x] = t->tab_ac[threadIdx.x]; }
    if (threadIdx.x < 64) s.zz[threadIdx.x] = c_dec_zz[threadIdx.x];
    __syncthreads();
}

// the entry (dec_entry) of the symbol at the head of `peek` in table tb (jdhuff.c jpeg_huff_decode)
__device__ __forceinline__ unsigned int dec_lookup(const DecSmem &s, unsigned int tb, unsigned int peek)
{
    unsigned int e = s.lut[tb][peek >> (32 - DEC_LUT_BITS)];
    if (e & 0x80000000u) e = s.lut2[tb][e & 0xffu][(peek >> (32 - DEC_LUT_BITS - 6)) & 63u];
    if (e) return e;
    for (int l = DEC_LUT_BITS + 1; l <= 16; l++) {
        const int code = (int)(peek >> (32 - l));
        if (code <= s.maxcode[tb][l]) return dec_entry((int)(tb >> 1), (unsigned int)l, s.huffval[tb][(code + s.valoff[tb][l]) & 255]);
    }
    return dec_entry((int)(tb >> 1), 16u, 0u);                  // not a code: corrupt data, decoded as a zero like IJG
}

// Decodes from state (pbit, bz = block-of-MCU * 64 + zigzag index) until a symbol starts at or after end_bit.
// WRITE: coefficients go to coef (block index blk counts from the caller's base; blocks >= nblocks are dropped).
__device__ __forceinline__ unsigned int st_n_rst(const DecParams &p) { return p.st->n_rst; }

__device__ __forceinline__ void dec_zero_block(int16_t *coef, unsigned int blk)
{
    uint4 *q = reinterpret_cast<uint4 *>(coef + (size_t)blk * 64);
    const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int i = 0; i < 8; i++) q[i] = z;
}

template <bool WRITE>
__device__ __forceinline__ void dec_run(const DecSmem &s, const unsigned char *bits, unsigned int &pbit, unsigned int &bz,
                                        unsigned int &nblk, unsigned int end_bit, int16_t *coef, unsigned int blk,
                                        unsigned int nblocks, DecState *st, unsigned int last_b, const DecParams &p,
                                        unsigned int next_base = 0xffffffffu, int16_t *buf = nullptr)
{
    // WRITE with buf: the coefficients of a block this thread decodes from its first symbol to its last are collected
    // in the thread's 128 bytes of shared memory and leave as eight 16-byte stores (zeros included); only the block the
    // thread finds half done at its start and the one it leaves half done at its end -- the next thread's first:
    // next_base -- take 2-byte stores, into memory k_dec_blocks cleared.
    bool direct = true;
    unsigned int b = bz >> 6, z = bz & 63u;
    unsigned int tdc = s.tab_dc[b], tac = s.tab_ac[b];          // the block's two tables (change once per block)
    DecBits br;
    br.open(bits, pbit);
    // Restart intervals (DRI): the interval behind marker k starts at byte rst_pos[k], byte aligned, with the first
    // block of an MCU; the exact decoder ends the interval before it inside the last byte in front (padding).  A
    // decoder that stands there between two MCUs, or finds itself past the boundary, restarts AT the boundary -- which
    // also makes every boundary a point where a wrong chain becomes exact.
    unsigned int rk = 0, rnext = 0xffffffffu;
    const unsigned int n_rst = p.ri > 0 ? min(st_n_rst(p), p.rst_cap) : 0u;
    if (n_rst) {
        unsigned int lo = 0, hi = n_rst;                         // first k with rst_pos[k] * 8 > pbit
        while (lo < hi) { const unsigned int mid = (lo + hi) >> 1; if (p.rst_pos[mid] * 8u > pbit) hi = mid; else lo = mid + 1; }
        rk = lo;
        if (rk < n_rst) rnext = p.rst_pos[rk] * 8u;
    }
    while (br.pos < end_bit) {
        if (WRITE && blk >= nblocks) break;
        if (br.pos + 7u >= rnext && (br.pos >= rnext || (z == 0 && b == 0))) {
            br.open(bits, rnext);
            z = 0; b = 0;
            tdc = s.tab_dc[0]; tac = s.tab_ac[0];
            rk++;
            if (WRITE) {
                const unsigned int nbk = rk * (unsigned int)p.ri * (last_b + 1u);
                if (nbk != blk) {                               // (only a wrong chain gets here)
                    if (buf) direct = true;                     // whatever it holds is not a whole block: scattered from here on
                    else if (nbk < nblocks && nbk != next_base) dec_zero_block(coef, nbk);
                }
                blk = nbk;
            }
            rnext = rk < n_rst ? p.rst_pos[rk] * 8u : 0xffffffffu;
            continue;
        }
        const unsigned int peek = br.peek();
        const unsigned int e = dec_lookup(s, z == 0 ? tdc : tac, peek);
        const unsigned int zadv = (e >> 6) & 127u;
        if (WRITE) {
            const unsigned int sz = (e >> 13) & 15u, at = z + zadv - 1u;     // (a DC symbol: z = 0, zadv = 1 -> 0)
            if (sz && at < 64u) {
                // the value bits follow the code inside the same 32-bit window (jdhuff.c HUFF_EXTEND)
                const unsigned int rb = (peek << ((e >> 17) & 31u)) >> (32u - sz);
                const int v = rb < (1u << (sz - 1)) ? (int)rb - (int)(1u << sz) + 1 : (int)rb;
                if (buf && !direct) buf[s.zz[at]] = (int16_t)v;
                else coef[(size_t)blk * 64 + s.zz[at]] = (int16_t)v;
            }
        }
        br.skip(e & 63u);
        z += zadv;                                               // coefficient / ZRL / end of block (64) alike
        if (z >= 64u) {                                          // block complete
            z = 0;
            b = b == last_b ? 0 : b + 1;
            tdc = s.tab_dc[b]; tac = s.tab_ac[b];
            nblk++;
            blk++;
            if (WRITE) {
                if (buf) {
                    if (!direct) {                              // a whole block: out it goes, and the buffer is clean again
                        uint4 *q = reinterpret_cast<uint4 *>(coef + (size_t)(blk - 1) * 64), *bq = reinterpret_cast<uint4 *>(buf);
                        const uint4 zero4 = make_uint4(0, 0, 0, 0);
#pragma unroll
                        for (int k8 = 0; k8 < 8; k8++) { q[k8] = bq[k8]; bq[k8] = zero4; }
                    }
                    direct = blk == next_base;
                } else if (blk < nblocks && blk != next_base) {
                    dec_zero_block(coef, blk);                  // (unbuffered form: clear each block as it starts)
                }
                if (blk == nblocks) st->end_bit = br.pos;
            }
        }
    }
    pbit = br.pos;
    bz = (b << 6) | z;
}

// One synchronisation round.  A thread whose entry state (the exit state of its left neighbour; thread 0: the
// start of the scan) differs from the one its record was decoded from decodes its subsequence again and carries on
// into the following ones -- at most DEC_CARRY of them per round -- until it meets a record that was decoded from
// exactly the state it arrives with: from there on the records are already what this thread would write.
__global__ void __launch_bounds__(DEC_THREADS) k_dec_sync(const DecParams *pp, int round)
{
    const DecParams p = *pp;
    __shared__ DecSmem s;
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int total_bits = p.st->total_bytes * 8u;
    const unsigned int S = p.sub_bits;
    const unsigned int nsub = min(p.nsub_max, (total_bits + S - 1) / S);
    unsigned int entry = DEC_STATE_NONE;
    bool active = false;
    if (i < nsub) {
        if (i == 0) entry = 0;                                   // offset 0, first block of an MCU, DC next
        else if (round == 0) entry = 0;                          // the guess
        else entry = dec_rec_exit(p.rec[i - 1]);
        active = round == 0 || entry != dec_rec_entry(p.rec[i]);
    }
    if (!__syncthreads_or(active ? 1 : 0)) return;
    dec_load_tables(s, p.tab);
    if (!active) return;
    if (round > 0) p.st->active[round] = 1;
    for (unsigned int k = i; k < nsub && k < i + p.carry; k++) {
        if (k > i && dec_rec_entry(p.rec[k]) == entry) break;   // joined what is already recorded
        const unsigned int start = k * S, end = min(start + S, total_bits);
        unsigned int pbit = start + (entry >> 9), bz = entry & 0x1ffu, nblk = 0;
        dec_run<false>(s, p.bits, pbit, bz, nblk, end, nullptr, 0, 0, nullptr, (unsigned int)p.nb - 1u, p);
        const unsigned int exit_ = (min(pbit - end, 63u) << 9) | bz;
        atomicMin(&p.rec[k], dec_rec(round, i, entry, exit_, nblk));
        entry = exit_;
    }
}

// blocks in front of every subsequence (exclusive prefix of the per-subsequence block counts), one CTA
__global__ void __launch_bounds__(1024) k_dec_blocks(const DecParams *pp)
{
    const DecParams p = *pp;
    __shared__ unsigned int s_part[1024];
    const int tid = threadIdx.x;
    const unsigned int total_bits = p.st->total_bytes * 8u;
    const unsigned int nsub = min(p.nsub_max, (total_bits + p.sub_bits - 1) / p.sub_bits);
    const unsigned int per = (nsub + 1023) / 1024;
    const unsigned int i0 = min(nsub, tid * per), i1 = min(nsub, i0 + per);
    unsigned int mine = 0;
    for (unsigned int i = i0; i < i1; i++) mine += dec_rec_nblk(p.rec[i]);
    s_part[tid] = mine;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const unsigned int t = tid >= o ? s_part[tid - o] : 0u;
        __syncthreads();
        s_part[tid] += t;
        __syncthreads();
    }
    unsigned int run = s_part[tid] - mine;
    for (unsigned int i = i0; i < i1; i++) {
        p.blk_base[i] = run;
        // the block in progress where a subsequence starts is shared with the thread to its left: zeroed here, once
        if (run < p.nblocks && (i == i0 || run != p.blk_base[i - 1])) dec_zero_block(p.coef, run);
        run += dec_rec_nblk(p.rec[i]);
    }
    if (tid == 1023) p.st->blocks_found = s_part[1023];
}

constexpr int DEC_WTHREADS = 128;           // threads per CTA of the write pass
constexpr int DEC_WBUF_STRIDE = 144;        // bytes of block buffer per thread: 128 + 16, so that the 16-byte reads of a warp spread over the banks
constexpr int DEC_WBUF_BYTES = DEC_WTHREADS * DEC_WBUF_STRIDE;

__global__ void __launch_bounds__(DEC_WTHREADS) k_dec_write(const DecParams *pp)
{
    const DecParams p = *pp;
    __shared__ DecSmem s;
    extern __shared__ __align__(16) unsigned char dec_wbuf[];
    dec_load_tables(s, p.tab);
    int16_t *buf = reinterpret_cast<int16_t *>(dec_wbuf + threadIdx.x * DEC_WBUF_STRIDE);
    {
        const uint4 zero4 = make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int k8 = 0; k8 < 8; k8++) reinterpret_cast<uint4 *>(buf)[k8] = zero4;
    }
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int total_bits = p.st->total_bytes * 8u;
    const unsigned int S = p.sub_bits;
    const unsigned int start = i * S;
    if (i >= p.nsub_max || start >= total_bits) return;
    const unsigned int entry = dec_rec_entry(p.rec[i]);
    unsigned int pbit = start + (entry >> 9), bz = entry & 0x1ffu, nblk = 0;
    const unsigned int end = min(start + S, total_bits);
    const unsigned int nsub = min(p.nsub_max, (total_bits + S - 1) / S);
    const unsigned int next_base = i + 1 < nsub ? p.blk_base[i + 1] : 0xffffffffu;
    dec_run<true>(s, p.bits, pbit, bz, nblk, end, p.coef, p.blk_base[i], p.nblocks, p.st, (unsigned int)p.nb - 1u, p, next_base,
                  p.wbuf ? buf : nullptr);
}

// ---- DC differences -> DC values (jdhuff.c last_dc_val): per component a prefix over the MCUs --------------
// k_dec_dc_local: one thread per MCU, the sums of its DC differences (Y: four blocks) scanned inside the CTA;
// k_dec_dc_cta: the CTA totals scanned by one CTA; the IDCT kernel adds CTA prefix + MCU prefix + the
// differences of the blocks in front of its own inside the MCU.
constexpr int DEC_DC_THREADS = 256;

__global__ void __launch_bounds__(DEC_DC_THREADS) k_dec_dc_local(const DecParams *pp)
{
    const DecParams p = *pp;
    __shared__ int s_w[3][DEC_DC_THREADS / 32];
    const unsigned int nm = p.nblocks / (unsigned int)p.nb;
    const unsigned int m = blockIdx.x * DEC_DC_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ny = p.nb - 2;
    int v[3] = {0, 0, 0};
    if (m < nm) {
        const int16_t *c = p.coef + (size_t)m * p.nb * 64;
        for (int j = 0; j < ny; j++) v[0] += c[64 * j];
        v[1] = c[64 * ny];
        v[2] = c[64 * ny + 64];
    }
    int incl[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        int x = v[k];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += t;
        }
        incl[k] = x;
        if (lane == 31) s_w[k][warp] = x;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 3; k++) {
        int base = 0, tot = 0;
        for (int w = 0; w < DEC_DC_THREADS / 32; w++) { if (w < warp) base += s_w[k][w]; tot += s_w[k][w]; }
        if (m < nm) p.mcu_pref[(size_t)m * 3 + k] = base + incl[k] - v[k];
        if (threadIdx.x == 0) p.cta_pref[(size_t)blockIdx.x * 3 + k] = tot;
    }
}

__global__ void __launch_bounds__(1024) k_dec_dc_cta(const DecParams *pp)
{
    const DecParams p = *pp;
    const int nctas = (int)((p.nblocks / (unsigned int)p.nb + DEC_DC_THREADS - 1) / DEC_DC_THREADS);
    __shared__ int s_part[3][1024];
    const int tid = threadIdx.x;
    const int per = (nctas + 1023) / 1024;
    const int i0 = min(nctas, tid * per), i1 = min(nctas, i0 + per);
    int mine[3] = {0, 0, 0};
    for (int i = i0; i < i1; i++)
        for (int k = 0; k < 3; k++) mine[k] += p.cta_pref[(size_t)i * 3 + k];
    for (int k = 0; k < 3; k++) s_part[k][tid] = mine[k];
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        int t[3] = {0, 0, 0};
        if (tid >= o) for (int k = 0; k < 3; k++) t[k] = s_part[k][tid - o];
        __syncthreads();
        for (int k = 0; k < 3; k++) s_part[k][tid] += t[k];
        __syncthreads();
    }
    int run[3];
    for (int k = 0; k < 3; k++) run[k] = s_part[k][tid] - mine[k];
    for (int i = i0; i < i1; i++)
        for (int k = 0; k < 3; k++) { const int c = p.cta_pref[(size_t)i * 3 + k]; p.cta_pref[(size_t)i * 3 + k] = run[k]; run[k] += c; }
}

// ---- jidctint.c jpeg_idct_islow ----------------------------------------------------------------------
__device__ __forceinline__ void dec_idct_1d(const int (&in)[8], int (&out)[8], int shift)
{
    int z1, z2, z3, z4, z5, tmp0, tmp1, tmp2, tmp3, tmp10, tmp11, tmp12, tmp13;
    z2 = in[2]; z3 = in[6];
    z1 = (z2 + z3) * 4433;
    tmp2 = z1 + z3 * (-15137);
    tmp3 = z1 + z2 * 6270;
    z2 = in[0]; z3 = in[4];
    tmp0 = (z2 + z3) << 13;
    tmp1 = (z2 - z3) << 13;
    tmp10 = tmp0 + tmp3; tmp13 = tmp0 - tmp3; tmp11 = tmp1 + tmp2; tmp12 = tmp1 - tmp2;
    tmp0 = in[7]; tmp1 = in[5]; tmp2 = in[3]; tmp3 = in[1];
    z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2; z4 = tmp1 + tmp3;
    z5 = (z3 + z4) * 9633;
    tmp0 *= 2446; tmp1 *= 16819; tmp2 *= 25172; tmp3 *= 12299;
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
    z3 += z5; z4 += z5;
    tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
    const int rnd = 1 << (shift - 1);
    out[0] = (tmp10 + tmp3 + rnd) >> shift; out[7] = (tmp10 - tmp3 + rnd) >> shift;
    out[1] = (tmp11 + tmp2 + rnd) >> shift; out[6] = (tmp11 - tmp2 + rnd) >> shift;
    out[2] = (tmp12 + tmp1 + rnd) >> shift; out[5] = (tmp12 - tmp1 + rnd) >> shift;
    out[3] = (tmp13 + tmp0 + rnd) >> shift; out[4] = (tmp13 - tmp0 + rnd) >> shift;
}

// jdmaster.c range-limit table at index x & RANGE_MASK (1023): with y = the low ten bits of x read as a SIGNED number
// (-512 .. 511) the table is clamp(y + 128, 0, 255) -- 0..127 -> +128, 128..511 -> 255, -512..-129 -> 0, -128..-1 -> +128
__device__ __forceinline__ unsigned int dec_range_limit(int x)
{
    const int y = (x << 22) >> 22;
    return (unsigned int)min(max(y + 128, 0), 255);
}

// one thread = one block: 64 coefficients (natural order) -> 8 x 8 samples of the block's plane
__global__ void __launch_bounds__(128) k_dec_idct(const DecParams *pp)
{
    const DecParams p = *pp;
    __shared__ short s_qt[3][64];
    for (int i = threadIdx.x; i < 192; i += blockDim.x) (&s_qt[0][0])[i] = (short)(&p.tab->qt[0][0])[i];
    __syncthreads();
    const unsigned int blk = blockIdx.x * blockDim.x + threadIdx.x;
    if (blk >= p.nblocks) return;
    const unsigned int nbm = (unsigned int)p.nb, ny = nbm - 2u;
    const unsigned int m = blk / nbm, k = blk - m * nbm;
    const int mx = (int)(m % (unsigned int)p.mcu_cols), my = (int)(m / (unsigned int)p.mcu_cols);
    const int comp = k < ny ? 0 : (int)(k - ny) + 1;
    const int yw = p.mcu_cols * 8 * p.hs, cw = p.mcu_cols * 8;
    unsigned char *dst;
    int pitch;
    if (k < ny) {
        const int ky = p.hs == 2 ? (int)(k >> 1) : (int)k, kx = p.hs == 2 ? (int)(k & 1u) : 0;     // (hs == 1: one block column)
        dst = p.planes + (size_t)(my * 8 * p.vs + (p.vs == 2 ? ky : 0) * 8) * yw + mx * 8 * p.hs + kx * 8;
        pitch = yw;
    } else {
        const size_t ysz = (size_t)yw * p.mcu_rows * 8 * p.vs, csz = (size_t)cw * p.mcu_rows * 8;
        dst = p.planes + ysz + (k == ny + 1u ? csz : 0) + (size_t)(my * 8) * cw + mx * 8;
        pitch = cw;
    }
    const uint4 *src = reinterpret_cast<const uint4 *>(p.coef + (size_t)blk * 64);
    int ws[64];
    // columns first (pass 1), on dequantised coefficients; rows of the input are 8 shorts = one uint4
    short cf[64];
#pragma unroll
    for (int r = 0; r < 8; r++) {
        const uint4 v = src[r];
        const unsigned int w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; j++) {
            cf[8 * r + 2 * j] = (short)(w4[j] & 0xffffu);
            cf[8 * r + 2 * j + 1] = (short)(w4[j] >> 16);
        }
    }
    {   // DC value = differences of every earlier block of the component + its own (stored back as the JCOEF short)
        int dc = p.cta_pref[(size_t)(m / DEC_DC_THREADS) * 3 + comp] + p.mcu_pref[(size_t)m * 3 + comp];
        if (p.ri > 0) {                                     // the predictions start again at every restart interval
            const unsigned int hm = m / (unsigned int)p.ri * (unsigned int)p.ri;
            dc -= p.cta_pref[(size_t)(hm / DEC_DC_THREADS) * 3 + comp] + p.mcu_pref[(size_t)hm * 3 + comp];
        }
        const int16_t *mc = p.coef + (size_t)m * nbm * 64;
        if (k < ny) { for (unsigned int j = 0; j <= k; j++) dc += mc[64 * j]; }
        else dc += mc[64 * k];
        cf[0] = (short)dc;
    }
#pragma unroll
    for (int c = 0; c < 8; c++) {
        int in[8], o[8];
#pragma unroll
        for (int r = 0; r < 8; r++) in[r] = (int)cf[8 * r + c] * (int)s_qt[comp][8 * r + c];
        dec_idct_1d(in, o, 11);
#pragma unroll
        for (int r = 0; r < 8; r++) ws[8 * r + c] = o[r];
    }
#pragma unroll
    for (int r = 0; r < 8; r++) {
        int in[8], o[8];
#pragma unroll
        for (int c = 0; c < 8; c++) in[c] = ws[8 * r + c];
        dec_idct_1d(in, o, 18);
        uint2 out;
        out.x = dec_range_limit(o[0]) | (dec_range_limit(o[1]) << 8) | (dec_range_limit(o[2]) << 16) | (dec_range_limit(o[3]) << 24);
        out.y = dec_range_limit(o[4]) | (dec_range_limit(o[5]) << 8) | (dec_range_limit(o[6]) << 16) | (dec_range_limit(o[7]) << 24);
        *reinterpret_cast<uint2 *>(dst + (size_t)r * pitch) = out;
    }
}

// ---- jdsample.c h2v2_fancy_upsample + jdcolor.c ycc_rgb_convert: one thread = 8 pixels of a row ----------
// HS, VS = luma sampling: 2,2 h2v2 fancy (context rows); 2,1 h2v1 fancy; 1,1 no upsampling (fullsize_upsample)
template <int HS, int VS>
__global__ void __launch_bounds__(256) k_dec_color(const DecParams *pp)
{
    const DecParams p = *pp;
    const int x0 = (blockIdx.x * 32 + (threadIdx.x & 31)) * 8;
    const int y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x0 >= p.w || y >= p.h) return;
    const int yw = p.mcu_cols * 8 * HS, cw = p.mcu_cols * 8;
    const size_t ysz = (size_t)yw * p.mcu_rows * 8 * VS, csz = (size_t)cw * p.mcu_rows * 8;
    const unsigned char *Y = p.planes + (size_t)y * yw + x0;
    const int dsw = (p.w + HS - 1) / HS, dsh = (p.h + VS - 1) / VS;
    const int cy = VS == 2 ? y >> 1 : y;
    int cn = (y & 1) ? cy + 1 : cy - 1;                      // jdmainct.c context rows: clamped to the real rows
    cn = max(0, min(cn, dsh - 1));
    const bool fancy = dsw > 2;
    const int cx0 = HS == 2 ? x0 >> 1 : x0;
    int cb[8], cr[8];
#pragma unroll
    for (int pl = 0; pl < 2; pl++) {
        const unsigned char *P = p.planes + ysz + (pl ? csz : 0);
        const unsigned char *r0 = P + (size_t)cy * cw, *r1 = P + (size_t)cn * cw;
        int *res = pl ? cr : cb;
        if (HS == 1) {
#pragma unroll
            for (int i = 0; i < 8; i++) res[i] = r0[min(cx0 + i, cw - 1)];
        } else if (!fancy) {
#pragma unroll
            for (int i = 0; i < 8; i++) res[i] = r0[min(cx0 + (i >> 1), cw - 1)];
        } else if (VS == 1) {
            // jdsample.c h2v1_fancy_upsample: 3/4 of the nearer sample + 1/4 of the further one, rounding 1 towards the left
            // neighbour and 2 towards the right; the first and the last output column copy their sample
            int sv[6];
#pragma unroll
            for (int j = 0; j < 6; j++) sv[j] = r0[min(max(cx0 - 1 + j, 0), cw - 1)];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int cx = cx0 + j, cur = sv[j + 1];
                res[2 * j] = cx == 0 ? cur : (3 * cur + sv[j] + 1) >> 2;
                res[2 * j + 1] = cx == dsw - 1 ? cur : (3 * cur + sv[j + 2] + 2) >> 2;
            }
        } else {
            int sum[6];                                       // 3 * cur + near for columns cx0 - 1 .. cx0 + 4
#pragma unroll
            for (int j = 0; j < 6; j++) {
                const int cx = min(max(cx0 - 1 + j, 0), cw - 1);
                sum[j] = 3 * r0[cx] + r1[cx];
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int cx = cx0 + j, cur = sum[j + 1];
                res[2 * j] = cx == 0 ? (cur * 4 + 8) >> 4 : (cur * 3 + sum[j] + 8) >> 4;
                res[2 * j + 1] = cx == dsw - 1 ? (cur * 4 + 7) >> 4 : (cur * 3 + sum[j + 2] + 7) >> 4;
            }
        }
    }
    const uint2 yv = *reinterpret_cast<const uint2 *>(Y);
    unsigned int c[24];                                      // the 24 output bytes, one per register
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int yy = (int)(((i < 4 ? yv.x : yv.y) >> (8 * (i & 3))) & 0xffu);
        const int cbx = cb[i] - 128, crx = cr[i] - 128;
        const int r = min(max(yy + ((91881 * crx + 32768) >> 16), 0), 255);
        const int b = min(max(yy + ((116130 * cbx + 32768) >> 16), 0), 255);
        const int g = min(max(yy + ((-22554 * cbx + 32768 - 46802 * crx) >> 16), 0), 255);
        c[3 * i] = (unsigned int)(p.bgr ? b : r); c[3 * i + 1] = (unsigned int)g; c[3 * i + 2] = (unsigned int)(p.bgr ? r : b);
    }
    unsigned int wv[6];
#pragma unroll
    for (int j = 0; j < 6; j++) wv[j] = c[4 * j] | (c[4 * j + 1] << 8) | (c[4 * j + 2] << 16) | (c[4 * j + 3] << 24);
    unsigned char *o = p.rgb + (size_t)y * p.rgb_pitch + (size_t)x0 * 3;
    if (x0 + 8 <= p.w) {
        uint2 *o8 = reinterpret_cast<uint2 *>(o);
        o8[0] = make_uint2(wv[0], wv[1]); o8[1] = make_uint2(wv[2], wv[3]); o8[2] = make_uint2(wv[4], wv[5]);
    } else {
        const int nb = (p.w - x0) * 3;                       // ragged right edge: byte stores, fully unrolled
#pragma unroll
        for (int i = 0; i < 21; i++)
            if (i < nb) o[i] = (unsigned char)(wv[i >> 2] >> (8 * (i & 3)));
    }
}

}  // namespace icb
