// k_png.cuh -- SURVEY section 8 row f3: the PNG file that follows the resize kernel, on the GPU.
//
// Replaces ImageIO.write(resized, "png", file) of core/ImageCompressionPng.java:70, i.e. the JDK's
// PNGImageWriter: per-row filter choice by the minimum sum of absolute filtered values (the libpng
// heuristic the JDK writer also uses [JDK-mem]), a zlib stream, and IDAT chunks of 8192 bytes like the
// JDK's IDATOutputStream [JDK-mem].  A zlib stream is not unique, so the parity bar of this row is:
// the file decodes (any inflater) to exactly the input pixels, and its size is within a stated
// tolerance of zlib's on the same filtered bytes (tests/test_gpu_png.py).
//
// Deflate here = ONE dynamic-Huffman block for the whole image: literals plus distance-1 run matches
// found inside 32-byte windows (flat areas; natural images after PNG filtering are entropy-, not
// match-compressible).  Pipeline, all on one stream:
//   k_png_filter   one warp per row: five filter sums, pick, write the filtered row
//   k_png_hist     one warp per 1 KB piece of the filtered stream: tokens -> symbol histogram, Adler-32 partials
//   k_png_tree     one thread: length-limited Huffman code lengths (15 / 7 bits), canonical codes,
//                  the block header, the Adler-32 of the whole stream
//   k_png_encode   one warp per piece: tokens -> codes -> a piece-local bit stream (LSB first)
//   k_png_scan     bit prefix over the pieces
//   k_png_merge    one warp per piece: its bits at their absolute bit offset, written as bytes
//                  straight into the IDAT chunks of the file
//   k_png_finish   IHDR / IEND, one thread per IDAT chunk for its length, type and CRC-32
// Byte and bit work throughout, HBM/L2 traffic of a few passes over <= 8 MB: latency, not bandwidth.
#pragma once

#include "common.cuh"

namespace icb {

constexpr int PNG_PIECE = 1024;             // filtered bytes per warp and pass
constexpr int PNG_PIECE_WORDS = 1024;       // piece-local bit stream: the last piece has up to 2047 bytes x 15 bits + EOB < 4096 B
constexpr int PNG_IDAT = 8192;              // data bytes per IDAT chunk
constexpr int PNG_NLIT = 286;
constexpr int PNG_HDR_WORDS = 128;          // zlib header + block header bits: <= 16 + 17 + 57 + 287 * 7 bits

struct PngParams {
    const unsigned char *src;               // resized pixels in HBM
    int w, h, pitch;
    int sbpp;                               // bytes per source pixel
    int pbpp;                               // bytes per PNG pixel: 1 gray, 3 RGB, 4 RGBA
    int perm;                               // source byte of PNG channel c = (perm >> (4 * c)) & 15
    unsigned char *filt;                    // [h][1 + w * pbpp] filtered rows
    long long nfilt;                        // bytes of filt
    int npieces;                            // ceil-ish(nfilt / PNG_PIECE): the last piece takes the remainder
    unsigned int *hist;                     // [288 + 32] symbol counts (lit/len, then dist)
    uint2 *adler;                           // [npieces] Adler partials (sum, weighted sum)
    uint32_t *codes;                        // [288 + 32] (bit-reversed code << 8) | length
    uint32_t *seg_bits;                     // [(npieces + 1)][PNG_PIECE_WORDS] LSB-first bit streams; segment 0 = headers
    unsigned int *seg_len;                  // [npieces + 1] bits per segment
    unsigned long long *seg_pos;            // [npieces + 2] exclusive bit prefix; [npieces + 1] = total bits
    unsigned int *adler_out;                // [1]
    unsigned char *out;                     // the PNG file
    long long out_cap;
    unsigned long long *out_len;            // [1] file bytes
};

// first byte and length of piece i: equal pieces, the last one also takes the remainder (< PNG_PIECE bytes)
__device__ __forceinline__ void png_piece_span(const PngParams &p, int i, long long &b0, int &n)
{
    b0 = (long long)i * PNG_PIECE;
    n = i == p.npieces - 1 ? (int)(p.nfilt - b0) : PNG_PIECE;
}

__device__ __forceinline__ int png_paeth(int a, int b, int c)
{
    const int pa = abs(b - c), pb = abs(a - c), pc = abs(a + b - 2 * c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// ---- rows -> filtered rows ----------------------------------------------------------------
// One CTA per row (256 threads walk its pixels).  PNG sample order is R G B [A]; the source raster may be
// BGR / ABGR / BGRA (perm).
__global__ void __launch_bounds__(256) k_png_filter(const PngParams p)
{
    __shared__ uint32_t s_sum[8][5];
    __shared__ int s_best;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int y = blockIdx.x;
    const unsigned char *cur = p.src + (size_t)y * p.pitch;
    const unsigned char *up = y > 0 ? cur - p.pitch : nullptr;
    const int rowbytes = p.w * p.pbpp;
    unsigned char *o = p.filt + (size_t)y * (rowbytes + 1);
    uint32_t sum[5] = {0, 0, 0, 0, 0};
    auto sample = [&](const unsigned char *row, int x, int c) -> int {
        return (row && x >= 0) ? (int)row[(size_t)x * p.sbpp + ((p.perm >> (4 * c)) & 15)] : 0;
    };
    auto mag = [](int v) -> uint32_t { const int s = (int)(signed char)(v & 0xff); return (uint32_t)abs(s); };
    for (int x = tid; x < p.w; x += 256) {
        for (int c = 0; c < p.pbpp; c++) {
            const int v = sample(cur, x, c), a = sample(cur, x - 1, c), b = sample(up, x, c), cc = sample(up, x - 1, c);
            sum[0] += mag(v);
            sum[1] += mag(v - a);
            sum[2] += mag(v - b);
            sum[3] += mag(v - ((a + b) >> 1));
            sum[4] += mag(v - png_paeth(a, b, cc));
        }
    }
#pragma unroll
    for (int f = 0; f < 5; f++) {
        const uint32_t t = __reduce_add_sync(0xffffffffu, sum[f]);
        if (lane == 0) s_sum[warp][f] = t;
    }
    __syncthreads();
    if (tid == 0) {
        int best = 0;
        uint32_t best_sum = 0;
        for (int f = 0; f < 5; f++) {
            uint32_t t = 0;
            for (int w = 0; w < 8; w++) t += s_sum[w][f];
            if (f == 0 || t < best_sum) { best = f; best_sum = t; }     // strict <: ties keep the lower filter type
        }
        s_best = best;
        o[0] = (unsigned char)best;
    }
    __syncthreads();
    const int best = s_best;
    for (int x = tid; x < p.w; x += 256) {
        for (int c = 0; c < p.pbpp; c++) {
            const int v = sample(cur, x, c), a = sample(cur, x - 1, c), b = sample(up, x, c), cc = sample(up, x - 1, c);
            int pr = 0;
            if (best == 1) pr = a; else if (best == 2) pr = b; else if (best == 3) pr = (a + b) >> 1;
            else if (best == 4) pr = png_paeth(a, b, cc);
            o[1 + x * p.pbpp + c] = (unsigned char)(v - pr);
        }
    }
}

// ---- tokens -------------------------------------------------------------------------------
// A window = 32 consecutive filtered bytes, one per lane.  eq = lanes whose byte equals the byte before
// it (lane 0 looks at the last byte of the previous window of the same piece).  A maximal stretch of
// m >= 3 set bits becomes ONE match (length m, distance 1) emitted by its first lane; every other byte
// is a literal.  Returns this lane's token: kind 0 none, 1 literal, 2 match (len = match length).
__device__ __forceinline__ int png_token(uint32_t eq, int lane, int &len)
{
    len = 0;
    const uint32_t me = 1u << lane;
    if (!(eq & me)) return 1;
    const uint32_t below = ~eq & (me - 1u);                         // zero bits under this lane
    const int start = below ? 32 - __clz(below) : 0;                // first lane of this stretch
    const uint32_t above = ~eq & ~((me << 1) - 1u);                 // zero bits over this lane
    const int end = above ? __ffs(above) - 1 : 32;                  // one past its last lane
    const int m = end - start;
    if (m < 3) return 1;
    if (lane != start) return 0;
    len = m;
    return 2;
}

// deflate length symbol and extra bits for a match length 3..32
__device__ __forceinline__ void png_len_code(int len, int &sym, int &ebits, int &eval)
{
    if (len <= 10) { sym = 254 + len; ebits = 0; eval = 0; return; }
    if (len <= 18) { sym = 265 + ((len - 11) >> 1); ebits = 1; eval = (len - 11) & 1; return; }
    if (len <= 34) { sym = 269 + ((len - 19) >> 2); ebits = 2; eval = (len - 19) & 3; return; }
    sym = 273; ebits = 3; eval = 0;                                 // not reached: windows are 32 bytes
}

// eq mask of one window of a piece (bytes b0 + off .. + 31, n bytes in the piece)
__device__ __forceinline__ uint32_t png_window(const PngParams &p, long long b0, int n, int off, int lane, int &byte,
                                               bool &valid)
{
    const int i = off + lane;
    valid = i < n;
    byte = valid ? (int)p.filt[b0 + i] : -1;
    int prev = __shfl_up_sync(0xffffffffu, byte, 1);
    if (lane == 0) prev = off > 0 ? (int)p.filt[b0 + off - 1] : -2;
    return __ballot_sync(0xffffffffu, valid && byte == prev);
}

__global__ void __launch_bounds__(256) k_png_hist(const PngParams p)
{
    __shared__ unsigned int s_hist[288 + 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 288 + 32; i += 256) s_hist[i] = 0;
    __syncthreads();
    const int piece = blockIdx.x * 8 + warp;
    if (piece < p.npieces) {
        long long b0;
        int n;
        png_piece_span(p, piece, b0, n);
        uint32_t sa = 0, sb = 0;
        for (int off = 0; off < n; off += 32) {
            int byte;
            bool valid;
            const uint32_t eq = png_window(p, b0, n, off, lane, byte, valid);
            if (valid) {
                sa += (uint32_t)byte;
                sb += (uint32_t)(n - (off + lane)) * (uint32_t)byte;
            }
            int len;
            const int kind = valid ? png_token(eq, lane, len) : 0;
            if (kind == 1) atomicAdd(&s_hist[byte], 1u);
            else if (kind == 2) {
                int sym, eb, ev;
                png_len_code(len, sym, eb, ev);
                atomicAdd(&s_hist[sym], 1u);
                atomicAdd(&s_hist[288], 1u);
            }
        }
        sa = __reduce_add_sync(0xffffffffu, sa);
        sb = __reduce_add_sync(0xffffffffu, sb % 65521u);           // each lane's share is < 2^32 already
        if (lane == 0) p.adler[piece] = make_uint2(sa % 65521u, sb % 65521u);
    }
    __syncthreads();
    for (int i = tid; i < 288 + 32; i += 256)
        if (s_hist[i]) atomicAdd(&p.hist[i], s_hist[i]);
}

// ---- Huffman code lengths ------------------------------------------------------------------
// Code lengths of n symbols from their counts, longest allowed length maxbits, by one thread:
// sort, two-queue Huffman merge, then the length limit: lengths above the limit are cut to it and the
// Kraft sum is restored on the per-length counts before the lengths are handed out again by rank.
// work: caller's 8-byte aligned scratch of >= 8 * n + 16 ints.  Symbols with count 0 get length 0.
// sorted != nullptr: the symbols with a non-zero count already in (count, symbol) order, m of them (the caller's
// threads ranked them in parallel); else an insertion sort runs here (small alphabets).
__device__ void png_code_lengths(const unsigned int *freq, int n, int maxbits, unsigned char *len, int *work,
                                 const int *sorted = nullptr, int m_sorted = 0)
{
    // work: sym [0, n) | parent [n, 3n) | depth of the leaves by rank [3n, 4n) | internal weights, 64 bit [4n, 6n) |
    //       depth per node [6n + 8, 8n + 8)
    int *sym = work, *parent = work + n, *depth = work + 3 * n;
    int m = 0;
    for (int i = 0; i < n; i++) len[i] = 0;
    if (sorted) { m = m_sorted; for (int i = 0; i < m; i++) sym[i] = sorted[i]; }
    else for (int i = 0; i < n; i++) if (freq[i]) sym[m++] = i;
    if (m == 0) return;
    if (m == 1) { len[sym[0]] = 1; return; }
    for (int i = 1; i < m && !sorted; i++) {                        // insertion sort by (count, symbol)
        const int s = sym[i];
        int j = i - 1;
        while (j >= 0 && (freq[sym[j]] > freq[s] || (freq[sym[j]] == freq[s] && sym[j] > s))) { sym[j + 1] = sym[j]; j--; }
        sym[j + 1] = s;
    }
    // nodes 0 .. m-1 = leaves in sorted order, m .. 2m-2 = internal nodes in creation (= weight) order
    unsigned long long *iw = reinterpret_cast<unsigned long long *>(work + 4 * n);
    int leaf = 0, inode = 0, made = 0;
    auto take = [&](unsigned long long &w) -> int {
        const bool use_leaf = leaf < m && (inode >= made || (unsigned long long)freq[sym[leaf]] <= iw[inode]);
        if (use_leaf) { w = freq[sym[leaf]]; return leaf++; }
        w = iw[inode];
        return m + inode++;
    };
    for (int k = 0; k < m - 1; k++) {
        unsigned long long wa, wb;
        const int a = take(wa), b = take(wb);
        iw[made] = wa + wb;
        parent[a] = m + made;
        parent[b] = m + made;
        made++;
    }
    // depth of every node, root (node 2m-2) first
    int *nd = work + 4 * n + 2 * n + 8;                             // depth per node, 2 n entries
    nd[2 * m - 2] = 0;
    for (int v = 2 * m - 3; v >= 0; v--) nd[v] = nd[parent[v]] + 1;
    // count the leaves per length with everything deeper than the limit cut to the limit, then restore the Kraft
    // sum (in units of 2^-maxbits): lengthen the longest codes still under the limit until the code fits, and give
    // any slack back by shortening the longest codes
    int bl_count[32];
    for (int b = 0; b < 32; b++) bl_count[b] = 0;
    unsigned long long kraft = 0;
    const unsigned long long full = 1ull << maxbits;
    for (int r = 0; r < m; r++) {
        const int d = nd[r] > maxbits ? maxbits : nd[r];
        depth[r] = d;
        bl_count[d]++;
        kraft += 1ull << (maxbits - d);
    }
    while (kraft > full) {
        int bits = maxbits - 1;
        while (bl_count[bits] == 0) bits--;
        bl_count[bits]--;
        bl_count[bits + 1]++;
        kraft -= 1ull << (maxbits - bits - 1);
    }
    for (int bits = maxbits; bits > 1 && kraft < full; bits--)
        while (bl_count[bits] > 0 && kraft + (1ull << (maxbits - bits)) <= full) {
            bl_count[bits]--;
            bl_count[bits - 1]++;
            kraft += 1ull << (maxbits - bits);
        }
    // hand the lengths out again: the rarest symbols (front of the sorted order) get the longest codes
    int r = 0;
    for (int bits = maxbits; bits >= 1; bits--)
        for (int c = bl_count[bits]; c > 0; c--) len[sym[r++]] = (unsigned char)bits;
}

__device__ __forceinline__ uint32_t png_bitrev(uint32_t v, int nbits) { return __brev(v) >> (32 - nbits); }

// canonical codes (RFC 1951 3.2.2) from lengths, stored bit-reversed for LSB-first packing: (code << 8) | length
__device__ void png_canonical(const unsigned char *len, int n, int maxbits, uint32_t *codes)
{
    int bl_count[17], next[17];
    for (int b = 0; b <= 16; b++) bl_count[b] = 0;
    for (int i = 0; i < n; i++) bl_count[len[i]]++;
    bl_count[0] = 0;
    int code = 0;
    for (int b = 1; b <= maxbits; b++) { code = (code + bl_count[b - 1]) << 1; next[b] = code; }
    for (int i = 0; i < n; i++)
        codes[i] = len[i] ? (png_bitrev((uint32_t)next[len[i]]++, len[i]) << 8) | len[i] : 0u;
}

struct PngBitWriter {
    uint32_t *w;
    unsigned int pos;
    __device__ void put(uint32_t v, int n)
    {
        if (!n) return;
        const unsigned int wi = pos >> 5, sh = pos & 31;
        w[wi] |= v << sh;
        if (sh + n > 32) w[wi + 1] |= v >> (32 - sh);
        pos += n;
    }
};

// One CTA: the threads rank the symbols by count in parallel and sum the Adler-32 partials; thread 0 then builds
// the code lengths and codes of the block and segment 0 (zlib header + block header).
constexpr int PNG_TREE_THREADS = 320;
__global__ void __launch_bounds__(PNG_TREE_THREADS) k_png_tree(const PngParams p)
{
    __shared__ __align__(8) int s_work[8 * 288 + 64];
    __shared__ unsigned char s_len[288 + 32];
    __shared__ unsigned int s_clfreq[19];
    __shared__ unsigned char s_cllen[19];
    __shared__ uint32_t s_clcode[19];
    __shared__ unsigned int s_freq[288];
    __shared__ int s_sorted[288];
    __shared__ int s_m;
    const int tid = threadIdx.x;
    if (tid < 288) s_freq[tid] = tid < PNG_NLIT ? p.hist[tid] + (tid == 256 ? 1u : 0u) : 0u;   // + the end-of-block symbol
    if (tid == 0) s_m = 0;
    __syncthreads();
    if (tid < PNG_NLIT && s_freq[tid]) {                            // rank among the used symbols by (count, symbol)
        const unsigned int f = s_freq[tid];
        int r = 0;
        for (int j = 0; j < PNG_NLIT; j++) {
            const unsigned int g = s_freq[j];
            r += (g && (g < f || (g == f && j < tid))) ? 1 : 0;
        }
        s_sorted[r] = tid;
        atomicAdd(&s_m, 1);
    }
    // Adler-32 of the whole filtered stream from the pieces' partial sums (a, b): A = 1 + sum a_c,
    // B = sum_c [ b_c + n_c * (1 + sum_{c' < c} a_c') ], everything mod 65521 -- a prefix over the pieces
    {
        __shared__ unsigned long long s_pa[PNG_TREE_THREADS], s_pb[PNG_TREE_THREADS];
        const int per = (p.npieces + PNG_TREE_THREADS - 1) / PNG_TREE_THREADS;
        const int c0 = min(p.npieces, tid * per), c1 = min(p.npieces, c0 + per);
        unsigned long long sa = 0;
        for (int c = c0; c < c1; c++) sa += p.adler[c].x;
        s_pa[tid] = sa;
        __syncthreads();
        if (tid == 0) {                                             // 320 partial sums: a serial prefix is enough
            unsigned long long run = 1;
            for (int t = 0; t < PNG_TREE_THREADS; t++) { const unsigned long long v = s_pa[t]; s_pa[t] = run % 65521ull; run += v; }
            s_pb[0] = run % 65521ull;                               // A
        }
        __syncthreads();
        const unsigned long long A_total = s_pb[0];
        __syncthreads();
        unsigned long long before = s_pa[tid], sb = 0;              // 1 + sum of a over the pieces in front
        for (int c = c0; c < c1; c++) {
            const uint2 v = p.adler[c];
            const unsigned long long n = c == p.npieces - 1 ? (unsigned long long)(p.nfilt - (long long)c * PNG_PIECE) : PNG_PIECE;
            sb = (sb + v.y + n * before) % 65521ull;
            before = (before + v.x) % 65521ull;
        }
        s_pb[tid] = sb;
        __syncthreads();
        if (tid == 0) {
            unsigned long long B = 0;
            for (int t = 0; t < PNG_TREE_THREADS; t++) B += s_pb[t];
            p.adler_out[0] = (unsigned int)(((B % 65521ull) << 16) | A_total);
        }
    }
    if (tid != 0) return;
    png_code_lengths(s_freq, PNG_NLIT, 15, s_len, s_work, s_sorted, s_m);
    unsigned char *dlen = s_len + 288;
    for (int i = 0; i < 30; i++) dlen[i] = 0;
    dlen[0] = 1;                                                    // the one distance the encoder uses (1)
    png_canonical(s_len, PNG_NLIT, 15, p.codes);
    png_canonical(dlen, 30, 15, p.codes + 288);
    int nlit = PNG_NLIT;
    while (nlit > 257 && s_len[nlit - 1] == 0) nlit--;
    const int ndist = 1;
    // the code-length alphabet: plain lengths 0..15, no repeat codes (a few hundred bits per image)
    for (int i = 0; i < 19; i++) s_clfreq[i] = 0;
    for (int i = 0; i < nlit; i++) s_clfreq[s_len[i]]++;
    for (int i = 0; i < ndist; i++) s_clfreq[dlen[i]]++;
    png_code_lengths(s_clfreq, 19, 7, s_cllen, s_work);
    png_canonical(s_cllen, 19, 7, s_clcode);
    const unsigned char order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    int nclen = 19;
    while (nclen > 4 && s_cllen[order[nclen - 1]] == 0) nclen--;
    uint32_t *seg = p.seg_bits;                                     // segment 0
    for (int i = 0; i < PNG_HDR_WORDS; i++) seg[i] = 0;
    PngBitWriter bw{seg, 0};
    bw.put(0x78, 8);                                                // CMF: deflate, 32 K window
    bw.put(0x9C, 8);                                                // FLG: default level, (0x78 << 8 | 0x9C) % 31 == 0
    bw.put(1, 1);                                                   // BFINAL
    bw.put(2, 2);                                                   // BTYPE = dynamic Huffman
    bw.put((uint32_t)(nlit - 257), 5);
    bw.put((uint32_t)(ndist - 1), 5);
    bw.put((uint32_t)(nclen - 4), 4);
    for (int i = 0; i < nclen; i++) bw.put(s_cllen[order[i]], 3);
    for (int i = 0; i < nlit; i++) bw.put(s_clcode[s_len[i]] >> 8, (int)(s_clcode[s_len[i]] & 0xff));
    for (int i = 0; i < ndist; i++) bw.put(s_clcode[dlen[i]] >> 8, (int)(s_clcode[dlen[i]] & 0xff));
    p.seg_len[0] = bw.pos;
}

// ---- tokens -> bits ------------------------------------------------------------------------
// One warp per piece: every window's tokens become codes, one warp scan of their lengths places them, and
// they are ORed (LSB first) into the piece's own bit stream in global memory (L2), which starts zeroed.
__global__ void __launch_bounds__(256) k_png_encode(const PngParams p)
{
    __shared__ uint32_t s_codes[288 + 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 288 + 32; i += 256) s_codes[i] = p.codes[i];
    __syncthreads();
    const int piece = blockIdx.x * 8 + warp;
    if (piece >= p.npieces) return;
    long long b0;
    int n;
    png_piece_span(p, piece, b0, n);
    uint32_t *seg = p.seg_bits + (size_t)(piece + 1) * PNG_PIECE_WORDS;
    for (int i = lane; i < PNG_PIECE_WORDS; i += 32) seg[i] = 0;
    __syncwarp();
    uint32_t running = 0;
    const uint32_t dist1 = s_codes[288];
    for (int off = 0; off < n; off += 32) {
        int byte;
        bool valid;
        const uint32_t eq = png_window(p, b0, n, off, lane, byte, valid);
        int mlen;
        const int kind = valid ? png_token(eq, lane, mlen) : 0;
        unsigned long long bits = 0;
        uint32_t nb = 0;
        if (kind == 1) {
            const uint32_t c = s_codes[byte];
            bits = c >> 8; nb = c & 0xffu;
        } else if (kind == 2) {
            int sym, eb, ev;
            png_len_code(mlen, sym, eb, ev);
            const uint32_t c = s_codes[sym];
            bits = c >> 8; nb = c & 0xffu;
            bits |= (unsigned long long)ev << nb; nb += eb;
            bits |= (unsigned long long)(dist1 >> 8) << nb; nb += dist1 & 0xffu;     // distance 1: no extra bits
        }
        uint32_t incl = nb;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        const uint32_t pos = running + incl - nb;
        running += __shfl_sync(0xffffffffu, incl, 31);
        if (nb) {
            const uint32_t wi = pos >> 5, sh = pos & 31;
            const uint32_t lo = (uint32_t)bits, hi = (uint32_t)(bits >> 32);
            atomicOr(&seg[wi], lo << sh);
            const uint32_t w1 = sh ? (lo >> (32 - sh)) | (hi << sh) : hi;
            if (sh + nb > 32 && w1) atomicOr(&seg[wi + 1], w1);
            if (sh + nb > 64) atomicOr(&seg[wi + 2], hi >> (32 - sh));
        }
    }
    if (piece == p.npieces - 1) {                                   // end of block
        const uint32_t c = s_codes[256];
        if (lane == 0) {
            const uint32_t wi = running >> 5, sh = running & 31;
            atomicOr(&seg[wi], (c >> 8) << sh);
            if (sh + (c & 0xffu) > 32) atomicOr(&seg[wi + 1], (c >> 8) >> (32 - sh));
        }
        running += c & 0xffu;
    }
    if (lane == 0) p.seg_len[piece + 1] = running;
}

// exclusive bit prefix over the segments (one CTA; a few thousand segments)
__global__ void __launch_bounds__(1024) k_png_scan(const PngParams p)
{
    __shared__ unsigned long long s_part[1024];
    const int tid = threadIdx.x, nseg = p.npieces + 1;
    const int per = (nseg + 1023) / 1024;
    const int i0 = min(nseg, tid * per), i1 = min(nseg, i0 + per);
    unsigned long long mine = 0;
    for (int i = i0; i < i1; i++) mine += p.seg_len[i];
    s_part[tid] = mine;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const unsigned long long t = tid >= o ? s_part[tid - o] : 0ull;
        __syncthreads();
        s_part[tid] += t;
        __syncthreads();
    }
    unsigned long long pos = s_part[tid] - mine;
    for (int i = i0; i < i1; i++) { p.seg_pos[i] = pos; pos += p.seg_len[i]; }
    if (tid == 1023) {
        const unsigned long long total_bits = s_part[1023];
        p.seg_pos[nseg] = total_bits;
        const unsigned long long zbytes = ((total_bits + 7) >> 3) + 4;          // + Adler-32
        const unsigned long long nidat = (zbytes + PNG_IDAT - 1) / PNG_IDAT;
        p.out_len[0] = 8 + 25 + zbytes + 12 * nidat + 12;
    }
}

// file offset of byte sb of the zlib stream: signature 8, IHDR 25, then per IDAT chunk 8 bytes in front, 4 behind
__device__ __forceinline__ long long png_file_off(unsigned long long sb)
{
    return 33 + 8 + (long long)sb + 12 * (long long)(sb / PNG_IDAT);
}

// One warp per segment: the bytes that BEGIN inside the segment, each completed with the first bits of the
// next segment where it straddles the boundary (a segment is never shorter than 8 bits).
__global__ void __launch_bounds__(256) k_png_merge(const PngParams p)
{
    const int lane = threadIdx.x & 31;
    const int seg = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int nseg = p.npieces + 1;
    if (seg >= nseg) return;
    const unsigned long long P = p.seg_pos[seg], L = p.seg_len[seg];
    const uint32_t *w = p.seg_bits + (size_t)seg * PNG_PIECE_WORDS;
    const uint32_t *wn = seg + 1 < nseg ? w + PNG_PIECE_WORDS : nullptr;
    const unsigned long long b_first = (P + 7) >> 3, b_end = (P + L + 7) >> 3;
    for (unsigned long long b = b_first + lane; b < b_end; b += 32) {
        const uint32_t o = (uint32_t)(8 * b - P);                   // local bit offset of the byte
        const uint32_t wi = o >> 5, sh = o & 31;
        uint32_t v = w[wi] >> sh;
        if (sh > 24) v |= w[wi + 1] << (32 - sh);                   // (the stream is zero past its end)
        v &= 0xffu;
        const uint32_t have = (uint32_t)min((unsigned long long)8, L - o);
        if (have < 8) {
            v &= (1u << have) - 1u;
            if (wn) v |= (wn[0] << have) & 0xffu;
        }
        const long long fo = png_file_off(b);
        if (fo < p.out_cap) p.out[fo] = (unsigned char)v;
    }
    if (seg == nseg - 1 && lane == 0) {                             // Adler-32, big-endian, after the last byte
        const unsigned long long zb = (p.seg_pos[nseg] + 7) >> 3;
        const unsigned int a = p.adler_out[0];
        for (int i = 0; i < 4; i++) {
            const long long fo = png_file_off(zb + i);
            if (fo < p.out_cap) p.out[fo] = (unsigned char)(a >> (24 - 8 * i));
        }
    }
}

// CRC-32 tables, filled by the host at init: [0..255] the byte table (reflected 0xEDB88320), [256 + 256 j + b] =
// the register after 256 zero bytes when it starts as b << 8 j -- the CRC register is linear over GF(2), so
// "advance by one 256-byte segment" is four look-ups, and a chunk's CRC can be folded from 32 lanes' segments.
__device__ uint32_t g_crc[256 + 4 * 256];
constexpr int PNG_CRC_SEG = 256;

__device__ __forceinline__ uint32_t png_crc_step(const uint32_t *tab, uint32_t crc, uint32_t byte)
{
    return tab[(crc ^ byte) & 0xffu] ^ (crc >> 8);
}

__device__ __forceinline__ uint32_t png_crc_adv(const uint32_t *tab, uint32_t c)       // c after PNG_CRC_SEG zero bytes
{
    return tab[256 + (c & 0xffu)] ^ tab[512 + ((c >> 8) & 0xffu)] ^ tab[768 + ((c >> 16) & 0xffu)] ^ tab[1024 + (c >> 24)];
}

// Signature, IHDR, IEND, and one warp per IDAT chunk: length, type, and the CRC-32 of type + data folded from
// the lanes' 256-byte segments.
__global__ void __launch_bounds__(256) k_png_finish(const PngParams p)
{
    __shared__ uint32_t s_tab[256 + 4 * 256];
    for (int i = threadIdx.x; i < 256 + 4 * 256; i += 256) s_tab[i] = g_crc[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const unsigned long long zbytes = ((p.seg_pos[p.npieces + 1] + 7) >> 3) + 4;
    const unsigned long long nidat = (zbytes + PNG_IDAT - 1) / PNG_IDAT;
    auto put32 = [&](long long off, uint32_t v) {
        for (int i = 0; i < 4; i++)
            if (off + i < p.out_cap) p.out[off + i] = (unsigned char)(v >> (24 - 8 * i));
    };
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
        unsigned char ihdr[17] = {'I', 'H', 'D', 'R', 0, 0, 0, 0, 0, 0, 0, 0, 8, 0, 0, 0, 0};
        for (int i = 0; i < 4; i++) { ihdr[4 + i] = (unsigned char)(p.w >> (24 - 8 * i)); ihdr[8 + i] = (unsigned char)(p.h >> (24 - 8 * i)); }
        ihdr[13] = p.pbpp == 1 ? 0 : (p.pbpp == 3 ? 2 : 6);         // colour type: gray, RGB, RGBA; 8 bits, no interlace
        for (int i = 0; i < 8; i++) if (i < p.out_cap) p.out[i] = sig[i];
        put32(8, 13);
        uint32_t crc = 0xffffffffu;
        for (int i = 0; i < 17; i++) { if (12 + i < p.out_cap) p.out[12 + i] = ihdr[i]; crc = png_crc_step(s_tab, crc, ihdr[i]); }
        put32(29, ~crc);
        const long long e = 33 + (long long)zbytes + 12 * (long long)nidat;      // IEND
        put32(e, 0);
        const unsigned char iend[4] = {'I', 'E', 'N', 'D'};
        crc = 0xffffffffu;
        for (int i = 0; i < 4; i++) { if (e + 4 + i < p.out_cap) p.out[e + 4 + i] = iend[i]; crc = png_crc_step(s_tab, crc, iend[i]); }
        put32(e + 8, ~crc);
    }
    const unsigned long long t = (unsigned long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (t >= nidat) return;
    const unsigned long long sb0 = t * PNG_IDAT;
    const uint32_t n = (uint32_t)min((unsigned long long)PNG_IDAT, zbytes - sb0);
    const long long base = 33 + (long long)t * (PNG_IDAT + 12);
    if (base + 12 + (long long)n > p.out_cap) return;
    const unsigned char *d = p.out + base + 8;
    static_assert(PNG_IDAT == 32 * PNG_CRC_SEG, "one lane per segment");
    const uint32_t seg0 = lane * PNG_CRC_SEG;
    const uint32_t len = n > seg0 ? min((uint32_t)PNG_CRC_SEG, n - seg0) : 0u;
    uint32_t c = 0;                                                 // this lane's segment from a zero register
    for (uint32_t i = 0; i < len; i++) c = png_crc_step(s_tab, c, d[seg0 + i]);
    uint32_t acc = 0xffffffffu;                                     // lane 0 folds: type bytes, then the segments in order
    const unsigned char idat[4] = {'I', 'D', 'A', 'T'};
    for (int i = 0; i < 4; i++) acc = png_crc_step(s_tab, acc, idat[i]);
    for (int l = 0; l < 32; l++) {
        const uint32_t cl = __shfl_sync(0xffffffffu, c, l), ll = __shfl_sync(0xffffffffu, len, l);
        if (ll == PNG_CRC_SEG) acc = png_crc_adv(s_tab, acc) ^ cl;
        else if (ll) { for (uint32_t i = 0; i < ll; i++) acc = png_crc_step(s_tab, acc, 0); acc ^= cl; }
    }
    if (lane == 0) {
        put32(base, n);
        for (int i = 0; i < 4; i++) p.out[base + 4 + i] = idat[i];
        put32(base + 8 + n, ~acc);
    }
}

}  // namespace icb
