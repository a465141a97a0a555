// k_entropy.cuh -- K2/K3: the entropy pass = quantise + zigzag run-lengths +
// baseline Huffman + exact byte stuffing at ONE quality, reading the coefficients
// K1 left in HBM/L2, and the final emission of the chosen stream.
//
// Replaces the back half of each javax.imageio encode of the reference's search
// (core/ImageCompressionJpg.java:167-190) and of saveCompressedImage (:255-260):
// IJG jcdctmgr.c quantisation (a2.5), jchuff.c encode_one_block / emit_bits /
// flush_bits (a2.7).  The reference compares the STUFFED stream size with the
// target (:176-180), so the pass computes exact sizes: a bit-length sum is not
// enough, every 0xFF data byte adds a stuffed 0x00 and that depends on the
// absolute bit alignment of each code.
//
// Three kernels, no inter-CTA waiting anywhere:
//   k_entropy_symbols  one warp = 32 consecutive blocks (MCU order).  Phase A, 8
//       lanes per block: quantise, DC-predict, and append every coefficient that
//       codes a symbol to a compact ring.  Phase B, all 32 lanes busy: run length
//       = distance to the previous ring entry, Huffman lookup, ZRL prefix, EOB
//       suffix, warp scan of code lengths, OR into a shared-memory bit window that
//       is flushed to a per-tile scratch stream (tile-local alignment 0).  The last
//       CTA to finish turns the per-tile bit counts into exclusive prefixes.
//   k_entropy_count    one warp per tile: re-reads the tile's bits through a funnel
//       shift at its absolute alignment (borrowing the predecessor's last bits for
//       the shared byte) and counts 0xFF bytes.  The last CTA to finish runs the
//       reference's bisection step on the exact size and sets up the next pass.
//   k_entropy_emit     once per image, for the chosen quality only: same shifted
//       read, writes the stuffed bytes at their scanned offsets, then EOI.
// Integer/bit work throughout: the integer pipes, not HBM, bound these kernels
// (coefficients stay L2-resident between passes).
#pragma once

#include "common.cuh"

namespace icb {

struct PassBufs {
    uint32_t *scratch;              // [ntiles][ENT_SCR_STRIDE] tile-local bit streams, big-endian words
    unsigned int *L;                // [ntiles] bits per tile
    unsigned long long *P;          // [ntiles] exclusive bit prefix
    unsigned int *bytes;            // [ntiles] stuffed bytes the tile completes
    unsigned long long *Q;          // [ntiles] exclusive byte prefix (k_finish, chosen pass only)
};

struct EntParams {
    const int16_t *coef;
    FitState *st;
    PassBufs buf[2];                // ping-pong: the best pass so far is never overwritten
    int nblocks, ntiles;
};

// select without dynamic indexing of the kernel parameter block (keeps it out of local memory)
__device__ __forceinline__ PassBufs pick_buf(const EntParams &p, int i)
{
    PassBufs b;
    b.scratch = i ? p.buf[1].scratch : p.buf[0].scratch;
    b.L = i ? p.buf[1].L : p.buf[0].L;
    b.P = i ? p.buf[1].P : p.buf[0].P;
    b.bytes = i ? p.buf[1].bytes : p.buf[0].bytes;
    b.Q = i ? p.buf[1].Q : p.buf[0].Q;
    return b;
}

// a2.1 / :158-200 -- defined in k_control.cuh
__device__ void device_step(FitState *st, int final_call);

// exclusive scan of src[0..n) by one CTA; returns the total in every thread
template <typename T>
__device__ __forceinline__ unsigned long long cta_exclusive_scan(const T *src, unsigned long long *dst, int n,
                                                                 unsigned long long *s_part)
{
    const int tid = threadIdx.x, nt = blockDim.x;
    const int chunk = (n + nt - 1) / nt;
    const int lo = min(n, tid * chunk), hi = min(n, lo + chunk);
    unsigned long long sum = 0;
    for (int i = lo; i < hi; i++) sum += (unsigned long long)__ldcg(src + i);
    s_part[tid] = sum;
    __syncthreads();
    if (tid == 0) {
        unsigned long long run = 0;
        for (int i = 0; i < nt; i++) {
            const unsigned long long v = s_part[i];
            s_part[i] = run;
            run += v;
        }
        s_part[nt] = run;
    }
    __syncthreads();
    unsigned long long run = s_part[tid];
    for (int i = lo; i < hi; i++) {
        dst[i] = run;
        run += (unsigned long long)__ldcg(src + i);
    }
    return s_part[nt];
}

__global__ void __launch_bounds__(ENT_THREADS) k_entropy_symbols(const EntParams p)
{
    __shared__ uint32_t s_ac[2][256];
    __shared__ uint32_t s_dc[2][16];
    __shared__ __align__(16) uint32_t s_recip[2][64];
    __shared__ __align__(16) unsigned short s_half[2][64];
    __shared__ unsigned long long s_part[ENT_THREADS + 1];
    __shared__ int s_last;
    extern __shared__ __align__(16) uint32_t s_dyn[];

    FitState *st = p.st;
    if (!st->cur_valid) return;
    const PassBufs pb = pick_buf(p, st->cur_buf);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 512; i += ENT_THREADS) s_ac[i >> 8][i & 255] = c_ac_enc[i >> 8][i & 255];
    if (tid < 32) s_dc[tid >> 4][tid & 15] = c_dc_enc[tid >> 4][tid & 15];
    if (tid < 128) {
        s_recip[tid >> 6][tid & 63] = st->recip[tid >> 6][tid & 63];
        s_half[tid >> 6][tid & 63] = st->half[tid >> 6][tid & 63];
    }
    uint32_t *stage = s_dyn + warp * (ENT_WIN + ENT_RING);
    uint32_t *ring = stage + ENT_WIN;
    for (int i = lane * 4; i < ENT_WIN; i += 128) *reinterpret_cast<uint4 *>(stage + i) = make_uint4(0, 0, 0, 0);
    __syncthreads();

    const int tile = blockIdx.x * ENT_WARPS + warp;
    if (tile < p.ntiles) {
        const int g = lane >> 3, j = lane & 7;
        uint32_t *scr = pb.scratch + (size_t)tile * ENT_SCR_STRIDE;
        uint32_t running = 0;       // bits produced so far (tile-local)
        uint32_t flushed = 0;       // whole words already copied to scratch
        uint32_t cnt = 0, head = 0; // ring entries appended / consumed
        const uint32_t zrl0 = s_ac[0][0xF0], zrl1 = s_ac[1][0xF0];
        const uint32_t eob0 = s_ac[0][0], eob1 = s_ac[1][0];

        // copy the complete words of the window to scratch and slide the window
        auto flush = [&](bool final_flush) {
            __syncwarp();
            const uint32_t have = (running >> 5) - flushed + (final_flush ? 2u : 0u);   // final: partial word + a zero
            const uint32_t carry = final_flush ? 0u : stage[have];                      // partial word stays
            __syncwarp();
            for (uint32_t w = lane; w < have; w += 32) scr[flushed + w] = stage[w];
            for (uint32_t w = lane; w <= have + 2; w += 32) stage[w] = 0;
            __syncwarp();
            if (lane == 0 && !final_flush) stage[0] = carry;
            flushed += final_flush ? 0u : have;
            __syncwarp();
        };

        auto process_round = [&](uint32_t n, bool at_end) {
            const uint32_t idx = head + lane;
            const bool active = (uint32_t)lane < n;
            const uint32_t e = ring[idx & (ENT_RING - 1)];
            const uint32_t pe = ring[(idx - 1) & (ENT_RING - 1)];
            const uint32_t ne = ring[(idx + 1) & (ENT_RING - 1)];
            const int val = (int)(short)(e & 0xffffu);
            const int pos = (int)((e >> 16) & 63u);
            const int tb = (int)((e >> 22) & 1u);
            const int nb = 32 - __clz(abs(val));
            const uint32_t vbits = (uint32_t)(val + (val >> 31)) & ((1u << nb) - 1u);
            uint32_t ent, nz = 0;
            if (pos == 0) {
                ent = s_dc[tb][nb & 15];
            } else {
                const int r = (pos - (int)((pe >> 16) & 63u) - 1) & 63;
                nz = (uint32_t)r >> 4;
                ent = s_ac[tb][((r & 15) << 4) | (nb & 15)];
            }
            unsigned long long sym = ((unsigned long long)(ent >> 8) << nb) | vbits;
            uint32_t len = (ent & 0xffu) + nb;
            if (nz && active) {                                 // runs of 16 zeros: ZRL codes in front
                const uint32_t z = tb ? zrl1 : zrl0;
                for (uint32_t k = 0; k < nz; k++) {
                    sym |= (unsigned long long)(z >> 8) << len;
                    len += z & 0xffu;
                }
            }
            const bool last_in_block =
                (at_end && idx + 1 == cnt) || (idx + 1 < cnt && ((ne >> 16) & 63u) == 0);
            if (last_in_block && pos < 63) {                    // end-of-block
                const uint32_t z = tb ? eob1 : eob0;
                sym = (sym << (z & 0xffu)) | (z >> 8);
                len += z & 0xffu;
            }
            if (!active) len = 0;
            uint32_t incl = len;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            const uint32_t off = running + incl - len;
            running += __shfl_sync(0xffffffffu, incl, 31);
            if (len) {
                const unsigned long long V = sym << (64 - len);
                const uint32_t hiw = (uint32_t)(V >> 32), low = (uint32_t)V;
                const uint32_t w = (off >> 5) - flushed, sh = off & 31;
                const uint32_t w0 = hiw >> sh;
                const uint32_t w1 = __funnelshift_r(low, hiw, sh);
                const uint32_t w2 = __funnelshift_r(0u, low, sh);
                if (w0) atomicOr(&stage[w], w0);
                if (w1) atomicOr(&stage[w + 1], w1);
                if (w2) atomicOr(&stage[w + 2], w2);
            }
            head += n;
            if ((running >> 5) - flushed > (uint32_t)(ENT_WIN - 72)) flush(false);   // a round adds <= 63 words
        };

        for (int it = 0; it < ENT_TILE_BLOCKS / 4; it++) {
            const int b = tile * ENT_TILE_BLOCKS + it * 4 + g;
            const bool valid = b < p.nblocks;
            const int cidx = b % 6;
            const int tb = cidx >= 4 ? 1 : 0;
            // DC predictor = previous block of the same component in scan order
            int prev = b - 1;
            bool has_prev = true;
            if (cidx == 0) { prev = b - 3; has_prev = b >= 6; }
            else if (cidx >= 4) { prev = b - 6; has_prev = b >= 6; }
            uint4 raw = make_uint4(0, 0, 0, 0);
            int xp = 0;
            if (valid) {
                raw = __ldg(reinterpret_cast<const uint4 *>(p.coef) + (size_t)b * 8 + j);
                if (j == 0 && has_prev) xp = (int)__ldg(p.coef + (size_t)prev * 64);
            }
            const uint4 m0 = *reinterpret_cast<const uint4 *>(&s_recip[tb][8 * j]);
            const uint4 m1 = *reinterpret_cast<const uint4 *>(&s_recip[tb][8 * j + 4]);
            const uint4 hh = *reinterpret_cast<const uint4 *>(&s_half[tb][8 * j]);
            const uint32_t rw[4] = {raw.x, raw.y, raw.z, raw.w};
            const uint32_t mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
            const uint32_t hw[4] = {hh.x, hh.y, hh.z, hh.w};

            // a2.5 quantise: sign(x) * ((|x| + 4q) / 8q), division as a multiply-high
            int v[8];
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int x = (i & 1) ? ((int)rw[i >> 1] >> 16) : ((int)(rw[i >> 1] << 16) >> 16);
                const uint32_t half = (i & 1) ? (hw[i >> 1] >> 16) : (hw[i >> 1] & 0xffffu);
                const uint32_t q = __umulhi((uint32_t)abs(x) + half, mm[i]);
                v[i] = x < 0 ? -(int)q : (int)q;
            }
            if (j == 0) {
                const uint32_t q = __umulhi((uint32_t)abs(xp) + (hw[0] & 0xffffu), mm[0]);
                v[0] -= xp < 0 ? -(int)q : (int)q;
            }
            uint32_t m8 = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) m8 |= (v[i] != 0 ? 1u : 0u) << i;
            if (j == 0) m8 |= 1u;                               // DC always codes a symbol
            if (!valid) m8 = 0;
            const uint32_t npop = __popc(m8);
            uint32_t incl = npop;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            uint32_t rank = cnt + incl - npop;
            cnt += __shfl_sync(0xffffffffu, incl, 31);
            const uint32_t tag = ((uint32_t)tb << 22) | ((uint32_t)(8 * j) << 16);
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if ((m8 >> i) & 1u) {
                    ring[rank & (ENT_RING - 1)] = ((uint32_t)v[i] & 0xffffu) | (tag + ((uint32_t)i << 16));
                    rank++;
                }
            }
            __syncwarp();
            while (cnt - head > 32) process_round(32, false);
        }
        while (head < cnt) {
            const uint32_t n = min(32u, cnt - head);
            process_round(n, head + n == cnt);
        }
        flush(true);
        if (lane == 0) pb.L[tile] = running;
    }

    // ---- last CTA: per-tile bit counts -> exclusive prefixes (replaces a look-back chain)
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(&st->done_a, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const unsigned long long total = cta_exclusive_scan(pb.L, pb.P, p.ntiles, s_part);
    if (tid == 0) st->total_bits = total;
}

// ---- a tile's bits at their absolute alignment --------------------------------
struct TileView {
    const uint32_t *scr;
    uint32_t r;             // absolute bit offset of the tile mod 8
    uint32_t prev_tail;     // the predecessor's last r bits (the byte the two tiles share)
    uint32_t nbytes;        // bytes this tile completes (the last tile also owns the padded one)
    uint32_t pad_word, pad_mask;
    bool last;
};

__device__ __forceinline__ TileView tile_view(const PassBufs &pb, int tile, int ntiles)
{
    TileView v;
    const unsigned long long L = __ldcg(pb.L + tile), P = __ldcg(pb.P + tile);
    v.scr = pb.scratch + (size_t)tile * ENT_SCR_STRIDE;
    v.r = (uint32_t)(P & 7ull);
    v.nbytes = (uint32_t)(((P + L) >> 3) - (P >> 3));
    v.last = tile == ntiles - 1;
    v.pad_word = 0xffffffffu;
    v.pad_mask = 0;
    if (v.last) {
        const uint32_t rem = (uint32_t)((P + L) & 7ull);
        if (rem) {                                              // jchuff.c flush_bits: pad with 1s
            v.nbytes += 1;
            const uint32_t bi = (v.nbytes - 1) & 3u;
            v.pad_word = (v.nbytes - 1) >> 2;
            v.pad_mask = ((1u << (8 - rem)) - 1u) << (24 - 8 * bi);
        }
    }
    v.prev_tail = 0;
    if (v.r && tile > 0) {
        const uint32_t Lp = __ldcg(pb.L + tile - 1);
        const uint32_t *sp = pb.scratch + (size_t)(tile - 1) * ENT_SCR_STRIDE;
        const uint32_t o = Lp - v.r, w = o >> 5, sh = o & 31;
        const unsigned long long two = ((unsigned long long)__ldcg(sp + w) << 32) | __ldcg(sp + w + 1);
        v.prev_tail = (uint32_t)(two >> (64 - v.r - sh)) & ((1u << v.r) - 1u);
    }
    return v;
}

// word w of the tile's byte stream (4 bytes, big-endian), and how many of them are real
__device__ __forceinline__ uint32_t tile_word(const TileView &v, uint32_t w, uint32_t &eq_ff, uint32_t &nvalid)
{
    const uint32_t hi = w ? __ldcg(v.scr + w - 1) : v.prev_tail;
    uint32_t val = __funnelshift_r(__ldcg(v.scr + w), hi, v.r);
    if (w == v.pad_word) val |= v.pad_mask;
    nvalid = min(4u, v.nbytes - 4 * w);
    eq_ff = __vcmpeq4(val, 0xffffffffu);
    if (nvalid < 4) eq_ff &= 0xffffffffu << (8 * (4 - nvalid));
    return val;
}

__global__ void __launch_bounds__(ENT_THREADS) k_entropy_count(const EntParams p)
{
    __shared__ unsigned int s_ff[ENT_WARPS];
    __shared__ int s_last;
    FitState *st = p.st;
    if (!st->cur_valid) return;
    const PassBufs pb = pick_buf(p, st->cur_buf);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x * ENT_WARPS + warp;
    uint32_t ff = 0;
    if (tile < p.ntiles) {
        const TileView v = tile_view(pb, tile, p.ntiles);
        const uint32_t nwords = (v.nbytes + 3) >> 2;
        for (uint32_t w = lane; w < nwords; w += 32) {
            uint32_t eq, nvalid;
            tile_word(v, w, eq, nvalid);
            ff += __popc(eq) >> 3;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ff += __shfl_xor_sync(0xffffffffu, ff, o);
        if (lane == 0) pb.bytes[tile] = v.nbytes + ff;
    }
    if (lane == 0) s_ff[warp] = ff;
    __syncthreads();
    if (tid == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < ENT_WARPS; w++) t += s_ff[w];
        if (t) atomicAdd(&st->ff_total, t);
        __threadfence();
        s_last = atomicAdd(&st->done_c, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    device_step(st, 0);         // :176-189 on the exact size, then tables of the next pass
}

// once per image: byte prefixes of the chosen pass + the 623-byte header (a2.8)
__global__ void __launch_bounds__(ENT_THREADS) k_finish(const EntParams p, unsigned char *out, long long out_cap,
                                                        int w, int h);

// once per image: a2.7 emit_bits byte stuffing of the chosen pass, then EOI
__global__ void __launch_bounds__(ENT_THREADS) k_entropy_emit(const EntParams p, unsigned char *out, long long cap)
{
    const FitState *st = p.st;
    if (!st->emit_valid) return;
    const PassBufs pb = pick_buf(p, st->best_buf);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile = blockIdx.x * ENT_WARPS + warp;
    if (tile >= p.ntiles) return;
    const TileView v = tile_view(pb, tile, p.ntiles);
    const uint32_t nwords = (v.nbytes + 3) >> 2;
    const long long pos0 = (long long)HEADER_BYTES + (long long)__ldcg(pb.Q + tile);
    uint32_t ff_before = 0;
    for (uint32_t wb = 0; wb < nwords; wb += 32) {
        const uint32_t w = wb + lane;
        uint32_t val = 0, cntf = 0, nvalid = 0;
        if (w < nwords) {
            uint32_t eq;
            val = tile_word(v, w, eq, nvalid);
            cntf = __popc(eq) >> 3;
        }
        uint32_t incl = cntf;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        long long pos = pos0 + 4ll * w + ff_before + (incl - cntf);
        for (uint32_t bi = 0; bi < nvalid; bi++) {
            const uint32_t byte = (val >> (24 - 8 * bi)) & 0xffu;
            if (pos < cap) out[pos] = (unsigned char)byte;
            pos++;
            if (byte == 0xffu) {
                if (pos < cap) out[pos] = 0;
                pos++;
            }
        }
        ff_before += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (v.last && lane == 0) {                                  // EOI
        const long long e = pos0 + (long long)__ldcg(pb.bytes + tile);
        if (e < cap) out[e] = 0xff;
        if (e + 1 < cap) out[e + 1] = 0xd9;
    }
}

constexpr int ENT_SMEM_BYTES = ENT_WARPS * (ENT_WIN + ENT_RING) * 4;

// per-slot pass arena: 2 x (scratch + L + P + bytes + Q)
__host__ __device__ inline size_t ent_pass_bytes(int ntiles)
{
    const size_t n = (size_t)ntiles;
    return n * ENT_SCR_STRIDE * 4 + n * 4 + n * 8 + n * 4 + n * 8 + 64;
}

}  // namespace icb
