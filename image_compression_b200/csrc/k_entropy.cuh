// k_entropy.cuh -- K2/K3: one entropy pass = quantise + zigzag run-lengths +
// baseline Huffman + exact byte stuffing + (optionally) the emitted scan, at ONE
// quality, reading the coefficients K1 left in HBM/L2.
//
// Replaces the back half of each javax.imageio encode of the reference's search
// (core/ImageCompressionJpg.java:167-190) and of saveCompressedImage (:255-260):
// IJG jcdctmgr.c quantisation (a2.5), jchuff.c encode_one_block / emit_bits /
// flush_bits (a2.7).  The reference compares the STUFFED stream size with the
// target (:176-180), so the pass computes exact sizes: a bit-length sum is not
// enough, every 0xFF data byte adds a stuffed 0x00 and that depends on the
// absolute bit alignment of each code.
//
// Shape: one warp owns a tile of 32 consecutive blocks (MCU order).  8 lanes per
// block, 8 zigzag coefficients per lane: ballot-free 64-bit nonzero masks built
// with shuffles give every lane its zero run (clz), code lengths are scanned
// across the warp, codes are OR-ed into a shared-memory bit stage at tile-local
// offsets.  A decoupled look-back over per-tile descriptors (bits, plus the last
// 7 bits of the tile so the successor can complete the shared byte) gives the
// tile's absolute bit offset; the stage is then read back through a funnel
// shift, 0xFF bytes are counted, and a second look-back (stuffed bytes) places
// the tile's bytes in the output.  Integer/bit work throughout: the integer pipes,
// not HBM, bound this kernel (coefficients are L2-resident between passes).
#pragma once

#include "common.cuh"

namespace icb {

struct EntParams {
    const int16_t *coef;
    FitState *st;
    unsigned long long *desc_bits;    // [ntiles] epoch(5) | status(2) | tail7(7) | bits(50)
    unsigned long long *desc_bytes;   // [ntiles] epoch(5) | status(2) | bytes(57)
    unsigned char *out0, *out1;       // ping-pong output streams (header included)
    long long out_cap;
    int nblocks, ntiles;
};

constexpr unsigned long long ST_AGG = 1ull, ST_INC = 2ull;

__device__ __forceinline__ unsigned long long ld_desc(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_desc(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// OR `len` bits (right-aligned in `bits`, len 1..32) into the big-endian bit stage at bit offset o
__device__ __forceinline__ void put_bits(uint32_t *stage, uint32_t o, uint32_t bits, int len)
{
    const uint32_t w = o >> 5, sh = o & 31;
    const unsigned long long v = (unsigned long long)bits << (64 - len - sh);
    const uint32_t hi = (uint32_t)(v >> 32), lo = (uint32_t)v;
    atomicOr(&stage[w], hi);
    if (lo) atomicOr(&stage[w + 1], lo);
}

// Warp-cooperative exclusive prefix over tiles [0, tile) by decoupled look-back.
// Lane 0 looks at tile-1, lane 1 at tile-2, ...  Returns the prefix in every lane;
// *first_desc receives tile-1's descriptor (for its tail bits).
__device__ __forceinline__ unsigned long long lookback(const unsigned long long *desc, int tile, int lane,
                                                       unsigned long long epoch, unsigned long long value_mask,
                                                       unsigned long long *first_desc)
{
    unsigned long long prefix = 0;
    bool first = true;
    int base = tile - 1;
    while (true) {
        const int idx = base - lane;
        unsigned long long d = (epoch << 59) | (ST_INC << 57);           // "before tile 0": inclusive 0
        if (idx >= 0) {
            do {
                d = ld_desc(desc + idx);
            } while ((d >> 59) != epoch);
        }
        __syncwarp();
        if (first) {
            *first_desc = __shfl_sync(0xffffffffu, d, 0);
            first = false;
        }
        const unsigned status = (unsigned)(d >> 57) & 3u;
        const unsigned inc = __ballot_sync(0xffffffffu, status == (unsigned)ST_INC);
        const int kstop = inc ? (__ffs(inc) - 1) : 32;
        unsigned long long v = (lane <= kstop) ? (d & value_mask) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        prefix += v;
        if (inc) break;
        base -= 32;
    }
    return prefix;
}

__global__ void __launch_bounds__(ENT_THREADS, 4) k_entropy(const EntParams p)
{
    __shared__ uint32_t s_ac[2][256];
    __shared__ uint32_t s_dc[2][16];
    __shared__ __align__(16) uint32_t s_recip[2][64];
    __shared__ __align__(16) unsigned short s_half[2][64];
    __shared__ unsigned s_ticket;
    extern __shared__ __align__(16) uint32_t s_stage_all[];

    FitState *st = p.st;
    if (!st->cur_valid) return;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 512; i += ENT_THREADS) s_ac[i >> 8][i & 255] = c_ac_enc[i >> 8][i & 255];
    if (tid < 32) s_dc[tid >> 4][tid & 15] = c_dc_enc[tid >> 4][tid & 15];
    if (tid < 128) {
        s_recip[tid >> 6][tid & 63] = st->recip[tid >> 6][tid & 63];
        s_half[tid >> 6][tid & 63] = st->half[tid >> 6][tid & 63];
    }
    if (tid == 0) s_ticket = atomicAdd(&st->ticket, 1u);
    uint32_t *stage = s_stage_all + warp * ENT_STAGE_WORDS;
    for (int i = lane * 4; i < ENT_STAGE_WORDS; i += 128)
        *reinterpret_cast<uint4 *>(stage + i) = make_uint4(0, 0, 0, 0);
    __syncthreads();

    const int tile = (int)s_ticket * ENT_WARPS + warp;
    if (tile >= p.ntiles) return;
    const unsigned long long epoch = (unsigned long long)st->epoch;
    const int g = lane >> 3, j = lane & 7;

    // ------------------------------------------------ symbols -> bit stage
    uint32_t running = 0;
    for (int it = 0; it < ENT_TILE_BLOCKS / 4; it++) {
        const int b = tile * ENT_TILE_BLOCKS + it * 4 + g;
        const bool valid = b < p.nblocks;
        const int cidx = b % 6;
        const int tb = cidx >= 4 ? 1 : 0;
        uint4 raw = make_uint4(0, 0, 0, 0);
        if (valid) raw = __ldg(reinterpret_cast<const uint4 *>(p.coef) + (size_t)b * 8 + j);
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
        // DC prediction: previous block of the same component in scan order
        if (j == 0 && valid) {
            int prev;
            bool has_prev = true;
            if (cidx == 0) { prev = b - 3; has_prev = b >= 6; }
            else if (cidx < 4) prev = b - 1;
            else { prev = b - 6; has_prev = b >= 6; }
            int pq = 0;
            if (has_prev) {
                const int xp = (int)__ldg(p.coef + (size_t)prev * 64);
                const uint32_t half = hw[0] & 0xffffu;
                const uint32_t q = __umulhi((uint32_t)abs(xp) + half, mm[0]);
                pq = xp < 0 ? -(int)q : (int)q;
            }
            v[0] -= pq;
        }

        // 64-bit "symbol present" mask of the block (bit 0 = DC, always present)
        uint32_t m8 = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) m8 |= (v[i] != 0 ? 1u : 0u) << i;
        if (j == 0) m8 |= 1u;
        if (!valid) m8 = 0;
        uint32_t mlo = j < 4 ? (m8 << (8 * j)) : 0u;
        uint32_t mhi = j >= 4 ? (m8 << (8 * (j - 4))) : 0u;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            mlo |= __shfl_xor_sync(0xffffffffu, mlo, o);
            mhi |= __shfl_xor_sync(0xffffffffu, mhi, o);
        }

        // per-coefficient symbols: sym[i] = code+value bits, meta[i] = len | nzrl<<8
        uint32_t sym[8], meta[8];
        uint32_t lane_bits = 0;
        const uint32_t zrl = s_ac[tb][0xF0];
        const uint32_t zrl_len = zrl & 0xffu;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            sym[i] = 0;
            meta[i] = 0;
            if ((m8 >> i) & 1u) {
                const int pos = 8 * j + i;
                const int val = v[i];
                const int nb = 32 - __clz(abs(val));
                const uint32_t vbits = (uint32_t)(val + (val >> 31)) & ((1u << nb) - 1u);
                uint32_t e, nz = 0;
                if (pos == 0) {
                    e = s_dc[tb][nb];
                } else {
                    uint32_t blo, bhi;
                    if (pos < 32) { blo = mlo & ((1u << pos) - 1u); bhi = 0; }
                    else { blo = mlo; bhi = mhi & ((1u << (pos - 32)) - 1u); }
                    const int prevpos = bhi ? (63 - __clz(bhi)) : (31 - __clz(blo));
                    const int r = pos - prevpos - 1;
                    nz = (uint32_t)r >> 4;
                    e = s_ac[tb][((r & 15) << 4) | nb];
                }
                const uint32_t len = (e & 0xffu) + nb;
                sym[i] = ((e >> 8) << nb) | vbits;
                meta[i] = len | (nz << 8);
                lane_bits += len + nz * zrl_len;
            }
        }
        // end-of-block when the last coefficient is zero
        const bool eob = valid && j == 7 && !(mhi >> 31);
        const uint32_t eobe = s_ac[tb][0];
        if (eob) lane_bits += eobe & 0xffu;

        // warp scan of bit lengths (blocks and lanes are already in stream order)
        uint32_t incl = lane_bits;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        uint32_t off = running + incl - lane_bits;
        running += __shfl_sync(0xffffffffu, incl, 31);

#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (meta[i]) {
                for (uint32_t z = meta[i] >> 8; z > 0; z--) {
                    put_bits(stage, off, zrl >> 8, (int)zrl_len);
                    off += zrl_len;
                }
                const int len = (int)(meta[i] & 0xffu);
                put_bits(stage, off, sym[i], len);
                off += len;
            }
        }
        if (eob) put_bits(stage, off, eobe >> 8, (int)(eobe & 0xffu));
    }
    __syncwarp();

    // ------------------------------------------------ absolute bit offset
    const unsigned long long L = running;
    unsigned long long tail7 = 0;
    if (L >= 7) {
        const uint32_t o = (uint32_t)L - 7, w = o >> 5, sh = o & 31;
        const unsigned long long two = ((unsigned long long)stage[w] << 32) | stage[w + 1];
        tail7 = (two >> (64 - 7 - sh)) & 0x7full;
    }
    const unsigned long long BITS_MASK = (1ull << 50) - 1ull;
    if (lane == 0)
        st_desc(p.desc_bits + tile, (epoch << 59) | ((tile == 0 ? ST_INC : ST_AGG) << 57) | (tail7 << 50) | L);
    unsigned long long P = 0, prevd = 0;
    if (tile > 0) {
        P = lookback(p.desc_bits, tile, lane, epoch, BITS_MASK, &prevd);
        if (lane == 0) st_desc(p.desc_bits + tile, (epoch << 59) | (ST_INC << 57) | (tail7 << 50) | (P + L));
    }
    const uint32_t r = (uint32_t)(P & 7ull);
    const uint32_t prev_tail = (uint32_t)((prevd >> 50) & 0x7full) & ((1u << r) - 1u);
    const bool last = tile == p.ntiles - 1;
    unsigned long long nbytes = ((P + L) >> 3) - (P >> 3);       // bytes this tile completes
    if (last) {
        const uint32_t rem = (uint32_t)((P + L) & 7ull);
        if (rem) {                                              // jchuff.c flush_bits: pad with 1s
            if (lane == 0) put_bits(stage, (uint32_t)L, (1u << (8 - rem)) - 1u, (int)(8 - rem));
            nbytes += 1;
        }
        if (lane == 0) st->total_bits = P + L;
    }
    __syncwarp();

    // ------------------------------------------------ count 0xFF bytes (stuffing)
    const uint32_t nb32 = (uint32_t)nbytes;
    const uint32_t nwords = (nb32 + 3) >> 2;
    uint32_t ff = 0;
    for (uint32_t w = lane; w < nwords; w += 32) {
        const uint32_t hi = w ? stage[w - 1] : prev_tail;
        const uint32_t val = __funnelshift_r(stage[w], hi, r);
        uint32_t eq = __vcmpeq4(val, 0xffffffffu);
        const uint32_t nvalid = min(4u, nb32 - 4 * w);
        if (nvalid < 4) eq &= 0xffffffffu << (8 * (4 - nvalid));
        ff += __popc(eq) >> 3;
    }
    uint32_t ff_tile = ff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ff_tile += __shfl_xor_sync(0xffffffffu, ff_tile, o);
    if (lane == 0 && ff_tile) atomicAdd(&st->ff_total, (unsigned long long)ff_tile);

    // ------------------------------------------------ emit (a2.7 emit_bits byte stuffing)
    unsigned char *out = st->cur_buf ? p.out1 : p.out0;
    if (out == nullptr) return;
    const unsigned long long mybytes = nbytes + ff_tile;
    const unsigned long long BYTES_MASK = (1ull << 57) - 1ull;
    if (lane == 0)
        st_desc(p.desc_bytes + tile, (epoch << 59) | ((tile == 0 ? ST_INC : ST_AGG) << 57) | mybytes);
    unsigned long long Q = 0, dummy = 0;
    if (tile > 0) {
        Q = lookback(p.desc_bytes, tile, lane, epoch, BYTES_MASK, &dummy);
        if (lane == 0) st_desc(p.desc_bytes + tile, (epoch << 59) | (ST_INC << 57) | (Q + mybytes));
    }
    const long long cap = p.out_cap;
    long long pos0 = (long long)HEADER_BYTES + (long long)Q;
    uint32_t ff_before = 0;
    for (uint32_t wb = 0; wb < nwords; wb += 32) {
        const uint32_t w = wb + lane;
        uint32_t val = 0, cnt = 0, nvalid = 0;
        if (w < nwords) {
            const uint32_t hi = w ? stage[w - 1] : prev_tail;
            val = __funnelshift_r(stage[w], hi, r);
            nvalid = min(4u, nb32 - 4 * w);
            uint32_t eq = __vcmpeq4(val, 0xffffffffu);
            if (nvalid < 4) eq &= 0xffffffffu << (8 * (4 - nvalid));
            cnt = __popc(eq) >> 3;
        }
        uint32_t incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        long long pos = pos0 + 4ll * w + ff_before + (incl - cnt);
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
    if (last && lane == 0) {                                    // EOI
        const long long e = pos0 + (long long)mybytes;
        if (e < cap) out[e] = 0xff;
        if (e + 1 < cap) out[e + 1] = 0xd9;
    }
}

constexpr int ENT_SMEM_BYTES = ENT_WARPS * ENT_STAGE_WORDS * 4;

}  // namespace icb
