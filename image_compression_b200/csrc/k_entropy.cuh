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

// Warp-cooperative exclusive prefix over descriptors [0, idx0) by decoupled
// look-back.  Lane 0 looks at idx0-1, lane 1 at idx0-2, ...; four 32-wide windows
// are requested at once so a long walk costs a quarter of the dependent L2 round
// trips.  Returns the prefix in every lane; *first_desc = descriptor idx0-1.
__device__ __forceinline__ unsigned long long lookback(const unsigned long long *desc, int idx0, int lane,
                                                       unsigned long long epoch, unsigned long long value_mask,
                                                       unsigned long long *first_desc)
{
    constexpr int U = 4;
    unsigned long long prefix = 0;
    bool first = true;
    int base = idx0 - 1;
    const unsigned long long before0 = (epoch << 59) | (ST_INC << 57);   // "before 0": inclusive 0
    while (true) {
        unsigned long long d[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int idx = base - 32 * u - lane;
            d[u] = idx >= 0 ? ld_desc(desc + idx) : before0;
        }
        bool done = false;
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (done) break;
            const int idx = base - 32 * u - lane;
            while ((d[u] >> 59) != epoch) d[u] = ld_desc(desc + idx);
            __syncwarp();
            if (first) {
                *first_desc = __shfl_sync(0xffffffffu, d[u], 0);
                first = false;
            }
            const unsigned status = (unsigned)(d[u] >> 57) & 3u;
            const unsigned inc = __ballot_sync(0xffffffffu, status == (unsigned)ST_INC);
            const int kstop = inc ? (__ffs(inc) - 1) : 32;
            unsigned long long v = (lane <= kstop) ? (d[u] & value_mask) : 0ull;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            prefix += v;
            if (inc) done = true;
        }
        if (done) break;
        base -= 32 * U;
    }
    return prefix;
}

__global__ void __launch_bounds__(ENT_THREADS) k_entropy(const EntParams p)
{
    __shared__ uint32_t s_ac[2][256];
    __shared__ uint32_t s_dc[2][16];
    __shared__ __align__(16) uint32_t s_recip[2][64];
    __shared__ __align__(16) unsigned short s_half[2][64];
    __shared__ unsigned s_ticket;
    __shared__ unsigned long long s_L[ENT_WARPS], s_bytes[ENT_WARPS], s_P, s_Q;
    __shared__ uint32_t s_tail[ENT_WARPS], s_prevtail;
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
    uint32_t *stage = s_stage_all + warp * (ENT_STAGE_WORDS + ENT_RING);
    uint32_t *ring = stage + ENT_STAGE_WORDS;
    for (int i = lane * 4; i < ENT_STAGE_WORDS; i += 128)
        *reinterpret_cast<uint4 *>(stage + i) = make_uint4(0, 0, 0, 0);
    __syncthreads();

    const int tile = (int)s_ticket * ENT_WARPS + warp;
    const unsigned long long epoch = (unsigned long long)st->epoch;
    const int g = lane >> 3, j = lane & 7;

    // ------------------------------------------------ symbols -> bit stage
    // Phase A (per 4 blocks): quantise, DC-predict, and append every coefficient that
    // produces a symbol (DC always, nonzero AC) to a compact ring of entries
    // val16 | pos<<16 | table<<22, in stream order.  Phase B (per 32 entries, all lanes
    // busy): run length = distance to the previous entry, Huffman lookup, ZRL prefix,
    // EOB suffix when the next entry starts a new block, warp scan of lengths, OR into
    // the bit stage.
    uint32_t running = 0;       // bits staged so far (tile-local)
    uint32_t cnt = 0, head = 0; // ring entries appended / consumed
    const uint32_t zrl0 = s_ac[0][0xF0], zrl1 = s_ac[1][0xF0];
    const uint32_t eob0 = s_ac[0][0], eob1 = s_ac[1][0];

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
        const bool last_in_block = (at_end && idx + 1 == cnt) || (idx + 1 < cnt && ((ne >> 16) & 63u) == 0);
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
            const uint32_t w = off >> 5, sh = off & 31;
            const uint32_t w0 = hiw >> sh;
            const uint32_t w1 = __funnelshift_r(low, hiw, sh);
            const uint32_t w2 = __funnelshift_r(0u, low, sh);
            if (w0) atomicOr(&stage[w], w0);
            if (w1) atomicOr(&stage[w + 1], w1);
            if (w2) atomicOr(&stage[w + 2], w2);
        }
        head += n;
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
    __syncwarp();

    // ------------------------------------------------ absolute bit offset (CTA-level chain)
    // The look-back chain runs over CTAs, not warps: a chain of N descriptors costs up
    // to N/32 dependent L2 round trips for the first wave, so it has to be short.
    const int cta = (int)s_ticket;
    const unsigned long long L = running;
    unsigned long long tail7 = 0;
    if (L >= 7) {
        const uint32_t o = (uint32_t)L - 7, w = o >> 5, sh = o & 31;
        const unsigned long long two = ((unsigned long long)stage[w] << 32) | stage[w + 1];
        tail7 = (two >> (64 - 7 - sh)) & 0x7full;
    }
    if (lane == 0) { s_L[warp] = L; s_tail[warp] = (uint32_t)tail7; }
    __syncthreads();
    const unsigned long long BITS_MASK = (1ull << 50) - 1ull;
    if (warp == 0) {
        unsigned long long Lc = 0, tc = 0;
#pragma unroll
        for (int w = 0; w < ENT_WARPS; w++) {
            if (s_L[w]) tc = s_tail[w];                         // tail of the last warp that has bits
            Lc += s_L[w];
        }
        if (lane == 0)
            st_desc(p.desc_bits + cta, (epoch << 59) | ((cta == 0 ? ST_INC : ST_AGG) << 57) | (tc << 50) | Lc);
        unsigned long long Pc = 0, prevd = 0;
        if (cta > 0) {
            Pc = lookback(p.desc_bits, cta, lane, epoch, BITS_MASK, &prevd);
            if (lane == 0) st_desc(p.desc_bits + cta, (epoch << 59) | (ST_INC << 57) | (tc << 50) | (Pc + Lc));
        }
        if (lane == 0) { s_P = Pc; s_prevtail = (uint32_t)((prevd >> 50) & 0x7full); }
    }
    __syncthreads();
    unsigned long long P = s_P;
    uint32_t ptail = s_prevtail;
    for (int w = 0; w < warp; w++) {
        if (s_L[w]) ptail = s_tail[w];
        P += s_L[w];
    }
    const uint32_t r = (uint32_t)(P & 7ull);
    const uint32_t prev_tail = ptail & ((1u << r) - 1u);
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
    const unsigned long long mybytes = nbytes + ff_tile;
    if (lane == 0) s_bytes[warp] = mybytes;

    // ------------------------------------------------ emit (a2.7 emit_bits byte stuffing)
    unsigned char *out = st->cur_buf ? p.out1 : p.out0;
    __syncthreads();
    if (warp == 0) {
        unsigned long long Bc = 0, Fc = 0;
#pragma unroll
        for (int w = 0; w < ENT_WARPS; w++) Bc += s_bytes[w];
        // stuffed zeros of the whole CTA = bytes - completed data bytes; one atomic per CTA
        if (lane == 0) {
            unsigned long long Lc = 0;
            for (int w = 0; w < ENT_WARPS; w++) Lc += s_L[w];
            const unsigned long long P0 = s_P;
            unsigned long long data = ((P0 + Lc) >> 3) - (P0 >> 3);
            if (cta == (p.ntiles - 1) / ENT_WARPS && ((P0 + Lc) & 7ull)) data += 1;
            Fc = Bc - data;
            if (Fc) atomicAdd(&st->ff_total, Fc);
        }
        unsigned long long Qc = 0, dummy = 0;
        if (out != nullptr) {
            const unsigned long long BYTES_MASK = (1ull << 57) - 1ull;
            if (lane == 0)
                st_desc(p.desc_bytes + cta, (epoch << 59) | ((cta == 0 ? ST_INC : ST_AGG) << 57) | Bc);
            if (cta > 0) {
                Qc = lookback(p.desc_bytes, cta, lane, epoch, BYTES_MASK, &dummy);
                if (lane == 0) st_desc(p.desc_bytes + cta, (epoch << 59) | (ST_INC << 57) | (Qc + Bc));
            }
        }
        if (lane == 0) s_Q = Qc;
    }
    if (out == nullptr) return;
    __syncthreads();
    unsigned long long Q = s_Q;
    for (int w = 0; w < warp; w++) Q += s_bytes[w];
    const long long cap = p.out_cap;
    long long pos0 = (long long)HEADER_BYTES + (long long)Q;
    uint32_t ff_before = 0;
    for (uint32_t wb = 0; wb < nwords; wb += 32) {
        const uint32_t w = wb + lane;
        uint32_t val = 0, cntf = 0, nvalid = 0;
        if (w < nwords) {
            const uint32_t hi = w ? stage[w - 1] : prev_tail;
            val = __funnelshift_r(stage[w], hi, r);
            nvalid = min(4u, nb32 - 4 * w);
            uint32_t eq = __vcmpeq4(val, 0xffffffffu);
            if (nvalid < 4) eq &= 0xffffffffu << (8 * (4 - nvalid));
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
    if (last && lane == 0) {                                    // EOI
        const long long e = pos0 + (long long)mybytes;
        if (e < cap) out[e] = 0xff;
        if (e + 1 < cap) out[e + 1] = 0xd9;
    }
}

constexpr int ENT_SMEM_BYTES = ENT_WARPS * (ENT_STAGE_WORDS + ENT_RING) * 4;

}  // namespace icb
