// k_entropy.cuh -- K2/K3: survivor compaction, the entropy pass (quantise +
// run lengths + baseline Huffman + exact byte-stuffing count at ONE quality) and
// the emission of the chosen stream.
//
// Replaces the back half of each javax.imageio encode of the reference's search
// (core/ImageCompressionJpg.java:167-190) and of saveCompressedImage (:255-260):
// IJG jcdctmgr.c quantisation (a2.5), jchuff.c encode_one_block / emit_bits /
// flush_bits (a2.7).  The reference compares the STUFFED stream size with the
// target (:176-180), so every pass computes the exact size: a bit-length sum is
// not enough, every 0xFF data byte adds a stuffed 0x00 and that depends on the
// absolute bit alignment of each code.
//
// Three kernels, no inter-CTA waiting anywhere:
//   k_entropy_prepare  once per image (per scale): walks the dense coefficients K1
//       left in HBM/L2 and keeps only the "survivors" -- coefficients that are
//       non-zero at the HIGHEST quality any pass of this job can use
//       (|x| >= 4*q_table, tables are monotone in quality) -- as compact 32-bit
//       entries in scan order, one list per tile of 64 blocks.  Every block
//       contributes a DC entry that also carries its predictor's raw DC.  One warp
//       per tile, one lane per block (half a tile staged in shared memory at a time).
//   k_entropy_pass     once per search step: one warp per tile, two survivors per
//       lane and round: quantise at the current tables, drop what became zero, run
//       length = distance to the previous live entry, Huffman lookup, ZRL prefixes,
//       the end-of-block code as a prefix of the next block's DC symbol, ONE warp
//       scan of code lengths per 64 entries, OR into a shared-memory bit window
//       flushed to a per-tile scratch stream (tile-local alignment 0).  While
//       flushing, 0xFF bytes are counted for all 8 possible start alignments.  The
//       last warp of a chunk (the 8 tiles a CTA takes per round) to finish folds the
//       chunk from shared memory; a CTA draws one ticket when it is done and the last
//       CTA scans the chunk records into the exact stuffed size and runs the
//       reference's bisection step (k_control.cuh).  CTAs are persistent and nobody
//       waits for another CTA.
//   k_entropy_emit     once per image, for the chosen pass only: re-reads the
//       tile streams through a funnel shift at their absolute alignment, writes
//       the stuffed bytes at their final offsets, the 623-byte header and EOI.
// Integer/bit work throughout: the integer pipes, not HBM, bound these kernels
// (coefficients and survivor lists stay L2-resident between passes).
#pragma once

#include "common.cuh"

namespace icb {

struct PassBufs {
    uint32_t *scratch;              // [ntiles][scr_stride] tile-local bit streams, big-endian words
    unsigned int *L;                // [ntiles] bits per tile
    uint4 *ff;                      // [ntiles] 8 x u16: 0xFF bytes inside the tile per absolute start offset mod 8
    unsigned int *ones;             // [ntiles] leading 1-bits | trailing 1-bits << 8 (each capped at 8)
    uint4 *crec;                    // [nchunks][3] chunk record: the same 8 counters (u32) folded over the chunk's
                                    //              tiles (2 x uint4) | bits, ones, 0, 0    (chunk = the 8 tiles of a CTA round)
    unsigned long long *Pc;         // [nchunks] exclusive bit prefix            (last CTA of the pass)
    unsigned int *Fc;               // [nchunks] exclusive prefix of 0xFF bytes, relative to Fb  (last CTA of the pass)
    unsigned int *Fb;               // [ENT_THREADS] base of Fc for the chunks thread t of the last CTA scanned
};

struct EntParams {
    const int16_t *coef;
    FitState *st;
    uint32_t *ent;                  // [nunits][UNIT_CAP] survivor entries, one list per unit (common.cuh)
    unsigned int *ent_cnt;          // [nunits]
    unsigned short *unit_info;      // [nunits][16] links between the units' DC predictors (k_transform.cuh XfLists)
    PassBufs buf[2];                // ping-pong: the best pass so far is never overwritten
    int nblocks, nunits, ntiles, nchunks;
    int unit_group, scr_stride;     // units per entropy tile (Geom.unit_group), scratch words reserved per tile
    int units_per_row, mcu_cols;    // unit geometry (unit_span)
};

// select without dynamic indexing of the kernel parameter block (keeps it out of local memory)
__device__ __forceinline__ PassBufs pick_buf(const EntParams &p, int i)
{
    PassBufs b;
    b.scratch = i ? p.buf[1].scratch : p.buf[0].scratch;
    b.L = i ? p.buf[1].L : p.buf[0].L;
    b.ff = i ? p.buf[1].ff : p.buf[0].ff;
    b.ones = i ? p.buf[1].ones : p.buf[0].ones;
    b.crec = i ? p.buf[1].crec : p.buf[0].crec;
    b.Pc = i ? p.buf[1].Pc : p.buf[0].Pc;
    b.Fc = i ? p.buf[1].Fc : p.buf[0].Fc;
    b.Fb = i ? p.buf[1].Fb : p.buf[0].Fb;
    return b;
}

// what a fit job starts from (core/ImageCompressionJpg.java:158-162 or one explicit quality)
struct BeginArgs {
    int mode;                       // FitMode
    float q;                        // initial quality (search) or the quality to encode at (single)
    long long target;
    int table_mode, jfif_minor;
    float list_q;                   // quality the first survivor lists are built for (>= that of the first pass)
    int pad;
};

// a2.1 / :158-200 -- defined in k_control.cuh
__device__ void device_step(FitState *st, const FitState *snap, int first, unsigned long long tbits,
                            unsigned long long ffs);
__device__ __forceinline__ void precompute_next_tables(FitState *st, int tid);
__device__ __forceinline__ int scaled_qval(int base, float q, int table_mode);
__device__ void begin_job(FitState *st, const BeginArgs &b);

// ---- survivor entries ------------------------------------------------------------
// AC: val(16) | table << 17 | zigzag pos(6) << 18
// DC: dc(16)  | 1 << 16 | table << 17 | predictor dc(14) << 18      (raw DC = sum - 8192: 14 bits signed)
// so that (e >> 17) & 127 = 2 * pos + table is the key of an AC entry's quantiser
constexpr uint32_t ENTRY_DC = 0x10000u;

__device__ __forceinline__ uint32_t ff_select(const uint4 &f, uint32_t r)        // counter r of 8 x u16
{
    const uint32_t lo = (r & 4u) ? f.z : f.x, hi = (r & 4u) ? f.w : f.y;
    const uint32_t w = (r & 2u) ? hi : lo;
    return (r & 1u) ? (w >> 16) : (w & 0xffffu);
}

// number of 0xFF bytes in a word (bytes whose 8 bits are all ones)
__device__ __forceinline__ uint32_t ff_bytes(uint32_t v)
{
    return __popc(((v & 0x7f7f7f7fu) + 0x01010101u) & v & 0x80808080u);
}

// byte shared by two neighbouring tiles: r = start offset mod 8 of the second (r != 0)
__device__ __forceinline__ uint32_t span_ff(uint32_t ones_prev, uint32_t ones_cur, uint32_t r)
{
    return (r != 0u && (ones_prev >> 8) >= r && (ones_cur & 0xffu) >= 8u - r) ? 1u : 0u;
}

// Survivor extraction from the dense coefficients in HBM/L2 -- the stand-alone form (the transform kernel
// writes the same lists itself while a tile's coefficients are on chip, k_transform.cuh; this kernel
// serves callers that reuse coefficients: icb_jpg_sizes, the search after a missed cached try, and the
// rebuild below).  One warp per unit, taken 32 blocks at a time: a slice (4 KB of
// zigzag int16) is staged in shared memory with coalesced 16-byte loads, rows padded to 144 bytes
// so that the 16-byte reads of 32 lanes standing on 32 different rows touch every bank once.  Each
// lane then owns one whole block: 64 threshold compares into a 64-bit mask, one warp scan for the
// run starts, and a pop-lowest-bit loop that writes the block's entries.  The grid is one wave of
// resident warps; a warp strides over tiles when the image has more tiles than that.
constexpr int PREP_WARPS = 4;
constexpr int PREP_THREADS = PREP_WARPS * 32;
constexpr int PREP_CTAS_PER_SM = 8;
constexpr int PREP_ROW = 144;

//
// REBUILD = false: the first kernel of a job (also starts the search state).  REBUILD = true: run once
// after the first pass of a search whose lists were built for that pass's quality only (the host
// expects the first midpoint not to fit, and then no later pass looks above it); if
// the pass did fit, the coming midpoints lie above the lists' quality and the lists are rebuilt
// from the dense coefficients for hi, otherwise this is a no-op.
template <bool REBUILD>
__global__ void __launch_bounds__(PREP_THREADS, PREP_CTAS_PER_SM) k_entropy_prepare(const EntParams p, const BeginArgs ba)
{
    // thresholds as [chunk of 8][table][8]: both tables' 16-byte chunks sit in different banks
    __shared__ __align__(16) unsigned short s_thr[8][2][8];
    __shared__ __align__(16) unsigned char s_half[PREP_WARPS][32 * PREP_ROW];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_launch_dependents();
    float tq = ba.list_q;
    if (REBUILD) {
        __shared__ float s_q;
        pdl_wait();                 // the step of the first pass is visible
        if (tid == 0) {
            FitState *st = p.st;
            const float lq = st->list_q[0], hi = st->hi;
            const bool go = !st->done && st->cur_valid && st->cur_q > lq;
            s_q = go ? hi : -1.0f;
            if (blockIdx.x == 0) st->list_q[1] = go ? hi : lq;
        }
        __syncthreads();
        tq = s_q;
        if (tq < 0.0f) return;
    }
    // survivor thresholds: a coefficient is non-zero at quality q iff |x| >= 4 * table(q), and
    // table(q) never grows with q, so the tables of the highest quality the coming passes can reach
    // (every midpoint lies below hi) bound them all.  Zigzag order, computed by every CTA.
    {
        const int tb = tid >> 6, k = tid & 63;
        const int nat = c_zigzag[k];
        s_thr[k >> 3][tb][k & 7] =
            (unsigned short)(scaled_qval(tb ? c_base_chr[nat] : c_base_lum[nat], tq, ba.table_mode) << 2);
    }
    // the job's search state and first tables (consumed by the pass kernels that follow)
    // (nothing of the previous kernel of the stream, the transform, is touched by this)
#ifdef ICB_TIMING
    if (blockIdx.x == 0 && tid == 0) p.st->dbg_k[0] = gtime();
#endif
    if (!REBUILD && blockIdx.x == 0 && tid == 0) { p.st->kt[1][0] = gtime(); p.st->kt[1][1] = 0; }
    if (!REBUILD && blockIdx.x == 0 && warp == PREP_WARPS - 1) begin_job(p.st, ba);
    __syncthreads();
    if (!REBUILD) {
        pdl_wait();                 // the coefficients are complete and visible from here on
        if (blockIdx.x == 0 && tid == 0) p.st->kt[0][1] = gtime();      // = the moment the transform grid finished
    }
    unsigned char *ts = s_half[warp];
    const unsigned char *rp = ts + lane * PREP_ROW;
    for (int unit = blockIdx.x * PREP_WARPS + warp; unit < p.nunits; unit += gridDim.x * PREP_WARPS) {
        uint32_t *out = p.ent + (size_t)unit * UNIT_CAP;
        uint32_t base = 0;
        int ub0, unb;
        unit_span(unit, p.units_per_row, p.mcu_cols, ub0, unb);
#pragma unroll 1
        for (int half = 0; half * 32 < unb; half++) {
            const int b0 = ub0 + half * 32;
            const int nb = unb - half * 32;             // blocks of the unit from this slice on
            {
                const uint4 *src = reinterpret_cast<const uint4 *>(p.coef) + (size_t)b0 * 8;
                uint4 v[8];
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    const int idx = r * 32 + lane;
                    v[r] = (idx >> 3) < nb ? __ldcg(src + idx) : make_uint4(0, 0, 0, 0);
                }
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    const int idx = r * 32 + lane;
                    *reinterpret_cast<uint4 *>(ts + (idx >> 3) * PREP_ROW + (idx & 7) * 16) = v[r];
                }
            }
            __syncwarp();
            const int b = b0 + lane;
            const int cidx = b % 6;
            const uint32_t tb = cidx >= 4 ? 1u : 0u;
            uint32_t lo = 0, hi = 0;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const uint4 raw = *reinterpret_cast<const uint4 *>(rp + 16 * j);
                const uint4 tt = *reinterpret_cast<const uint4 *>(&s_thr[j][tb][0]);
                const uint32_t rw[4] = {raw.x, raw.y, raw.z, raw.w};
                const uint32_t tw[4] = {tt.x, tt.y, tt.z, tt.w};
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const int x = (i & 1) ? ((int)rw[i >> 1] >> 16) : ((int)(rw[i >> 1] << 16) >> 16);
                    const uint32_t thr = (i & 1) ? (tw[i >> 1] >> 16) : (tw[i >> 1] & 0xffffu);
                    const int bit = 8 * j + i;
                    if ((uint32_t)abs(x) >= thr) {
                        if (bit < 32) lo |= 1u << bit; else hi |= 1u << (bit - 32);
                    }
                }
            }
            lo |= 1u;                                       // DC always codes a symbol
            if (lane >= nb) { lo = 0; hi = 0; }
            const uint32_t cnt = __popc(lo) + __popc(hi);
            uint32_t incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            uint32_t rank = base + incl - cnt;
            base += __shfl_sync(0xffffffffu, incl, 31);
            if (lane < nb) {
                const uint32_t tag = tb << 17;
                // DC predictor = previous block of the same component in scan order
                int prev = b - 1;
                bool has_prev = true;
                if (cidx == 0) { prev = b - 3; has_prev = b >= 6; }
                else if (cidx >= 4) { prev = b - 6; has_prev = b >= 6; }
                int xp = 0;
                if (has_prev)
                    xp = prev >= b0 ? (int)*reinterpret_cast<const short *>(ts + (prev - b0) * PREP_ROW)
                                    : (int)__ldcg(p.coef + (size_t)prev * 64);
                out[rank++] = ENTRY_DC | tag | ((uint32_t)xp << 18) |
                              (uint32_t)*reinterpret_cast<const unsigned short *>(rp);
                uint32_t m = lo & ~1u;
                while (m) {
                    const uint32_t k = __ffs(m) - 1;
                    m &= m - 1;
                    out[rank++] = (uint32_t)*reinterpret_cast<const unsigned short *>(rp + 2 * k) | tag | (k << 18);
                }
                m = hi;
                while (m) {
                    const uint32_t k = __ffs(m) + 31;
                    m &= m - 1;
                    out[rank++] = (uint32_t)*reinterpret_cast<const unsigned short *>(rp + 2 * k) | tag | (k << 18);
                }
            }
            __syncwarp();                                   // the rows are free for the next half
        }
        // the list is padded with dead (zero) entries to a whole number of 64-entry rounds: the pass
        // reads rounds without masking their tails (UNIT_CAP is a multiple of 64)
        for (uint32_t i = base + lane; i < ((base + 63u) & ~63u); i += 32) out[i] = 0u;
        if (lane == 0) p.ent_cnt[unit] = base;
    }
}


// run-of-8 detector: bit (31 - p) of the result is set iff bits p .. p+7 of the 64-bit
// big-endian pair (a, b) are all ones, for starts p = 0..31 inside a
__device__ __forceinline__ uint32_t ff_starts(uint32_t a, uint32_t b)
{
    const uint32_t a1 = a & __funnelshift_l(b, a, 1), b1 = b & (b << 1);
    const uint32_t a2 = a1 & __funnelshift_l(b1, a1, 2), b2 = b1 & (b1 << 2);
    return a2 & __funnelshift_l(b2, a2, 4);
}

__global__ void __launch_bounds__(ENT_THREADS, 5) k_entropy_pass(const EntParams p)
{
    __shared__ __align__(16) uint32_t s_huf[HUF_WORDS];         // [plane = 2 * table + isdc][run or owed EOB][size] (common.cuh g_huf)
    __shared__ __align__(16) uint2 s_qt[128];       // quantiser by key = 2 * pos + table
    __shared__ int4 s_ctl[6];                       // control block of FitState as the pass found it
    __shared__ __align__(16) uint32_t s_win[ENT_WARPS * ENT_WSTRIDE];
    __shared__ __align__(16) uint32_t s_rec[ENT_RING][ENT_WARPS][8];    // tile records by CTA round: ff (4 words), L, ones
    __shared__ unsigned int s_arr[ENT_RING];        // warps of the round's chunk that have finished their tile
    __shared__ unsigned long long s_scan[ENT_WARPS];
    __shared__ unsigned long long s_total[2];
    __shared__ int s_last;

    FitState *st = p.st;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#ifdef ICB_TIMING
    int dbg_i = 0;
    const bool dbg_me = blockIdx.x == gridDim.x / 2 && warp == 3 && lane == 0;
    unsigned long long dbg_early[3] = {0, 0, 0};            // a pass with nothing to do must not overwrite the last real one
#define ICB_STAMP() do { if (dbg_me && dbg_i < 24) { if (dbg_i < 3) dbg_early[dbg_i++] = clock64(); else st->dbg_w[dbg_i++] = clock64(); } } while (0)
#else
#define ICB_STAMP() do { } while (0)
#endif
    ICB_STAMP();                    // 0: entry
    pdl_launch_dependents();
    // set-up that does not depend on the previous pass: runs while that pass drains
    static_assert(HUF_WORDS % 4 == 0 && HUF_WORDS / 4 <= ENT_THREADS && ENT_WSTRIDE % 4 == 0, "16-byte set-up copies");
    if (tid < HUF_WORDS / 4) reinterpret_cast<uint4 *>(s_huf)[tid] = __ldg(reinterpret_cast<const uint4 *>(g_huf) + tid);
    const uint32_t wbase = warp * ENT_WSTRIDE;
    const uint32_t win_base = (uint32_t)__cvta_generic_to_shared(s_win + wbase);     // this warp's bit window
    for (int i = lane; i < ENT_WSTRIDE / 4; i += 32) reinterpret_cast<uint4 *>(s_win + wbase)[i] = make_uint4(0, 0, 0, 0);
    if (tid < ENT_RING) s_arr[tid] = 0;
    ICB_STAMP();                    // 1: tables requested, window zeroed
    pdl_wait();                     // the previous pass ran its bisection step: st is final
    ICB_STAMP();                    // 2: after griddepcontrol.wait
    // one memory round trip, few requests: warp 0 alone fetches the control block (lanes 0-5) and
    // all three table slots, picks the slot and leaves both in shared memory (hundreds of CTAs
    // start together: every thread asking for the same lines would queue up at one L2 slice)
    if (warp == 0) {
        int4 cw = make_int4(0, 0, 0, 0);
        if (lane < 6) cw = __ldcg(reinterpret_cast<const int4 *>(st) + lane);
        uint4 ta[3], tb[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            ta[k] = __ldcg(reinterpret_cast<const uint4 *>(st->tab[k].rh) + lane);
            tb[k] = __ldcg(reinterpret_cast<const uint4 *>(st->tab[k].rh) + 32 + lane);
        }
        if (lane < 6) s_ctl[lane] = cw;
        const int tsel = __shfl_sync(0xffffffffu, cw.z, 2);                 // int4 #2 = cur_buf, best_buf, tsel, n_trace
        reinterpret_cast<uint4 *>(s_qt)[lane] = tsel == 0 ? ta[0] : (tsel == 1 ? ta[1] : ta[2]);
        reinterpret_cast<uint4 *>(s_qt)[32 + lane] = tsel == 0 ? tb[0] : (tsel == 1 ? tb[1] : tb[2]);
    }
    ICB_STAMP();                    // 3: control + tables arrived
    __syncthreads();
    ICB_STAMP();                    // 4: after the prologue barrier
    const int4 c1 = s_ctl[1];       // iter, done, cur_valid, mode
    const int4 c2 = s_ctl[2];       // cur_buf, best_buf, tsel, n_trace
    if (!c1.z) return;
#ifdef ICB_TIMING
    if (dbg_me) { st->dbg_w[0] = dbg_early[0]; st->dbg_w[1] = dbg_early[1]; st->dbg_w[2] = dbg_early[2]; }
#endif
    const PassBufs pb = pick_buf(p, c2.x);
    const int kt_slot = 2 + ((c1.x - 1) & 7);       // kernel clock of this pass (FitState.kt)
#ifndef ICB_NO_KCLK
    if (blockIdx.x == 0 && tid == 0) st->kt[kt_slot][0] = gtime();
#endif
#ifdef ICB_TIMING
    const int dbg_pass = (c1.x - 1) & 7;
    if (blockIdx.x == 0 && threadIdx.x == 0) st->dbg_t[dbg_pass][0] = gtime();
#endif
    // both possible next tables, by the CTA with the least work (the last one has the fewest chunks)
    if (blockIdx.x == gridDim.x - 1) {
        precompute_next_tables(st, tid);
        if (tid < 32)               // the tables this pass codes with, kept beside its stream for the header
            reinterpret_cast<unsigned int *>(&st->cand_qtab[c2.x][0][0])[tid] =
                reinterpret_cast<const unsigned int *>(&st->tab[c2.z].qtab[0][0])[tid];
    }

    const uint32_t zrl0 = s_huf[0xF0], zrl1 = s_huf[512 + 0xF0];
    const uint32_t eob0 = s_huf[0], eob1 = s_huf[512];
    const uint32_t lt_mask = (1u << lane) - 1u;

    // persistent CTAs: chunks of 8 consecutive tiles (one per warp) are dealt out statically (no
    // shared cursor a late CTA could overrun); nobody ever waits for another CTA
    int round = 0;
    for (int chunk = blockIdx.x; chunk < p.nchunks; chunk += gridDim.x, round++) {
      const int tile = chunk * ENT_WARPS + warp;
      const int slot = round & (ENT_RING - 1);
      if (round && slot == 0) __syncthreads();                  // the record ring starts a new lap (rare: > 16 rounds)
      // The tile's units (one or two lists, each zero-padded to whole 64-entry rounds) are coded as one
      // run of rounds.  Counts and the first round are requested together: every unit has at least one
      // DC entry, so its first round exists.  (Deeper look-ahead -- four rounds, or one tile ahead -- was
      // measured slower, alone and with several fits sharing the GPU: the loop is issue bound.)
      // Plain (.cg) loads, not the non-coherent path: the lists are rewritten by kernels of this same
      // stream chain that may still be running when this grid becomes resident.
#ifdef ICB_PASS_LDG          // A/B only: the non-coherent path the round-1 kernel used
#define ICB_ENT_LD __ldg
#else
#define ICB_ENT_LD __ldcg
#endif
      const int u0 = tile * p.unit_group;
      uint32_t rounds0 = 0, rounds = 0;
      uint2 e_first = make_uint2(0u, 0u);
      if (tile < p.ntiles) {
          rounds0 = (ICB_ENT_LD(p.ent_cnt + u0) + 63u) >> 6;
          rounds = rounds0;
          if (p.unit_group > 1 && u0 + 1 < p.nunits) rounds += (ICB_ENT_LD(p.ent_cnt + u0 + 1) + 63u) >> 6;
          e_first = ICB_ENT_LD(reinterpret_cast<const uint2 *>(p.ent + (size_t)u0 * UNIT_CAP) + lane);
      }
      ICB_STAMP();                  // 5 / 10: tile start
      if (tile < p.ntiles) {
        uint32_t *scr = pb.scratch + (size_t)tile * p.scr_stride;
        uint32_t running = 0;       // bits produced so far (tile-local)
        uint32_t flushed = 0;       // whole words already copied to scratch
        uint32_t acc[4] = {0, 0, 0, 0};     // 0xFF counters, 2 x 16 bits each: local start bit s = 0,1 | 2,3 | 4,5 | 6,7
                                            // (a lane sees <= 104 words of a 64-block tile, <= 4 starts per word and s)
        uint32_t carry = 126u;      // key of the last live entry (126 = pos 63: no EOB owed)
        uint32_t first_word = 0;    // word 0 of the tile stream (for the leading-ones count)

        // 0xFF bytes that START in window words [0, nw) -- word w+1 must be final already --
        // for the 8 local start bits s; limit = tile bits that exist (a byte must end <= limit)
        // ICB_PASS_BITS_ONLY (A/B build): the pass stops at the code lengths -- no bit window, no
        // flush, no 0xFF count.  Sizes are then wrong; the build exists to time what a bit-length-only sizing
        // pass (the north-star's "bit-length pass with a prefix scan") would cost next to the exact one.
        auto count_ff = [&](uint32_t nw, uint32_t limit) {
#ifdef ICB_PASS_BITS_ONLY
            return;
#endif
            for (uint32_t w = lane; w < nw; w += 32) {
                uint32_t m = ff_starts(s_win[wbase + w], s_win[wbase + w + 1]);
                const uint32_t bit0 = (flushed + w) << 5;
                const uint32_t rem = limit > bit0 ? limit - bit0 : 0u;      // bits from this word on
                if (rem < 40u) m &= rem < 8u ? 0u : 0xffffffffu << (39u - rem);
                if (m) {
#pragma unroll
                    for (uint32_t s = 0; s < 4; s++) {
                        acc[s >> 1] += __popc(m & (0x80808080u >> s)) << (16 * (s & 1));
                        acc[2 + (s >> 1)] += __popc(m & (0x08080808u >> s)) << (16 * (s & 1));
                    }
                }
            }
        };

        // copy finished words of the window to scratch and slide the window; one finished word
        // stays behind so that bytes straddling it can still be counted
        auto flush_some = [&]() {
#ifdef ICB_PASS_BITS_ONLY
            flushed = (running >> 5) > 2 ? (running >> 5) - 2 : 0;
            return;
#endif
            __syncwarp();
            const uint32_t have = (running >> 5) - flushed;         // finished words in the window
            if (have >= 2) {
                if (flushed == 0) first_word = s_win[wbase];
                const uint32_t nf = have - 1;
                count_ff(nf, 0xffffffffu);
                const uint32_t c0 = s_win[wbase + nf], c1 = s_win[wbase + nf + 1], c2 = s_win[wbase + nf + 2];
                for (uint32_t w = lane; w < nf; w += 32) scr[flushed + w] = s_win[wbase + w];
                __syncwarp();
                for (uint32_t w = 4 * lane; w < nf + 3; w += 128)           // (words past nf + 2 are zero already)
                    *reinterpret_cast<uint4 *>(s_win + wbase + w) = make_uint4(0, 0, 0, 0);
                __syncwarp();
                if (lane == 0) { s_win[wbase] = c0; s_win[wbase + 1] = c1; s_win[wbase + 2] = c2; }
                flushed += nf;
            }
            __syncwarp();
        };

        // OR a word into the window.  or_nz skips a zero word behind a branch (ptxas turns even a
        // predicated red into divergence barrier + branch + reconvergence); or_any always issues the
        // reduction -- cheaper where the word is rarely zero: the first two words of a 64-bit field
        // (measured: pass 25.6 -> 24.8 us alone; -DICB_PASS_OR_BRANCH restores the branches)
        auto or_nz = [&](uint32_t word_idx, uint32_t w) {
#ifdef ICB_PASS_BITS_ONLY
            return;
#endif
            if (w) atomicOr(&s_win[wbase + word_idx], w);
        };
        auto or_any = [&](uint32_t word_idx, uint32_t w) {
#ifdef ICB_PASS_BITS_ONLY
            return;
#endif
#ifndef ICB_PASS_OR_BRANCH
            asm volatile("red.shared.or.b32 [%0], %1;" :: "r"(win_base + 4u * word_idx), "r"(w) : "memory");
#else
            or_nz(word_idx, w);
#endif
        };

        // OR `len` (1..32) bits of sym into the window at tile bit offset off
        auto put = [&](uint32_t sym, uint32_t len, uint32_t off) {
            const uint32_t V = sym << (32u - len);
            const uint32_t a = (off >> 5) - flushed, sh = off & 31u;
            or_nz(a, V >> sh);
            or_nz(a + 1, __funnelshift_r(0u, V, sh));
        };

        // OR `len` (1..64) bits of sym into the window at tile bit offset off (two symbols of a lane)
        auto put64 = [&](unsigned long long sym, uint32_t len, uint32_t off) {
            const unsigned long long V = sym << (64u - len);
            const uint32_t hi = (uint32_t)(V >> 32), lo = (uint32_t)V;
            const uint32_t a = (off >> 5) - flushed, sh = off & 31u;
            or_any(a, hi >> sh);
            or_any(a + 1, __funnelshift_r(lo, hi, sh));
            or_nz(a + 2, __funnelshift_r(0u, lo, sh));
        };

        // a2.5 quantise one survivor at the current tables: sign(x) * ((|x| + 4q) / 8q), the division
        // as a multiply-high; a DC entry becomes the difference against its predictor block
        auto quant = [&](uint32_t e, bool &isdc, uint32_t &key, uint32_t &plane) -> int {
            isdc = (e & ENTRY_DC) != 0u;
            plane = (e >> 8) & 0x300u;                                   // (2 * table + isdc) << 8: the entry's code table
            key = (e >> 17) & (isdc ? 1u : 127u);                        // 2 * pos + table
            const int x = (int)(short)(e & 0xffffu);
            const uint2 qt = s_qt[key];
            const uint32_t qa = __umulhi((uint32_t)abs(x) + qt.y, qt.x);
            const int v = x < 0 ? -(int)qa : (int)qa;
            // straight-line on purpose (DC and AC lanes share every round: a branch would run both sides)
            const int xp = (int)e >> 18;                                 // predictor DC; position bits on an AC entry
            const uint32_t qp = __umulhi((uint32_t)abs(xp) + qt.y, qt.x);
            const int vp = xp < 0 ? -(int)qp : (int)qp;
            return isdc ? v - vp : v;
        };
        // a2.7 the code of one entry whose previous live entry had key pk: symbol bits, their length
        // (0 for a dead entry), and the number of ZRL codes (16-zero runs) that go in front of it
        auto mksym = [&](bool live, bool isdc, uint32_t key, uint32_t plane, int v, uint32_t pk, uint32_t &sym,
                         uint32_t &len, uint32_t &nz) {
            const uint32_t nb = 32u - (uint32_t)__clz(abs(v));
            const uint32_t vbits = (uint32_t)(v + (v >> 31)) & ((1u << nb) - 1u);
            const uint32_t r = ((key - pk) >> 1) - 1u;                   // zero run in front of an AC coefficient
            // second index of the code table: the run of an AC entry; for a DC entry whether the previous block
            // ended early and in which table (its EOB then sits in front of the DC code, pre-joined in the table)
            const uint32_t owed = pk < 126u ? 0x10u + ((pk & 1u) << 4) : 0u;
            const uint32_t code = s_huf[plane | (isdc ? owed : (r & 15u) << 4) | nb];
            sym = ((code >> 8) << nb) | vbits;
            len = (code & 0xffu) + nb;
            nz = (isdc || !live) ? 0u : r >> 4;
            len = live ? len : 0u;
            sym = live ? sym : 0u;
        };

        // one round = 64 consecutive entries, two per lane (lane l: entries 2l, 2l+1): the scan of the
        // code lengths, the look-up of the previous live key and the window bookkeeping are paid once
        // per pair, and a pair's two codes usually go into the window as one 64-bit field
        auto code_round = [&](const uint2 ee) {
            bool dc_a, dc_b;
            uint32_t key_a, key_b, pl_a, pl_b;
            const int v_a = quant(ee.x, dc_a, key_a, pl_a), v_b = quant(ee.y, dc_b, key_b, pl_b);
            const bool live_a = dc_a || v_a != 0, live_b = dc_b || v_b != 0;   // entries past n are 0: dead
            const uint32_t bal = __ballot_sync(0xffffffffu, live_a || live_b);
            if (bal == 0u) return;
            const uint32_t last_key = live_b ? key_b : key_a;           // of this lane, if it has a live entry
            const uint32_t below = bal & lt_mask;
            uint32_t pk_a = __shfl_sync(0xffffffffu, last_key, (31 - __clz(below)) & 31);
            if (!below) pk_a = carry;                                   // key of the previous live entry
            const uint32_t pk_b = live_a ? key_a : pk_a;
            uint32_t sym_a, len_a, nz_a, sym_b, len_b, nz_b;
            mksym(live_a, dc_a, key_a, pl_a, v_a, pk_a, sym_a, len_a, nz_a);
            mksym(live_b, dc_b, key_b, pl_b, v_b, pk_b, sym_b, len_b, nz_b);
            const uint32_t zl_a = nz_a * (((key_a & 1u) ? zrl1 : zrl0) & 0xffu);
            const uint32_t zl_b = nz_b * (((key_b & 1u) ? zrl1 : zrl0) & 0xffu);
            const uint32_t tot = len_a + zl_a + len_b + zl_b;
            uint32_t incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            uint32_t off = running + incl - tot;
            running += __shfl_sync(0xffffffffu, incl, 31);
            if (nz_a | nz_b) {                                          // rare: a run of 16+ zeros
                const uint32_t za = (key_a & 1u) ? zrl1 : zrl0, zb = (key_b & 1u) ? zrl1 : zrl0;
                for (uint32_t k = 0; k < nz_a; k++) { put(za >> 8, za & 0xffu, off); off += za & 0xffu; }
                if (len_a) { put(sym_a, len_a, off); off += len_a; }
                for (uint32_t k = 0; k < nz_b; k++) { put(zb >> 8, zb & 0xffu, off); off += zb & 0xffu; }
                if (len_b) put(sym_b, len_b, off);
            } else if (len_a + len_b) {
                put64(((unsigned long long)sym_a << len_b) | sym_b, len_a + len_b, off);
            }
            carry = __shfl_sync(0xffffffffu, last_key, 31 - __clz(bal));
            if ((running >> 5) - flushed > (uint32_t)(ENT_WIN - 128)) flush_some();   // a round adds <= 118 words
        };

        const uint2 *ent2 = reinterpret_cast<const uint2 *>(p.ent + (size_t)u0 * UNIT_CAP) + lane;
        uint2 e_next = e_first;
        for (uint32_t r = 0; r < rounds; r++) {
            const uint2 e = e_next;
            const uint32_t rn = r + 1;                                  // next round's entries fly while this one is coded
            if (rn < rounds) e_next = ICB_ENT_LD(rn < rounds0 ? ent2 + 32u * rn : ent2 + UNIT_CAP / 2 + 32u * (rn - rounds0));
            code_round(e);                                              // (tails are zero-padded by the list writers)
        }
        ICB_STAMP();                // 7 / 12: rounds done
        if (carry < 126u) {                                     // EOB of the tile's last block
            const uint32_t z = (carry & 1u) ? eob1 : eob0;
            if (lane == 0) put(z >> 8, z & 0xffu, running);
            running += z & 0xffu;
        }
        __syncwarp();
        // final flush: finished words + the partial one + a zero word (emit reads one word ahead)
        const uint32_t nw = ((running + 31) >> 5) - flushed;
        if (flushed == 0) first_word = s_win[wbase];
        count_ff(nw, running);
        for (uint32_t w = lane; w < nw + 1; w += 32) scr[flushed + w] = s_win[wbase + w];
        uint32_t f[8];
        if (__any_sync(0xffffffffu, (acc[0] | acc[1] | acc[2] | acc[3]) != 0u)) {
#pragma unroll
            for (int s = 0; s < 8; s++) {
                const uint32_t a = (acc[s >> 1] >> (16 * (s & 1))) & 0xffffu;
                f[s] = __reduce_add_sync(0xffffffffu, a);
            }
        } else {
#pragma unroll
            for (int s = 0; s < 8; s++) f[s] = 0;
        }
        uint32_t tail_bits = 0;
        if (lane == 0 && running) {
            const uint32_t eb = running >= 8u ? running - 8u : 0u;
            const uint32_t w = wbase + (eb >> 5) - flushed, sh = eb & 31;
            const unsigned long long two = ((unsigned long long)s_win[w] << 32) | s_win[w + 1];
            const uint32_t nbv = running >= 8u ? 8u : running;
            const uint32_t last = (uint32_t)(two >> (64 - nbv - sh)) & ((1u << nbv) - 1u);
            tail_bits = min(nbv, (uint32_t)(__ffs(~last) - 1));
        }
        __syncwarp();
        for (uint32_t w = 4 * lane; w < nw + 3; w += 128)                      // window clean for the next tile
            *reinterpret_cast<uint4 *>(s_win + wbase + w) = make_uint4(0, 0, 0, 0);   // (rounds up: the stride has room)
        if (lane == 0) {
            // counters are indexed by the tile's absolute start offset r = (8 - s) & 7
            const uint4 fpk = make_uint4(f[0] | (f[7] << 16), f[6] | (f[5] << 16), f[4] | (f[3] << 16),
                                         f[2] | (f[1] << 16));
            uint32_t head = min(8u, (uint32_t)__clz(~first_word));
            if (running < 8u) head = min(head, running);
            const uint32_t on = head | (tail_bits << 8);
            pb.ff[tile] = fpk;                                  // for k_entropy_emit (next kernel: no fence needed)
            pb.L[tile] = running;
            pb.ones[tile] = on;
            *reinterpret_cast<uint4 *>(&s_rec[slot][warp][0]) = fpk;
            s_rec[slot][warp][4] = running;
            s_rec[slot][warp][5] = on;
        }
        ICB_STAMP();                // 8 / 12: tile record written

        // ---- the last warp of the chunk to finish folds it; nobody waits (the warps of a CTA drift
        // apart freely: a barrier here costs throughput when several fits share the GPU)
        const int g_n = min(ENT_WARPS, p.ntiles - chunk * ENT_WARPS);
        uint32_t arrived = 0;
        if (lane == 0) {
            __threadfence_block();
            arrived = atomicAdd(&s_arr[slot], 1u);
            if (arrived == (uint32_t)(g_n - 1)) s_arr[slot] = 0;       // ready for the ring's next lap
            __threadfence_block();
        }
        arrived = __shfl_sync(0xffffffffu, arrived, 0);
        if (arrived == (uint32_t)(g_n - 1)) {
            const bool have = lane < g_n;
            const volatile uint32_t *rec = &s_rec[slot][lane & (ENT_WARPS - 1)][0];
            const uint4 f_l = have ? make_uint4(rec[0], rec[1], rec[2], rec[3]) : make_uint4(0, 0, 0, 0);
            const uint32_t L_l = have ? rec[4] : 0u;
            const uint32_t on_l = have ? rec[5] : 0u;
            uint32_t inclL = L_l;
#pragma unroll
            for (int o = 1; o < ENT_WARPS; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, inclL, o);
                if (lane >= o) inclL += t;
            }
            const uint32_t ex = inclL - L_l;
            const uint32_t on_prev = __shfl_up_sync(0xffffffffu, on_l, 1);
            uint32_t tot8[8];
#pragma unroll
            for (uint32_t R = 0; R < 8; R++) {                          // R = the chunk's absolute start offset mod 8
                const uint32_t r = (R + ex) & 7u;
                uint32_t c = have ? ff_select(f_l, r) : 0u;
                if (have && lane > 0) c += span_ff(on_prev, on_l, r);
                tot8[R] = __reduce_add_sync(0xffffffffu, c);
            }
            const uint32_t Lc = __shfl_sync(0xffffffffu, inclL, g_n - 1);
            const uint32_t head = __shfl_sync(0xffffffffu, on_l, 0) & 0xffu;
            const uint32_t tail = __shfl_sync(0xffffffffu, on_l, g_n - 1) >> 8;
            if (lane == 0) {                                    // up to 8 x 13312 stuffed bytes per counter: 32 bits
                pb.crec[3 * chunk] = make_uint4(tot8[0], tot8[1], tot8[2], tot8[3]);
                pb.crec[3 * chunk + 1] = make_uint4(tot8[4], tot8[5], tot8[6], tot8[7]);
                pb.crec[3 * chunk + 2] = make_uint4(Lc, head | (tail << 8), 0u, 0u);
            }
        }
      }
    }

    // ---- this CTA is done: one fence and one ticket per CTA (the barrier makes every warp's chunk
    // records part of what thread 0 releases); the last CTA of the pass scans the chunk records
    // into prefixes and the exact stuffed size, then runs the bisection step
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        s_last = atomicAdd(&st->done_a, 1u) == gridDim.x - 1u ? 1 : 0;
    }
    __syncthreads();
    ICB_STAMP();                    // CTA ticket drawn
    if (!s_last) return;
    __threadfence();
#ifdef ICB_TIMING
    if (tid == 0) st->dbg_t[dbg_pass][1] = gtime();
#endif
    {
        const int n_c = p.nchunks;
        const int per = (n_c + ENT_THREADS - 1) / ENT_THREADS;          // consecutive chunk records per thread
        const int c0 = min(n_c, tid * per), c1 = min(n_c, c0 + per);
        const uint32_t *rw = reinterpret_cast<const uint32_t *>(pb.crec);
        // bits in front of this thread's records
        unsigned long long mine = 0;
        for (int c = c0; c < c1; c++) mine += rw[12 * c + 8];            // L1 was invalidated by the fence above;
                                                                        // the second sweep below hits it
        unsigned long long incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_scan[warp] = incl;
        __syncthreads();
        unsigned long long P = incl - mine;
        for (int w = 0; w < warp; w++) P += s_scan[w];
        __syncthreads();
        // 0xFF bytes of each record at its alignment; first pass leaves thread-local prefixes in Fc
        uint32_t on_prev = c0 > 0 && c0 < n_c ? rw[12 * (c0 - 1) + 9] : 0u;
        uint32_t nff = 0;
        for (int c = c0; c < c1; c++) {
            const uint32_t Lc = rw[12 * c + 8], on = rw[12 * c + 9];
            const uint32_t r = (uint32_t)P & 7u;
            uint32_t cnt = rw[12 * c + r];                              // the counter for this alignment
            if (c > 0) cnt += span_ff(on_prev, on, r);
            pb.Pc[c] = P;
            pb.Fc[c] = nff;
            nff += cnt;
            P += Lc;
            on_prev = on;
        }
        uint32_t inclf = nff;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inclf, o);
            if (lane >= o) inclf += t;
        }
        if (lane == 31) s_scan[warp] = inclf;
        __syncthreads();
        uint32_t basef = inclf - nff;
        for (int w = 0; w < warp; w++) basef += (uint32_t)s_scan[w];
        pb.Fb[tid] = basef;                                             // Fc of a chunk = Fc[chunk] + Fb[chunk / per]
        if (c1 == n_c && c0 < n_c) {                                    // owner of the last record
            uint32_t total_ff = basef + nff;
            const uint32_t rem = (uint32_t)P & 7u;                      // jchuff.c flush_bits: last byte padded with 1s
            if (rem && (on_prev >> 8) >= rem) total_ff++;
            s_total[0] = P;
            s_total[1] = total_ff;
        }
        __syncthreads();
    }
#ifdef ICB_TIMING
    if (tid == 0) st->dbg_t[dbg_pass][2] = gtime();
#endif
    if (warp == 0) device_step(st, reinterpret_cast<const FitState *>(s_ctl), 0, s_total[0], s_total[1]);      // :176-189 on the exact size, then the next table slot
#ifndef ICB_NO_KCLK
    if (tid == 0) st->kt[kt_slot][1] = gtime();
#endif
#ifdef ICB_TIMING
    if (tid == 0) st->dbg_t[dbg_pass][3] = gtime();
#endif
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

__device__ __forceinline__ TileView tile_view(const PassBufs &pb, int tile, int ntiles, int scr_stride, unsigned long long P,
                                              uint32_t L, uint32_t Lp)
{
    TileView v;
    v.scr = pb.scratch + (size_t)tile * scr_stride;
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
        const uint32_t *sp = pb.scratch + (size_t)(tile - 1) * scr_stride;
        const uint32_t o = Lp - v.r, w = o >> 5, sh = o & 31;
        const unsigned long long two = ((unsigned long long)__ldcg(sp + w) << 32) | __ldcg(sp + w + 1);
        v.prev_tail = (uint32_t)(two >> (64 - v.r - sh)) & ((1u << v.r) - 1u);
    }
    return v;
}

// word w of the tile's byte stream (4 bytes, big-endian), and how many of them are real
__device__ __forceinline__ uint32_t tile_word(const TileView &v, uint32_t w, uint32_t &nvalid)
{
    const uint32_t hi = w ? __ldcg(v.scr + w - 1) : v.prev_tail;
    uint32_t val = __funnelshift_r(__ldcg(v.scr + w), hi, v.r);
    if (w == v.pad_word) val |= v.pad_mask;
    nvalid = min(4u, v.nbytes - 4 * w);
    if (nvalid < 4) val &= 0xffffffffu << (8 * (4 - nvalid));
    return val;
}

// header template (tables, dims and JFIF minor patched per image); filled by the host
__constant__ unsigned char c_hdr[HEADER_BYTES];
constexpr int HDR_OFF_MINOR = 12, HDR_OFF_DQT0 = 25, HDR_OFF_DQT1 = 94, HDR_OFF_H = 163, HDR_OFF_W = 165;

// one tile of the chosen pass: its bits at their absolute alignment, stuffed, at their final offsets (one warp)
__device__ __forceinline__ void emit_tile(const EntParams &p, const PassBufs &pb, const FitState *st, unsigned char *out,
                                          long long cap, int tile, int lane)
{
    // this tile's bit prefix and the 0xFF bytes completed before it: the chunk prefixes of the pass
    // plus the tiles in front of it inside its chunk of 8 (lane = tile of the chunk)
    const int chunk = tile / ENT_WARPS, t_first = chunk * ENT_WARPS, me = tile - t_first;
    const bool have = lane < ENT_WARPS && t_first + lane < p.ntiles;
    const uint32_t L_l = have ? __ldcg(pb.L + t_first + lane) : 0u;
    const uint4 f_l = have ? __ldcg(pb.ff + t_first + lane) : make_uint4(0, 0, 0, 0);
    const uint32_t on_l = have ? __ldcg(pb.ones + t_first + lane) : 0u;
    uint32_t inclL = L_l;
#pragma unroll
    for (int o = 1; o < ENT_WARPS; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inclL, o);
        if (lane >= o) inclL += t;
    }
    const unsigned long long P_l = __ldcg(pb.Pc + chunk) + (inclL - L_l);
    const uint32_t r_l = (uint32_t)P_l & 7u;
    uint32_t on_prev = __shfl_up_sync(0xffffffffu, on_l, 1);
    if (lane == 0) on_prev = t_first > 0 ? __ldcg(pb.ones + t_first - 1) : 0u;
    const uint32_t c_l = have ? ff_select(f_l, r_l) + ((t_first + lane > 0) ? span_ff(on_prev, on_l, r_l) : 0u) : 0u;
    uint32_t inclC = c_l;
#pragma unroll
    for (int o = 1; o < ENT_WARPS; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inclC, o);
        if (lane >= o) inclC += t;
    }
    const int per = (p.nchunks + ENT_THREADS - 1) / ENT_THREADS;        // chunk records per thread of the final scan
    const uint32_t ffb_l = __ldcg(pb.Fc + chunk) + __ldcg(pb.Fb + chunk / per) + (inclC - c_l);
    const unsigned long long P = __shfl_sync(0xffffffffu, P_l, me);
    const uint32_t L = __shfl_sync(0xffffffffu, L_l, me);
    const uint32_t ff_before = __shfl_sync(0xffffffffu, ffb_l, me);
    uint32_t Lp = __shfl_sync(0xffffffffu, L_l, (me + 31) & 31);
    if (me == 0 && tile > 0) Lp = __ldcg(pb.L + tile - 1);

    const TileView v = tile_view(pb, tile, p.ntiles, p.scr_stride, P, L, Lp);
    const uint32_t nwords = (v.nbytes + 3) >> 2;
    const long long pos0 = (long long)HEADER_BYTES + (long long)(P >> 3) + (long long)ff_before;
    uint32_t ff_run = 0;
    for (uint32_t wb = 0; wb < nwords; wb += 32) {
        const uint32_t w = wb + lane;
        uint32_t val = 0, cntf = 0, nvalid = 0;
        if (w < nwords) {
            val = tile_word(v, w, nvalid);
            cntf = ff_bytes(val);
        }
        uint32_t incl = cntf;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        long long pos = pos0 + 4ll * w + ff_run + (incl - cntf);
        for (uint32_t bi = 0; bi < nvalid; bi++) {
            const uint32_t byte = (val >> (24 - 8 * bi)) & 0xffu;
            if (pos < cap) out[pos] = (unsigned char)byte;
            pos++;
            if (byte == 0xffu) {
                if (pos < cap) out[pos] = 0;
                pos++;
            }
        }
        ff_run += __shfl_sync(0xffffffffu, incl, 31);
    }
#ifdef ICB_TIMING
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) const_cast<FitState *>(st)->dbg_k[2] = gtime();
#endif
    if (v.last && lane == 0) {                                  // EOI
        const long long e = st->best_size - 2;
        if (e >= 0 && e < cap) out[e] = 0xff;
        if (e + 1 >= 0 && e + 1 < cap) out[e + 1] = 0xd9;
    }
}

// once per image: header (a2.8), a2.7 emit_bits byte stuffing of the chosen pass, EOI
__global__ void __launch_bounds__(ENT_THREADS) k_entropy_emit(const EntParams p, unsigned char *out, long long cap,
                                                              int img_w, int img_h)
{
    const FitState *st = p.st;
    pdl_wait();
    if (st->best_size <= 0 || out == nullptr) return;
    if (blockIdx.x == 0 && threadIdx.x == 0) const_cast<FitState *>(st)->kt[10][0] = gtime();
#ifdef ICB_TIMING
    if (blockIdx.x == 0 && threadIdx.x == 0) const_cast<FitState *>(st)->dbg_k[1] = gtime();
#endif
    const int best_buf = st->best_buf;
    const PassBufs pb = pick_buf(p, best_buf);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (blockIdx.x == 0) {
        for (int i = tid; i < HEADER_BYTES; i += ENT_THREADS) {
            unsigned char b = c_hdr[i];
            if (i == HDR_OFF_MINOR) b = (unsigned char)st->jfif_minor;
            else if (i >= HDR_OFF_DQT0 && i < HDR_OFF_DQT0 + 64) b = st->cand_qtab[best_buf][0][i - HDR_OFF_DQT0];
            else if (i >= HDR_OFF_DQT1 && i < HDR_OFF_DQT1 + 64) b = st->cand_qtab[best_buf][1][i - HDR_OFF_DQT1];
            else if (i == HDR_OFF_H) b = (unsigned char)(img_h >> 8);
            else if (i == HDR_OFF_H + 1) b = (unsigned char)img_h;
            else if (i == HDR_OFF_W) b = (unsigned char)(img_w >> 8);
            else if (i == HDR_OFF_W + 1) b = (unsigned char)img_w;
            if (i < cap) out[i] = b;
        }
    }
    const int tile = blockIdx.x * ENT_WARPS + warp;
    if (tile < p.ntiles) emit_tile(p, pb, st, out, cap, tile, lane);
}

// After the emit kernel: the chosen stream goes to the caller -- exactly best_size bytes of it, and only if the
// phase succeeded (a missed cached try or a scale that did not fit sends nothing).  The size is known on the
// device only, so the copy is a kernel: 16-byte loads from HBM, 16-byte stores into page-locked host memory
// (zero-copy over PCIe); a cudaMemcpyAsync would have to be sized by the host after a round trip.
constexpr int FETCH_THREADS = 256;
__global__ void __launch_bounds__(FETCH_THREADS) k_fetch_stream(const FitState *st, const unsigned char *d_out, long long d_cap,
                                                                unsigned char *h_out, long long h_cap)
{
    pdl_wait();
    const long long n = st->best_size;
    if (n <= 0 || n > st->target || n > d_cap || n > h_cap) return;
    const long long nvec = n >> 4;
    const uint4 *src = reinterpret_cast<const uint4 *>(d_out);
    uint4 *dst = reinterpret_cast<uint4 *>(h_out);
    for (long long i = (long long)blockIdx.x * FETCH_THREADS + threadIdx.x; i < nvec; i += (long long)gridDim.x * FETCH_THREADS)
        dst[i] = __ldcg(src + i);
    if (blockIdx.x == 0 && threadIdx.x < (n & 15)) h_out[(nvec << 4) + threadIdx.x] = d_out[(nvec << 4) + threadIdx.x];
}

// the same for a byte string whose length sits in device memory (the PNG file, k_png.cuh)
__global__ void __launch_bounds__(FETCH_THREADS) k_fetch_len(const unsigned long long *len, const unsigned char *d_src,
                                                             long long d_cap, unsigned char *h_out, long long h_cap)
{
    const long long n = (long long)*len;
    if (n <= 0 || n > d_cap || n > h_cap) return;
    const long long nvec = n >> 4;
    const uint4 *src = reinterpret_cast<const uint4 *>(d_src);
    uint4 *dst = reinterpret_cast<uint4 *>(h_out);
    for (long long i = (long long)blockIdx.x * FETCH_THREADS + threadIdx.x; i < nvec; i += (long long)gridDim.x * FETCH_THREADS)
        dst[i] = __ldcg(src + i);
    if (blockIdx.x == 0 && threadIdx.x < (n & 15)) h_out[(nvec << 4) + threadIdx.x] = d_src[(nvec << 4) + threadIdx.x];
}

// per-slot arenas
__host__ __device__ inline size_t ent_pass_bytes(int ntiles, int nchunks, int scr_stride)
{
    const size_t n = (size_t)ntiles, c = (size_t)nchunks;
    return n * (size_t)scr_stride * 4 + n * 4 + n * 16 + n * 4 + c * 48 + c * 8 + c * 4 + ENT_THREADS * 4 + 256;
}

// survivor lists + counts + unit links of one image
__host__ __device__ inline size_t ent_list_bytes(int nunits)
{
    const size_t n = (size_t)nunits;
    return n * (size_t)UNIT_CAP * 4 + ((n + 3) / 4 * 4) * 4 + n * 32 + 256;
}

}  // namespace icb
