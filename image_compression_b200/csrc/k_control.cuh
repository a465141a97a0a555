// k_control.cuh -- device-side control of the quality search, so a whole fit
// (transform + up to 8 entropy passes + final stream) runs on one stream without a
// host round trip between passes.
//
// device_step restates core/ImageCompressionJpg.java:158-200
// (findBestQualityByBinarySearch) in float32 exactly as Java evaluates it:
// lo = 0, hi = initialQuality, mid = (lo + hi) / 2.0f, `mid < 0.01f` break before
// encoding, fit -> best = lo = mid else hi = mid, `(hi - lo) < 0.01f` break after.
// It runs inside the last CTA of k_entropy_pass (the exact stuffed size of the
// pass is known there) and once in begin_job (k_entropy_prepare), and builds the quantisation tables of
// the coming pass (a2.1: JDK JPEG.convertToLinearQuality +
// JPEGQTable.getScaledInstance in float32, or the IJG integer formula under the
// alternative profile).  Each pass keeps its bit stream in the scratch buffer that
// does not hold the best pass so far, so the reference's extra "save" encode
// (:255-260) costs only the final emission, not another entropy pass.
#pragma once

#include "common.cuh"
#include "k_entropy.cuh"

namespace icb {

// a2.1 -- one table entry for quality q
__device__ __forceinline__ int scaled_qval(int base, float q, int table_mode)
{
    int sv;
    if (table_mode == 0) {
        float qq = q;
        if (qq <= 0.0f) qq = 0.01f;
        if (qq > 1.00f) qq = 1.00f;
        const float s = qq < 0.5f ? __fdiv_rn(0.5f, qq) : __fsub_rn(2.0f, __fmul_rn(qq, 2.0f));
        sv = (int)__fadd_rn(__fmul_rn((float)base, s), 0.5f);
    } else {
        int quality = (int)__fadd_rn(__fmul_rn(q, 100.0f), 0.5f);
        if (quality <= 0) quality = 1;
        if (quality > 100) quality = 100;
        const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
        sv = (base * scale + 50) / 100;
    }
    if (sv < 1) sv = 1;
    if (sv > 255) sv = 255;
    return sv;
}

// one entry (idx = table * 64 + zigzag k) of the tables for quality q
__device__ __forceinline__ void table_entry(QTables *t, int idx, float q, int table_mode)
{
    const int tb = idx >> 6, k = idx & 63;
    const int nat = c_zigzag[k];
    const int qv = scaled_qval(tb ? c_base_chr[nat] : c_base_lum[nat], q, table_mode);
    const unsigned d = (unsigned)qv << 3;                       // jcdctmgr.c: divisor = quantval << 3
    // floor(2^32 / d) + 1 without a 64-bit division: umulhi(t, recip) == t / d for t < 2^17
    const unsigned recip = 0xffffffffu / d + 1u + ((d & (d - 1u)) == 0u ? 1u : 0u);
    t->rh[2 * k + tb] = make_uint2(recip, d >> 1);
    t->qtab[tb][k] = (unsigned char)qv;
}

// Both midpoints the bisection can ask for after the running pass (:168 with :180-185 applied),
// computed by 256 threads of one CTA while the pass runs: set 0 = this pass fits, set 1 = it does not.
__device__ __forceinline__ void precompute_next_tables(FitState *st, int tid)
{
    if (st->mode != MODE_SEARCH || tid >= 256) return;
    const int set = tid >> 7;
    const float lo = set ? st->lo : st->cur_q, hi = set ? st->cur_q : st->hi;
    const float mid = __fdiv_rn(__fadd_rn(lo, hi), 2.0f);
    table_entry(&st->tab[(st->tsel + 1 + set) % 3], tid & 127, mid, st->table_mode);
}

// Called by all 32 lanes of ONE warp.  Consumes the pass that just ran (if any),
// decides, and selects the tables of the next one.  Lane 0 reads every control field
// first and writes them back at the end, so the global accesses pipeline.
__device__ void device_step(FitState *st, const FitState *snap, int first, unsigned long long tbits,
                            unsigned long long ffs)
{
    const int lane = threadIdx.x & 31;
    int go = 0, better = 0;
    if (lane == 0) {
        // snap = the control block (first 88 bytes) as it was when the pass started: nothing but this
        // function writes it, so the pass hands over the copy it made in shared memory
        float lo = snap->lo, hi = snap->hi, best = snap->best, cur_q = snap->cur_q;
        int iter = snap->iter, done = snap->done, n_trace = snap->n_trace, passes = snap->passes;
        int cur_buf = snap->cur_buf, best_buf = snap->best_buf, tsel = snap->tsel;
        const int cur_valid = snap->cur_valid, mode = snap->mode;
        const long long target = snap->target;
        long long best_size = snap->best_size, cur_size = snap->cur_size;
        int fit = 0;
        if (cur_valid) {                                        // consume the pass that just ran
            const long long size = (long long)HEADER_BYTES + (long long)((tbits + 7ull) >> 3) + (long long)ffs + 2;
            cur_size = size;
            passes++;
            if (mode == MODE_SEARCH) {
                if (n_trace < 8) {
                    st->trace_q[n_trace] = cur_q;
                    st->trace_size[n_trace] = size;
                    n_trace++;
                }
                if (size <= target) {                           // :180-185
                    best = cur_q;
                    lo = cur_q;
                    better = 1;
                    fit = 1;
                } else {
                    hi = cur_q;
                }
                if (__fsub_rn(hi, lo) < 0.01f) done = 1;        // :187-189
            } else {
                best = cur_q;
                better = 1;
                done = 1;
            }
            if (better) {
                best_buf = cur_buf;
                best_size = size;
            }
        }
        if (!done) {
            if (mode == MODE_SEARCH) {
                if (iter >= 8) {                                // :167
                    done = 1;
                } else {
                    const float mid = __fdiv_rn(__fadd_rn(lo, hi), 2.0f);   // :168
                    if (mid < 0.01f) {                          // :170-172
                        done = 1;
                    } else {
                        cur_q = mid;
                        go = 1;
                    }
                }
            } else {
                go = iter == 0;
            }
        }
        if (go) {
            iter++;
            cur_buf = (best_size > 0) ? (1 - best_buf) : 0;
            if (!first) tsel = (tsel + (fit ? 1 : 2)) % 3;       // precompute_next_tables filled that slot
        }
        st->lo = lo; st->hi = hi; st->best = best; st->cur_q = cur_q;
        st->iter = iter; st->done = done; st->n_trace = n_trace; st->passes = passes;
        st->cur_buf = cur_buf; st->best_buf = best_buf; st->cur_valid = go; st->tsel = tsel;
        st->best_size = best_size; st->cur_size = cur_size;
        st->done_a = 0;
        st->ff_total = ffs;
        st->total_bits = tbits;
    }
    __threadfence();
    __syncwarp();
}

// Called by all 32 lanes of ONE warp of k_entropy_prepare: the state a job starts from
// (:160-162), the first bisection step and the tables of the first pass.
__device__ void begin_job(FitState *st, const BeginArgs &b)
{
    const int lane = threadIdx.x & 31;
    if (lane == 0) {
        st->list_q[0] = b.list_q;
        st->lo = 0.0f;
        st->hi = b.q;
        st->best = -1.0f;
        st->cur_q = b.q;
        st->iter = 0;
        st->done = 0;
        st->cur_valid = 0;
        st->mode = b.mode;
        st->cur_buf = 0;
        st->best_buf = 0;
        st->tsel = 0;
        st->n_trace = 0;
        st->table_mode = b.table_mode;
        st->jfif_minor = b.jfif_minor;
        st->passes = 0;
        st->target = b.target;
        st->cur_size = 0;
        st->best_size = 0;
    }
    __threadfence();
    __syncwarp();
    device_step(st, st, 1, 0ull, 0ull);
    int go = 0;
    float q0 = 0.0f;
    if (lane == 0) { go = st->cur_valid; q0 = st->cur_q; }
    go = __shfl_sync(0xffffffffu, go, 0);
    q0 = __shfl_sync(0xffffffffu, q0, 0);
    if (go)
        for (int t = lane; t < 128; t += 32) table_entry(&st->tab[0], t, q0, b.table_mode);
}

// Job start after a transform that wrote the survivor lists itself (k_transform.cuh): the search state and
// first tables, and the three DC predictors per unit that look into the previous unit (the first MCU's
// Y00, Cb and Cr follow the previous unit's last Y11, Cb and Cr in scan order).
constexpr int BEGIN_THREADS = 256;
__global__ void __launch_bounds__(BEGIN_THREADS) k_job_begin(const EntParams p, const BeginArgs ba)
{
    pdl_launch_dependents();
    if (blockIdx.x == 0 && threadIdx.x < 32) begin_job(p.st, ba);       // touches nothing the transform writes
    pdl_wait();                                                         // lists and unit_info are complete
    if (blockIdx.x == 0 && threadIdx.x == 0) p.st->kt[0][1] = gtime(); // = the moment the transform grid finished
    const int u = blockIdx.x * BEGIN_THREADS + threadIdx.x;
    if (u < 1 || u >= p.nunits) return;
    // two independent 32-byte reads, three plain stores: the entries come complete out of unit_info
    const uint4 *info = reinterpret_cast<const uint4 *>(p.unit_info);
    const uint4 prev = __ldcg(info + 2 * (u - 1));            // .y = Y11 | Cb << 16, .z = Cr (low half)
    const uint4 mine0 = __ldcg(info + 2 * u);                  // .x = Cb index | Cr index << 16, .w = Y00 entry
    const uint4 mine1 = __ldcg(info + 2 * u + 1);              // .x = Cb entry, .y = Cr entry
    uint32_t *list = p.ent + (size_t)u * UNIT_CAP;
    list[0] = mine0.w | ((prev.y & 0x3fffu) << 18);
    list[mine0.x & 0xffffu] = mine1.x | (((prev.y >> 16) & 0x3fffu) << 18);
    list[mine0.x >> 16] = mine1.y | ((prev.z & 0x3fffu) << 18);
}

// Between two passes of a long search: when the upper end of the bracket has dropped well below
// the quality the survivor lists were built for, every midpoint still to come quantises with
// tables at least as coarse as those of hi, so entries below the thresholds of hi are dead for the
// rest of the job.  One warp per unit compacts its list in place (DC entries always stay: they
// carry the block boundaries).  A no-op otherwise; only enqueued for searches the host expects to
// overshoot at their first midpoint (api.cu expect_first_pass_fails).
constexpr float PRUNE_RATIO = 0.8f;

__global__ void __launch_bounds__(ENT_THREADS) k_entropy_prune(const EntParams p, int k)
{
    __shared__ unsigned short s_thr[2][64];
    __shared__ float s_hi;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_launch_dependents();
    pdl_wait();                                     // the step of the pass before this is visible
    FitState *st = p.st;
    if (tid == 0) {
        const float hi = st->hi, lq = st->list_q[k];
        const bool go = !st->done && st->cur_valid && hi <= lq * PRUNE_RATIO;
        s_hi = go ? hi : -1.0f;
        if (blockIdx.x == 0) st->list_q[k + 1] = go ? hi : lq;
    }
    __syncthreads();
    const float hi = s_hi;
    if (hi < 0.0f) return;
    if (tid < 128) {
        const int tb = tid >> 6, z = tid & 63;
        const int nat = c_zigzag[z];
        s_thr[tb][z] = (unsigned short)(scaled_qval(tb ? c_base_chr[nat] : c_base_lum[nat], hi, st->table_mode) << 2);
    }
    __syncthreads();
    const int unit = blockIdx.x * ENT_WARPS + warp;
    if (unit >= p.nunits) return;
    uint32_t *list = p.ent + (size_t)unit * UNIT_CAP;
    const uint32_t n = p.ent_cnt[unit];
    uint32_t w = 0;
    // four chunks of 32 in flight; a chunk's stores land at or below its own first slot, and every
    // lane holds its entry of that chunk (it went into the ballot) before any of them
    for (uint32_t i0 = 0; i0 < n; i0 += 128) {
        uint32_t e[4];
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const uint32_t i = i0 + 32 * c + lane;
            e[c] = i < n ? list[i] : 0u;
        }
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const uint32_t i = i0 + 32 * c + lane;
            const int x = (int)(short)(e[c] & 0xffffu);
            const bool keep = i < n && ((e[c] & ENTRY_DC) != 0u ||
                                        (uint32_t)abs(x) >= s_thr[(e[c] >> 17) & 1u][(e[c] >> 18) & 63u]);
            const uint32_t m = __ballot_sync(0xffffffffu, keep);
            if (keep) list[w + __popc(m & ((1u << lane) - 1u))] = e[c];
            w += __popc(m);
        }
    }
    for (uint32_t i = w + lane; i < ((w + 63u) & ~63u); i += 32) list[i] = 0u;     // dead padding to a whole round (k_entropy_pass)
    if (lane == 0) p.ent_cnt[unit] = w;
}

}  // namespace icb
