// k_control.cuh -- device-side control of the quality search, so a whole fit
// (transform + up to 8 entropy passes + final stream) runs on one stream without a
// host round trip between passes.
//
// device_step restates core/ImageCompressionJpg.java:158-200
// (findBestQualityByBinarySearch) in float32 exactly as Java evaluates it:
// lo = 0, hi = initialQuality, mid = (lo + hi) / 2.0f, `mid < 0.01f` break before
// encoding, fit -> best = lo = mid else hi = mid, `(hi - lo) < 0.01f` break after.
// It runs inside the last CTA of k_entropy_count (the exact stuffed size of the
// pass is known there) and once in k_begin, and builds the quantisation tables of
// the coming pass (a2.1: JDK JPEG.convertToLinearQuality +
// JPEGQTable.getScaledInstance in float32, or the IJG integer formula under the
// alternative profile).  Each pass keeps its bit stream in the scratch buffer that
// does not hold the best pass so far, so the reference's extra "save" encode
// (:255-260) costs only the final emission, not another entropy pass.
#pragma once

#include "common.cuh"
#include "k_entropy.cuh"

namespace icb {

// header template (tables, dims and JFIF minor patched per image); filled by the host
__constant__ unsigned char c_hdr[HEADER_BYTES];
constexpr int HDR_OFF_MINOR = 12, HDR_OFF_DQT0 = 25, HDR_OFF_DQT1 = 94, HDR_OFF_H = 163, HDR_OFF_W = 165;

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

// Called by every thread of ONE CTA (>= 128 threads).  Consumes the pass that just
// ran (if any), decides, and prepares the next one.
__device__ void device_step(FitState *st, int final_call)
{
    __shared__ int s_go;
    if (threadIdx.x == 0) {
        if (st->cur_valid) {                                    // consume the pass that just ran
            const unsigned long long tbits = *(volatile unsigned long long *)&st->total_bits;
            const unsigned long long ffs = *(volatile unsigned long long *)&st->ff_total;
            const long long size = (long long)HEADER_BYTES + (long long)((tbits + 7ull) >> 3) + (long long)ffs + 2;
            bool better = false;
            st->cur_size = size;
            st->passes++;
            st->cur_valid = 0;
            if (st->mode == MODE_SEARCH) {
                if (st->n_trace < 8) {
                    st->trace_q[st->n_trace] = st->cur_q;
                    st->trace_size[st->n_trace] = size;
                    st->n_trace++;
                }
                if (size <= st->target) {                       // :180-185
                    st->best = st->cur_q;
                    st->lo = st->cur_q;
                    better = true;
                } else {
                    st->hi = st->cur_q;
                }
                if (__fsub_rn(st->hi, st->lo) < 0.01f) st->done = 1;   // :187-189
            } else {
                st->best = st->cur_q;
                better = true;
                st->done = 1;
            }
            if (better) {
                st->best_buf = st->cur_buf;
                st->best_size = size;
                for (int i = 0; i < 64; i++) {
                    st->best_qtab[0][i] = st->qtab[0][i];
                    st->best_qtab[1][i] = st->qtab[1][i];
                }
            }
        }
        int go = 0;
        if (!st->done && !final_call) {
            if (st->mode == MODE_SEARCH) {
                if (st->iter >= 8) {                            // :167
                    st->done = 1;
                } else {
                    const float mid = __fdiv_rn(__fadd_rn(st->lo, st->hi), 2.0f);   // :168
                    if (mid < 0.01f) {                          // :170-172
                        st->done = 1;
                    } else {
                        st->cur_q = mid;
                        go = 1;
                    }
                }
            } else {
                go = st->iter == 0;
            }
        }
        if (go) {
            st->iter++;
            st->cur_valid = 1;
            st->cur_buf = (st->best_size > 0) ? (1 - st->best_buf) : 0;
        }
        st->done_a = 0;
        st->done_c = 0;
        st->ff_total = 0;
        st->total_bits = 0;
        s_go = go;
    }
    __syncthreads();
    if (!s_go) return;
    const int t = threadIdx.x;
    if (t < 128) {                                              // 2 tables x 64 entries, zigzag order
        const float q = st->cur_q;
        const int tb = t >> 6, k = t & 63;
        const int nat = c_zigzag[k];
        const int qv = scaled_qval(tb ? c_base_chr[nat] : c_base_lum[nat], q, st->table_mode);
        const unsigned d = (unsigned)qv << 3;                   // jcdctmgr.c: divisor = quantval << 3
        st->qtab[tb][k] = (unsigned char)qv;
        st->recip[tb][k] = (unsigned)((1ull << 32) / d) + 1u;   // umulhi(t, recip) == t / d for t < 2^17
        st->half[tb][k] = (unsigned short)(d >> 1);
    }
}

__global__ void __launch_bounds__(128) k_begin(FitState *st, int mode, float q, long long target, int table_mode,
                                               int jfif_minor)
{
    if (threadIdx.x == 0) {
        st->lo = 0.0f;
        st->hi = q;
        st->best = -1.0f;
        st->cur_q = q;
        st->iter = 0;
        st->done = 0;
        st->cur_valid = 0;
        st->mode = mode;
        st->cur_buf = 0;
        st->best_buf = 0;
        st->epoch = 0;
        st->n_trace = 0;
        st->table_mode = table_mode;
        st->jfif_minor = jfif_minor;
        st->passes = 0;
        st->target = target;
        st->cur_size = 0;
        st->best_size = 0;
        st->emit_valid = 0;
    }
    __syncthreads();
    device_step(st, 0);
}

__global__ void __launch_bounds__(ENT_THREADS) k_finish(const EntParams p, unsigned char *out, long long out_cap,
                                                        int w, int h)
{
    __shared__ unsigned long long s_part[ENT_THREADS + 1];
    FitState *st = p.st;
    if (st->best_size <= 0 || out == nullptr) {
        if (threadIdx.x == 0) st->emit_valid = 0;
        return;
    }
    const PassBufs pb = pick_buf(p, st->best_buf);
    cta_exclusive_scan(pb.bytes, pb.Q, p.ntiles, s_part);
    for (int i = threadIdx.x; i < HEADER_BYTES; i += blockDim.x) {
        unsigned char b = c_hdr[i];
        if (i == HDR_OFF_MINOR) b = (unsigned char)st->jfif_minor;
        else if (i >= HDR_OFF_DQT0 && i < HDR_OFF_DQT0 + 64) b = st->best_qtab[0][i - HDR_OFF_DQT0];
        else if (i >= HDR_OFF_DQT1 && i < HDR_OFF_DQT1 + 64) b = st->best_qtab[1][i - HDR_OFF_DQT1];
        else if (i == HDR_OFF_H) b = (unsigned char)(h >> 8);
        else if (i == HDR_OFF_H + 1) b = (unsigned char)h;
        else if (i == HDR_OFF_W) b = (unsigned char)(w >> 8);
        else if (i == HDR_OFF_W + 1) b = (unsigned char)w;
        if (i < out_cap) out[i] = b;
    }
    if (threadIdx.x == 0) st->emit_valid = 1;
}

}  // namespace icb
