// k_control.cuh -- device-side control of the quality search, so a whole
// fit (transform + up to 8 entropy passes + final stream) runs on one stream
// without a host round trip between passes.
//
// k_step restates core/ImageCompressionJpg.java:158-200
// (findBestQualityByBinarySearch) in float32 exactly as Java evaluates it:
// lo = 0, hi = initialQuality, mid = (lo + hi) / 2.0f, `mid < 0.01f` break before
// encoding, fit -> best = lo = mid else hi = mid, `(hi - lo) < 0.01f` break after.
// It also builds the quantisation tables of the coming pass (a2.1: JDK
// JPEG.convertToLinearQuality + JPEGQTable.getScaledInstance in float32, or the
// IJG integer formula under the alternative profile) and writes the 623-byte
// JFIF header (a2.8) in front of the stream the pass will emit.  Because every
// pass emits its scan into the buffer that does not hold the best stream so far,
// the reference's extra "save" encode (:255-260) needs no pass of its own.
#pragma once

#include "common.cuh"

namespace icb {

// header template (tables, dims and JFIF minor patched per pass); filled by the host
__constant__ unsigned char c_hdr[HEADER_BYTES];
constexpr int HDR_OFF_MINOR = 12, HDR_OFF_DQT0 = 25, HDR_OFF_DQT1 = 94, HDR_OFF_H = 163, HDR_OFF_W = 165;

__global__ void k_init_state(FitState *st, int mode, float q, long long target, int table_mode, int jfif_minor)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
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
    st->ticket = 0;
    st->ff_total = 0;
    st->total_bits = 0;
}

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

// One CTA.  final != 0: only consume the last pass, start nothing.
__global__ void __launch_bounds__(128) k_step(FitState *st, unsigned char *out0, unsigned char *out1,
                                              long long out_cap, int w, int h, int final_call)
{
    __shared__ int s_go;
    if (threadIdx.x == 0) {
        if (st->cur_valid) {                                    // consume the pass that just ran
            const long long size = (long long)HEADER_BYTES + (long long)((st->total_bits + 7ull) >> 3) +
                                   (long long)st->ff_total + 2;
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
                    st->best_buf = st->cur_buf;
                    st->best_size = size;
                } else {
                    st->hi = st->cur_q;
                }
                if (__fsub_rn(st->hi, st->lo) < 0.01f) st->done = 1;   // :187-189
            } else {
                st->best = st->cur_q;
                st->best_buf = st->cur_buf;
                st->best_size = size;
                st->done = 1;
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
            st->cur_buf = (st->best > 0.0f) ? (1 - st->best_buf) : 0;
            st->epoch = st->epoch + 1;
            st->ticket = 0;
            st->ff_total = 0;
            st->total_bits = 0;
        }
        s_go = go;
    }
    __syncthreads();
    if (!s_go) return;

    const float q = st->cur_q;
    const int tmode = st->table_mode;
    {
        const int t = threadIdx.x;                              // 128 threads = 2 tables x 64
        const int tb = t >> 6, k = t & 63;
        const int nat = c_zigzag[k];
        const int qv = scaled_qval(tb ? c_base_chr[nat] : c_base_lum[nat], q, tmode);
        const unsigned d = (unsigned)qv << 3;                   // jcdctmgr.c: divisor = quantval << 3
        st->qtab[tb][k] = (unsigned char)qv;
        st->recip[tb][k] = (unsigned)((1ull << 32) / d) + 1u;   // umulhi(t, recip) == t / d for t < 2^17
        st->half[tb][k] = (unsigned short)(d >> 1);
    }
    __syncthreads();
    unsigned char *out = st->cur_buf ? out1 : out0;
    if (out == nullptr) return;
    for (int i = threadIdx.x; i < HEADER_BYTES; i += blockDim.x) {
        unsigned char b = c_hdr[i];
        if (i == HDR_OFF_MINOR) b = (unsigned char)st->jfif_minor;
        else if (i >= HDR_OFF_DQT0 && i < HDR_OFF_DQT0 + 64) b = st->qtab[0][i - HDR_OFF_DQT0];
        else if (i >= HDR_OFF_DQT1 && i < HDR_OFF_DQT1 + 64) b = st->qtab[1][i - HDR_OFF_DQT1];
        else if (i == HDR_OFF_H) b = (unsigned char)(h >> 8);
        else if (i == HDR_OFF_H + 1) b = (unsigned char)h;
        else if (i == HDR_OFF_W) b = (unsigned char)(w >> 8);
        else if (i == HDR_OFF_W + 1) b = (unsigned char)w;
        if (i < out_cap) out[i] = b;
    }
}

// Leaves the best stream in out0 (a no-op unless it sits in the other buffer).
__global__ void __launch_bounds__(256) k_publish(const FitState *st, unsigned char *out0, const unsigned char *out1,
                                                 long long out_cap)
{
    if (!(st->best > 0.0f) || st->best_buf == 0) return;
    long long n = st->best_size < out_cap ? st->best_size : out_cap;
    const long long nvec = (n + 15) >> 4;
    uint4 *d = reinterpret_cast<uint4 *>(out0);
    const uint4 *s = reinterpret_cast<const uint4 *>(out1);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
         i += (long long)gridDim.x * blockDim.x)
        d[i] = s[i];
}

}  // namespace icb
