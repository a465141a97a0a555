// api.cu -- host side of libimgcomp_b200.so: the C ABI of include/imgcomp_b200.h,
// the per-GPU slot pool (stream work queues that replace the reference's
// executor, CompressionBatch.java:64-108) and the per-image orchestration of
// core/ImageCompressionJpg.java:77-122 / core/ImageCompressionPng.java:37-75 on
// top of the sm_100a kernels.  There is no CPU fallback: without a CUDA device
// every entry point fails with ICB_ERR_NO_DEVICE.
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/imgcomp_b200.h"
#include "common.cuh"
#include "k_control.cuh"
#include "k_entropy.cuh"
#include "k_resize.cuh"
#include "k_transform.cuh"

using namespace icb;

namespace {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);

struct DeviceCtx {
    int dev = 0;
    int sms = 148;
    cudaStream_t timer_stream = nullptr;    // icb_mark: joins every slot stream of this GPU
    cudaEvent_t mark[2] = {nullptr, nullptr};
    cudaEvent_t join = nullptr;
};

enum Phase { PH_IDLE = 0, PH_CACHED, PH_SEARCH, PH_RESIZE };

struct Job {
    Phase phase = PH_IDLE;
    const unsigned char *d_src = nullptr;   // original pixels in HBM
    int w = 0, h = 0, pitch = 0, fmt = 0;
    float init_q = 0;
    int64_t target = 0;
    int has_cached = 0;
    float cached_q = 0;
    double cached_scale = 1.0;
    unsigned char *out = nullptr;
    size_t out_cap = 0;
    double scale = 1.0;                     // scale of the phase in flight
    bool coef_is_scale1 = false;            // coefficients in HBM belong to the unscaled image
    long long dev_cap = 0;                  // capacity handed to the kernels
    int passes = 0;
    icb_result res{};
};

struct Slot {
    int index = 0;
    int devi = 0;                           // index into handle->devs
    cudaStream_t stream = nullptr;
    unsigned char *d_rgb = nullptr;   size_t rgb_cap = 0;
    unsigned char *d_rgb2 = nullptr;  size_t rgb2_cap = 0;
    int16_t *d_coef = nullptr;        size_t coef_cap = 0;      // bytes
    unsigned char *d_pass = nullptr;  size_t pass_cap = 0;      // 2 x entropy pass arena (ping-pong)
    uint32_t *d_ent = nullptr;        size_t ent_cap = 0;       // survivor entries + per-tile counts
    unsigned char *d_out = nullptr;   size_t out_cap = 0;       // final JFIF stream
    FitState *d_state = nullptr;
    FitState *h_state = nullptr;            // pinned
    Geom geom{};
    bool busy = false;
    uint64_t gen = 0;
    Job job;
};

}  // namespace

struct icb_image {
    int devi;
    unsigned char *d_px;
    int w, h, pitch, fmt;
};

struct icb_handle {
    std::vector<DeviceCtx> devs;
    std::vector<Slot> slots;
    std::mutex mu;
    std::condition_variable cv;
    int rr = 0;
    int table_mode = 0, jfif_minor = 2;
    PFN_encodeTiled encode_tiled = nullptr;
    std::string last_error;
    std::atomic<uint64_t> images_done{0}, kernels{0}, passes{0}, h2d{0}, d2h{0};
    float last_xf_ms = 0, last_ent_ms = 0, last_total_ms = 0;
    int last_xf_launches = 0, last_ent_launches = 0;
    // what the first midpoint of earlier searches of the same (w, h, q, target) did: key << 1 | failed
    std::atomic<uint64_t> first_pass_seen[256];
    icb_handle() { for (auto &e : first_pass_seen) e.store(0); }
};

namespace {

int fail(icb_handle *h, int code, const char *what, cudaError_t e = cudaSuccess)
{
    if (h) {
        std::lock_guard<std::mutex> lk(h->mu);
        h->last_error = std::string(what) + (e != cudaSuccess ? std::string(": ") + cudaGetErrorString(e) : "");
    }
    if (e == cudaErrorMemoryAllocation) return ICB_ERR_OOM;
    return code;
}

#define CK(call)                                                              \
    do {                                                                      \
        cudaError_t e_ = (call);                                              \
        if (e_ != cudaSuccess) return fail(h, ICB_ERR_CUDA, #call, e_);       \
    } while (0)

// Launch on `st`; with pdl the kernel may start before its predecessor on the stream has finished
// (programmatic stream serialization) -- it calls griddepcontrol.wait before touching its results.
bool pdl_enabled()
{
    static const bool on = [] { const char *e = getenv("ICB_PDL"); return !(e && e[0] == '0'); }();
    return on;
}

template <typename... KArgs, typename... Args>
cudaError_t launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl, Args &&...args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = (pdl && pdl_enabled()) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

inline int bpp_of(int fmt) { return (fmt == ICB_FMT_RGBA8 || fmt == ICB_FMT_ABGR8) ? 4 : 3; }
inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
// device pitch: 16-byte multiple (TMA global stride) and at least one TMA box wide
inline int dev_pitch(int w, int bpp) { return (int)std::max<size_t>(round_up((size_t)w * bpp, 16), TILE_ROW_BYTES); }

// jchuff.c jpeg_make_c_derived_tbl
void derive(const unsigned char *bits, const unsigned char *vals, unsigned *enc, int n)
{
    for (int i = 0; i < n; i++) enc[i] = 0;
    unsigned code = 0;
    int k = 0;
    for (int l = 1; l <= 16; l++) {
        for (int i = 0; i < bits[l - 1]; i++, k++) enc[vals[k]] = (code++ << 8) | (unsigned)l;
        code <<= 1;
    }
}

// a2.8 header template: SOI, APP0 JFIF, DQT x2, SOF0, DHT x4 (DC0 AC0 DC1 AC1), SOS
void build_header_template(unsigned char *hdr, const unsigned char dc_bits[2][16], const unsigned char ac_bits[2][16],
                           const unsigned char ac_vals[2][162])
{
    int n = 0;
    auto b = [&](int v) { hdr[n++] = (unsigned char)v; };
    auto u16 = [&](int v) { b(v >> 8); b(v & 0xff); };
    u16(0xFFD8);
    u16(0xFFE0); u16(16); b('J'); b('F'); b('I'); b('F'); b(0); b(1); b(2); b(0); u16(1); u16(1); b(0); b(0);
    for (int t = 0; t < 2; t++) { u16(0xFFDB); u16(67); b(t); for (int i = 0; i < 64; i++) b(0); }
    u16(0xFFC0); u16(17); b(8); u16(0); u16(0); b(3);
    b(1); b(0x22); b(0); b(2); b(0x11); b(1); b(3); b(0x11); b(1);
    for (int t = 0; t < 2; t++) {
        u16(0xFFC4); u16(2 + 1 + 16 + 12); b(t);
        for (int i = 0; i < 16; i++) b(dc_bits[t][i]);
        for (int i = 0; i < 12; i++) b(i);
        u16(0xFFC4); u16(2 + 1 + 16 + 162); b(0x10 | t);
        for (int i = 0; i < 16; i++) b(ac_bits[t][i]);
        for (int i = 0; i < 162; i++) b(ac_vals[t][i]);
    }
    u16(0xFFDA); u16(12); b(3); b(1); b(0x00); b(2); b(0x11); b(3); b(0x11); b(0); b(0x3F); b(0);
}

int setup_device(icb_handle *h, DeviceCtx &d)
{
    CK(cudaSetDevice(d.dev));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, d.dev));
    d.sms = prop.multiProcessorCount;
    if (prop.major < 10) return fail(h, ICB_ERR_NO_DEVICE, "device is not sm_100 class");
    unsigned char dc_bits[2][16], ac_bits[2][16], ac_vals[2][162], dc_vals[12];
    CK(cudaMemcpyFromSymbol(dc_bits, c_dc_bits, sizeof(dc_bits)));
    CK(cudaMemcpyFromSymbol(ac_bits, c_ac_bits, sizeof(ac_bits)));
    CK(cudaMemcpyFromSymbol(ac_vals, c_ac_vals, sizeof(ac_vals)));
    for (int i = 0; i < 12; i++) dc_vals[i] = (unsigned char)i;
    unsigned dc_enc[2][16], ac_enc[2][256];
    for (int t = 0; t < 2; t++) {
        derive(dc_bits[t], dc_vals, dc_enc[t], 16);
        derive(ac_bits[t], ac_vals[t], ac_enc[t], 256);
    }
    unsigned huf[HUF_WORDS];
    memcpy(huf, ac_enc, sizeof(ac_enc));
    memcpy(huf + 512, dc_enc, sizeof(dc_enc));
    CK(cudaMemcpyToSymbol(g_huf, huf, sizeof(huf)));
    unsigned char hdr[HEADER_BYTES + 16];
    build_header_template(hdr, dc_bits, ac_bits, ac_vals);
    CK(cudaMemcpyToSymbol(c_hdr, hdr, HEADER_BYTES));
    CK(cudaFuncSetAttribute(k_transform<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, XF_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_transform<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, XF_SMEM_BYTES));
    return ICB_OK;
}

template <typename T>
int ensure(icb_handle *h, T *&p, size_t &cap, size_t need)
{
    if (need <= cap) return ICB_OK;
    if (p) CK(cudaFree(p));
    p = nullptr;
    cap = 0;
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, need);
    if (e != cudaSuccess) return fail(h, ICB_ERR_OOM, "cudaMalloc", e);
    p = (T *)q;
    cap = need;
    return ICB_OK;
}

Slot *acquire_slot(icb_handle *h, int devi = -1)
{
    std::unique_lock<std::mutex> lk(h->mu);
    while (true) {
        const int n = (int)h->slots.size();
        for (int k = 0; k < n; k++) {
            Slot &s = h->slots[(h->rr + k) % n];
            if (!s.busy && (devi < 0 || s.devi == devi)) {
                s.busy = true;
                s.gen++;
                h->rr = (s.index + 1) % n;
                return &s;
            }
        }
        h->cv.wait(lk);
    }
}

void release_slot(icb_handle *h, Slot *s)
{
    {
        std::lock_guard<std::mutex> lk(h->mu);
        s->busy = false;
        s->job = Job();
    }
    h->cv.notify_all();
}

// ---- enqueue helpers (all asynchronous on the slot's stream) ----

int enqueue_transform(icb_handle *h, Slot *s, const unsigned char *d_px, int w, int hgt, int pitch, int fmt)
{
    static const int forced_tile = [] { const char *e = getenv("ICB_ENT_TILE"); return e ? atoi(e) : 0; }();   // tests / A-B runs
    // a fit that has the GPU to itself (no other slot of this device in use) is a latency case: smaller
    // tiles give its passes more warps (25 MP alone: 1.22 ms with 128-block tiles, 1.36 ms with 256);
    // the tile size never changes a byte of the result
    int tile_pick = forced_tile;
    if (!tile_pick) {
        int busy = 0;
        {
            std::lock_guard<std::mutex> lk(h->mu);
            for (const Slot &o : h->slots) busy += (o.busy && o.devi == s->devi) ? 1 : 0;
        }
        if (busy <= 1) tile_pick = std::min(128, ent_tile_blocks(6 * ((w + 15) >> 4) * ((hgt + 15) >> 4), 0));
    }
    const Geom g = make_geom(w, hgt, pitch, fmt == ICB_FMT_BGR8, tile_pick);
    s->geom = g;
    int rc = ensure(h, s->d_coef, s->coef_cap, (size_t)g.nblocks * 128);
    if (rc) return rc;
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)(pitch / 4), (cuuint64_t)hgt};
    const cuuint64_t gstr[1] = {(cuuint64_t)pitch};
    const cuuint32_t box[2] = {TILE_ROW_BYTES / 4, 16};
    const cuuint32_t estr[2] = {1, 1};
    CUresult cr = h->encode_tiled(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, (void *)d_px, gdim, gstr, box, estr,
                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) return fail(h, ICB_ERR_CUDA, "cuTensorMapEncodeTiled failed");
    const int tiles_per_row = (g.mcu_cols + TILE_MCUS - 1) / TILE_MCUS;
    const int ntiles = tiles_per_row * g.mcu_rows;
    static const int xf_cps = [] { const char *e = getenv("ICB_XF_CPS"); return e ? atoi(e) : XF_CTAS; }();   // tuning knob
    const int grid = std::min(ntiles, h->devs[s->devi].sms * std::max(1, std::min(XF_CTAS, xf_cps)));
    if (g.bgr)
        k_transform<true><<<grid, XF_THREADS, XF_SMEM_BYTES, s->stream>>>(tmap, g, s->d_coef, tiles_per_row, ntiles);
    else
        k_transform<false><<<grid, XF_THREADS, XF_SMEM_BYTES, s->stream>>>(tmap, g, s->d_coef, tiles_per_row, ntiles);
    CK(cudaGetLastError());
    h->kernels++;
    return ICB_OK;
}

// Number of bisection iterations Java can start for this initial quality (:167-189): hi - lo halves
// every iteration whatever the path, and the loop stops once it is below 0.01f.  Midpoints can round
// (by an ulp) when init_q is not dyadic, so one more pass is queued -- it finds cur_valid == 0 and
// returns at once -- only if some q / 2^k comes that close to the 0.01f threshold.
int max_search_iters(float init_q)
{
    int n = 0;
    bool near = false;
    float d = init_q;
    while (n < 8 && !(d < 0.01f)) {
        d *= 0.5f;
        n++;
        if (d > 0.0099f && d < 0.0101f) near = true;
    }
    return std::min(8, n + (near ? 1 : 0));
}

// persistent grid of the entropy pass: c CTAs per SM (c <= 4 by registers) such that the statically
// dealt tiles fill the last round of every warp as evenly as possible
int pass_grid(int ntiles, int sms)
{
    const int need = (ntiles + ENT_WARPS - 1) / ENT_WARPS;
    static const int forced = [] { const char *e = getenv("ICB_PASS_CPS"); return e ? atoi(e) : 0; }();
    if (forced > 0) return std::min(need, sms * forced);     // tuning knob: CTAs per SM of the pass
    if (need <= sms * 4) return need;
    int best_grid = sms * 4;
    double best_eff = 0.0;
    for (int c = 4; c >= 2; c--) {
        const long long warps = (long long)sms * c * ENT_WARPS;
        const long long rounds = (ntiles + warps - 1) / warps;
        const double eff = (double)ntiles / (double)(rounds * warps);
        if (eff > best_eff + 0.02) { best_eff = eff; best_grid = sms * c; }
    }
    return best_grid;
}

// Will the first midpoint (q / 2) of this search overshoot the target?  If so every later pass
// looks below it, and the survivor lists are built for q / 2 and pruned as hi walks down; if not,
// they are built for q once.  Only time depends on the answer, never a byte of the result: a wrong
// guess costs one list rebuild.  The guess is what the same (w, h, q, target) did last time --
// batches are runs of like images, the reason the reference keeps its LearnedParams cache
// (CacheTools.java:14-21) -- else a bytes-per-pixel rule of thumb.
uint64_t first_pass_key(int w, int hgt, float q, long long target)
{
    uint32_t qb;
    memcpy(&qb, &q, 4);
    uint64_t k = ((uint64_t)(uint32_t)w << 32 | (uint32_t)hgt) * 0x9E3779B97F4A7C15ull;
    k ^= ((uint64_t)qb << 32 | (uint64_t)(target & 0xffffffffll)) * 0xC2B2AE3D27D4EB4Full;
    k ^= k >> 29;
    return (k | 1ull) & 0x7fffffffffffffffull;
}

bool expect_first_pass_fails(icb_handle *h, int w, int hgt, float q, long long target)
{
    static const int forced = [] { const char *e = getenv("ICB_PRUNE"); return e ? atoi(e) : -1; }();   // tuning knob
    if (forced >= 0) return forced != 0;
    const uint64_t k = first_pass_key(w, hgt, q, target);
    const uint64_t v = h->first_pass_seen[k & 255].load(std::memory_order_relaxed);
    if ((v >> 1) == k) return (v & 1) != 0;
    return 0.4 * (double)(q * 0.5f) * (double)w * (double)hgt > (double)target;
}

void note_first_pass(icb_handle *h, int w, int hgt, float q, long long target, bool failed)
{
    const uint64_t k = first_pass_key(w, hgt, q, target);
    h->first_pass_seen[k & 255].store(k << 1 | (failed ? 1u : 0u), std::memory_order_relaxed);
}

int enqueue_passes(icb_handle *h, Slot *s, int mode, float q, long long target, int npasses, bool emit,
                   long long dev_cap, cudaEvent_t *pass_ev = nullptr, bool expect_fail = false)
{
    const Geom &g = s->geom;
    const int nctas = (g.ntiles + ENT_WARPS - 1) / ENT_WARPS;
    const int nchunks = nctas;                       // chunk = the 8 tiles one CTA takes per round
    const int cap_t = ent_cap(g.tile_blocks), stride_t = ent_scr_stride(g.tile_blocks);
    const size_t per = round_up(ent_pass_bytes(g.ntiles, nchunks, stride_t), 256);
    int rc = ensure(h, s->d_pass, s->pass_cap, 2 * per + 256);
    if (rc) return rc;
    rc = ensure(h, s->d_ent, s->ent_cap, ((size_t)g.ntiles * cap_t + (size_t)g.ntiles) * 4 + 256);
    if (rc) return rc;
    EntParams ep;
    ep.coef = s->d_coef;
    ep.st = s->d_state;
    ep.ent = s->d_ent;
    ep.ent_cnt = s->d_ent + (size_t)g.ntiles * cap_t;
    ep.tile_blocks = g.tile_blocks; ep.ent_cap = cap_t; ep.scr_stride = stride_t;
    ep.nblocks = g.nblocks;
    ep.ntiles = g.ntiles;
    ep.nchunks = nchunks;
    for (int b = 0; b < 2; b++) {
        unsigned char *base = s->d_pass + (size_t)b * per;
        const size_t n = (size_t)g.ntiles, c = (size_t)nchunks;
        ep.buf[b].ff = reinterpret_cast<uint4 *>(base);                 base += n * 16;
        ep.buf[b].crec = reinterpret_cast<uint4 *>(base);               base += c * 48;
        ep.buf[b].Pc = reinterpret_cast<unsigned long long *>(base);    base += c * 8;
        ep.buf[b].scratch = reinterpret_cast<uint32_t *>(base);         base += n * (size_t)stride_t * 4;
        ep.buf[b].L = reinterpret_cast<unsigned int *>(base);           base += n * 4;
        ep.buf[b].ones = reinterpret_cast<unsigned int *>(base);        base += n * 4;
        ep.buf[b].Fc = reinterpret_cast<unsigned int *>(base);          base += c * 4;
        ep.buf[b].Fb = reinterpret_cast<unsigned int *>(base);
    }
    // searches expected to overshoot at their first midpoint walk hi down from there: the first
    // lists are built for that midpoint only (rebuilt for hi if the pass fits after all) and pruned
    // between passes as hi drops
    const bool prune = mode == MODE_SEARCH && expect_fail && npasses > 1;
    BeginArgs ba;
    ba.mode = mode; ba.q = q; ba.target = target; ba.table_mode = h->table_mode; ba.jfif_minor = h->jfif_minor;
    ba.list_q = prune ? (0.0f + q) / 2.0f : q;                   // = the first midpoint (:168)
    ba.pad = 0;
    const int pneed = (g.ntiles + PREP_WARPS - 1) / PREP_WARPS;
    const int pwave = h->devs[s->devi].sms * PREP_CTAS_PER_SM;
    const int prounds = (pneed + pwave - 1) / pwave;             // tiles a warp takes
    const dim3 prep_grid((pneed + prounds - 1) / prounds);
    CK(launch(k_entropy_prepare<false>, prep_grid, dim3(PREP_THREADS), 0, s->stream, true, ep, ba));
    h->kernels += 1;
    const int pgrid = pass_grid(g.ntiles, h->devs[s->devi].sms);
    const bool pdl = pass_ev == nullptr;
    for (int i = 0; i < npasses; i++) {
        if (pass_ev) cudaEventRecord(pass_ev[2 * i], s->stream);
        CK(launch(k_entropy_pass, dim3(pgrid), dim3(ENT_THREADS), 0, s->stream, pdl, ep));
        if (pass_ev) cudaEventRecord(pass_ev[2 * i + 1], s->stream);
        h->kernels += 1;
        if (prune && i + 1 < npasses) {
            if (i == 0) {
                CK(launch(k_entropy_prepare<true>, prep_grid, dim3(PREP_THREADS), 0, s->stream, pdl, ep, ba));
                h->kernels += 1;
            }
            CK(launch(k_entropy_prune, dim3(nctas), dim3(ENT_THREADS), 0, s->stream, pdl, ep, i + 1));
            h->kernels += 1;
        }
    }
    if (emit) {
        CK(launch(k_entropy_emit, dim3(nctas), dim3(ENT_THREADS), 0, s->stream, pass_ev == nullptr, ep, s->d_out,
                  dev_cap, g.w, g.h));
        h->kernels += 1;
    }
    CK(cudaGetLastError());
    return ICB_OK;
}

void resize_consts(int w, int hgt, int dw, int dh, ResizeParams &rp)
{
    // DrawImage.scaleImage builds scale(dw/w, dh/h); renderImageXform inverts it;
    // TransformHelper.c converts with DblToLong = (jlong)(d * 2^32)   [JDK-mem]
    const double m00 = (double)dw / w, m11 = (double)dh / hgt;
    const double ixx = 1.0 / m00, iyy = 1.0 / m11;
    rp.dxl = (long long)(ixx * 4294967296.0);
    rp.dyl = (long long)(iyy * 4294967296.0);
    rp.xbase = (long long)((0.5 * ixx) * 4294967296.0);
    rp.ybase = (long long)((0.5 * iyy) * 4294967296.0);
}

int enqueue_resize(icb_handle *h, Slot *s, const unsigned char *d_src, int w, int hgt, int pitch, int fmt,
                   unsigned char *d_dst, int dw, int dh, int dpitch)
{
    ResizeParams rp;
    rp.src = d_src; rp.dst = d_dst;
    rp.w = w; rp.h = hgt; rp.pitch = pitch;
    rp.dw = dw; rp.dh = dh; rp.dpitch = dpitch;
    resize_consts(w, hgt, dw, dh, rp);
    dim3 grid((dw + 31) / 32, (dh + 7) / 8);
    if (fmt == ICB_FMT_RGBA8) k_resize<4, 3><<<grid, 256, 0, s->stream>>>(rp);
    else if (fmt == ICB_FMT_ABGR8) k_resize<4, 0><<<grid, 256, 0, s->stream>>>(rp);
    else k_resize<3, -1><<<grid, 256, 0, s->stream>>>(rp);
    CK(cudaGetLastError());
    h->kernels++;
    return ICB_OK;
}

int upload(icb_handle *h, cudaStream_t st, unsigned char *d, int dpitch, const unsigned char *px, int w, int hgt,
           int pitch, int bpp)
{
    if (pitch == dpitch)
        CK(cudaMemcpyAsync(d, px, (size_t)(hgt - 1) * pitch + (size_t)w * bpp, cudaMemcpyHostToDevice, st));
    else
        CK(cudaMemcpy2DAsync(d, dpitch, px, pitch, (size_t)w * bpp, hgt, cudaMemcpyHostToDevice, st));
    h->h2d += (uint64_t)w * bpp * hgt;
    return ICB_OK;
}

}  // namespace

// ======================================================================
//  JPG fit state machine (core/ImageCompressionJpg.java:77-122)
// ======================================================================
namespace {

long long worst_stream_bytes(const Geom &g)
{
    return (long long)HEADER_BYTES + 2 + (long long)g.nblocks * 416;     // every byte stuffed
}

int fit_enqueue_phase(icb_handle *h, Slot *s)
{
    Job &j = s->job;
    CK(cudaSetDevice(h->devs[s->devi].dev));
    const unsigned char *src = j.d_src;
    int w = j.w, hgt = j.h, pitch = j.pitch;
    const double scale = j.phase == PH_CACHED ? j.cached_scale : j.scale;
    bool need_xf = true;
    if (scale < 1.0) {                                   // :97-102 / :221-224, always from the original
        int nw, nh;
        icb_resize_dims(j.w, j.h, scale, &nw, &nh);
        const int np = dev_pitch(nw, 3);
        int rc = ensure(h, s->d_rgb2, s->rgb2_cap, (size_t)np * nh + 16);
        if (rc) return rc;
        rc = enqueue_resize(h, s, src, j.w, j.h, j.pitch, j.fmt, s->d_rgb2, nw, nh, np);
        if (rc) return rc;
        src = s->d_rgb2; w = nw; hgt = nh; pitch = np;
        j.coef_is_scale1 = false;
    } else if (j.coef_is_scale1) {
        need_xf = false;                                 // cached miss at scale 1: coefficients still valid
    }
    if (need_xf) {
        int rc = enqueue_transform(h, s, src, w, hgt, pitch, j.fmt);
        if (rc) return rc;
        if (!(scale < 1.0)) j.coef_is_scale1 = true;
    }
    long long cap = std::min<long long>(j.target, worst_stream_bytes(s->geom));
    if (cap < 0) cap = 0;
    j.dev_cap = cap;
    {
        int rc = ensure(h, s->d_out, s->out_cap, round_up((size_t)cap, 16) + 32);
        if (rc) return rc;
    }
    int rc;
    if (j.phase == PH_CACHED)
        rc = enqueue_passes(h, s, MODE_SINGLE, j.cached_q, j.target, 1, true, cap);
    else
        rc = enqueue_passes(h, s, MODE_SEARCH, j.init_q, j.target, max_search_iters(j.init_q), true, cap, nullptr,
                            scale < 1.0 || expect_first_pass_fails(h, j.w, j.h, j.init_q, j.target));
    if (rc) return rc;
    CK(cudaMemcpyAsync(s->h_state, s->d_state, sizeof(FitState), cudaMemcpyDeviceToHost, s->stream));
    const size_t ncopy = (size_t)std::min<long long>((long long)j.out_cap, cap);
    if (ncopy && j.out) {
        CK(cudaMemcpyAsync(j.out, s->d_out, ncopy, cudaMemcpyDeviceToHost, s->stream));
        h->d2h += ncopy;
    }
    h->d2h += sizeof(FitState);
    return ICB_OK;
}

// after the stream drained: 0 = job finished (j.res filled), 1 = another phase was queued, <0 error
int fit_advance(icb_handle *h, Slot *s)
{
    Job &j = s->job;
    const FitState &st = *s->h_state;
    j.passes += st.passes;
    h->passes += (uint64_t)st.passes;
    j.res.n_passes = j.passes;
    j.res.device = h->devs[s->devi].dev;
    bool ok = false;
    if (j.phase == PH_CACHED) {
        if (st.best_size <= j.target) {                 // :228-231
            ok = true;
            j.res.best_q = j.cached_q;
            j.res.scale = j.cached_scale;
            j.res.used_cache = 1;
        } else {                                         // :86-88 fall back to the standard flow
            j.phase = PH_SEARCH;
            j.scale = 1.0;
        }
    } else {
        if (j.scale == 1.0) note_first_pass(h, j.w, j.h, j.init_q, j.target, !(st.best >= (0.0f + j.init_q) / 2.0f));
        if (st.best > 0.0f) {                            // :109-113
            ok = true;
            j.res.best_q = st.best;
            j.res.scale = j.scale;
            j.res.used_cache = 0;
        } else {                                         // :96 next scale
            j.scale = (j.scale == 1.0) ? 0.85 : j.scale * 0.85;
            if (!(j.scale > 0.1)) {
                j.res.status = ICB_NO_FIT;               // :120-121
                j.res.out_len = 0;
                j.res.best_q = -1.0f;
                j.res.scale = 0.0;
                return 0;
            }
        }
    }
    if (ok) {
        j.res.out_len = (size_t)st.best_size;
        j.res.status = (size_t)st.best_size <= j.out_cap ? ICB_OK : ICB_ERR_CAPACITY;
        return 0;
    }
    int rc = fit_enqueue_phase(h, s);
    return rc ? rc : 1;
}

int start_fit(icb_handle *h, Slot *s, const unsigned char *d_src, int w, int hgt, int pitch, int fmt, float init_q,
              int64_t target, int has_cached, float cached_q, double cached_scale, unsigned char *out, size_t out_cap)
{
    Job &j = s->job;
    j = Job();
    j.d_src = d_src; j.w = w; j.h = hgt; j.pitch = pitch; j.fmt = fmt;
    j.init_q = init_q; j.target = target;
    j.has_cached = has_cached; j.cached_q = cached_q; j.cached_scale = cached_scale;
    j.out = out; j.out_cap = out ? out_cap : 0;
    j.phase = has_cached ? PH_CACHED : PH_SEARCH;
    j.scale = 1.0;
    return fit_enqueue_phase(h, s);
}

int check_jpg_args(icb_handle *h, const void *px, int w, int hgt, int pitch, int fmt)
{
    if (!h) return ICB_ERR_INVALID;
    if (!px || w < 1 || hgt < 1 || w > 65535 || hgt > 65535) return fail(h, ICB_ERR_INVALID, "bad image arguments");
    if (fmt != ICB_FMT_RGB8 && fmt != ICB_FMT_BGR8)
        return fail(h, ICB_ERR_UNSUPPORTED, "JPG path takes RGB8/BGR8 (gray, CMYK, alpha: use the stock writer)");
    if (pitch < w * 3) return fail(h, ICB_ERR_INVALID, "pitch smaller than a row");
    return ICB_OK;
}

int64_t make_ticket(const Slot *s) { return (int64_t)((s->gen << 16) | (uint64_t)s->index); }

}  // namespace

// ======================================================================
//  C ABI
// ======================================================================
extern "C" {

int icb_init(const int *devices, int n_devices, int slots_per_gpu, int max_w, int max_h, icb_handle **out_handle)
{
    if (!out_handle) return ICB_ERR_INVALID;
    *out_handle = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count < 1) return ICB_ERR_NO_DEVICE;
    icb_handle *h = new icb_handle();
    std::vector<int> devs;
    if (devices && n_devices > 0) {
        for (int i = 0; i < n_devices; i++) {
            if (devices[i] < 0 || devices[i] >= count) { delete h; return ICB_ERR_INVALID; }
            devs.push_back(devices[i]);
        }
    } else {
        for (int i = 0; i < count; i++) devs.push_back(i);
    }
    if (slots_per_gpu < 1) slots_per_gpu = 4;
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) {
        delete h;
        return ICB_ERR_CUDA;
    }
    h->encode_tiled = (PFN_encodeTiled)fn;
    for (int d : devs) {
        DeviceCtx dc;
        dc.dev = d;
        int rc = setup_device(h, dc);
        if (rc) { fprintf(stderr, "imgcomp_b200: %s\n", h->last_error.c_str()); delete h; return rc; }
        h->devs.push_back(dc);
    }
    h->slots.resize(devs.size() * (size_t)slots_per_gpu);
    // interleave devices so round-robin acquisition spreads images over GPUs
    for (size_t i = 0; i < h->slots.size(); i++) {
        Slot &s = h->slots[i];
        s.index = (int)i;
        s.devi = (int)(i % devs.size());
        if (cudaSetDevice(h->devs[s.devi].dev) != cudaSuccess ||
            cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaMalloc((void **)&s.d_state, sizeof(FitState)) != cudaSuccess ||
            cudaMallocHost((void **)&s.h_state, sizeof(FitState)) != cudaSuccess) {
            icb_shutdown(h);
            return ICB_ERR_OOM;
        }
        if (max_w > 0 && max_h > 0) {
            const Geom g = make_geom(max_w, max_h, dev_pitch(max_w, 3), 0);
            if (ensure(h, s.d_rgb, s.rgb_cap, (size_t)g.pitch * max_h + 16) ||
                ensure(h, s.d_coef, s.coef_cap, (size_t)g.nblocks * 128)) {
                icb_shutdown(h);
                return ICB_ERR_OOM;
            }
        }
    }
    *out_handle = h;
    return ICB_OK;
}

void icb_shutdown(icb_handle *h)
{
    if (!h) return;
#ifdef ICB_TIMING
    if (getenv("ICB_DUMP_TIMELINE")) {          // the last job of every slot, on the GPU's global timer
        unsigned long long t0 = ~0ull;
        for (Slot &s : h->slots) if (s.h_state && s.h_state->dbg_k[0] && s.h_state->dbg_k[0] < t0) t0 = s.h_state->dbg_k[0];
        for (Slot &s : h->slots) {
            if (!s.h_state || !s.h_state->dbg_k[0]) continue;
            const FitState &fs = *s.h_state;
            fprintf(stderr, "slot %d: prepare@%.1f", s.index, (fs.dbg_k[0] - t0) * 1e-3);
            for (int i = 0; i < fs.passes && i < 8; i++)
                fprintf(stderr, " | pass%d %.1f..%.1f", i, (fs.dbg_t[i][0] - t0) * 1e-3, (fs.dbg_t[i][3] - t0) * 1e-3);
            fprintf(stderr, " | emit %.1f..%.1f\n", (fs.dbg_k[1] - t0) * 1e-3, (fs.dbg_k[2] - t0) * 1e-3);
        }
    }
#endif
    for (Slot &s : h->slots) {
        if (s.devi < (int)h->devs.size()) cudaSetDevice(h->devs[s.devi].dev);
        if (s.stream) { cudaStreamSynchronize(s.stream); cudaStreamDestroy(s.stream); }
        cudaFree(s.d_rgb); cudaFree(s.d_rgb2); cudaFree(s.d_coef); cudaFree(s.d_pass);
        cudaFree(s.d_out); cudaFree(s.d_state); cudaFree(s.d_ent);
        if (s.h_state) cudaFreeHost(s.h_state);
    }
    delete h;
}

int icb_device_count(const icb_handle *h) { return h ? (int)h->devs.size() : 0; }

const char *icb_last_error(const icb_handle *h) { return h ? h->last_error.c_str() : "no handle"; }

int icb_set_profile(icb_handle *h, int table_mode, int jfif_minor)
{
    if (!h || table_mode < 0 || table_mode > 1 || jfif_minor < 0 || jfif_minor > 255) return ICB_ERR_INVALID;
    h->table_mode = table_mode;
    h->jfif_minor = jfif_minor;
    return ICB_OK;
}

void *icb_alloc_pinned(size_t bytes)
{
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) return nullptr;
    return p;
}

void icb_free_pinned(void *p) { if (p) cudaFreeHost(p); }

void icb_resize_dims(int w, int hgt, double scale, int *dw, int *dh)
{
    const int a = (int)(w * scale), b = (int)(hgt * scale);       // tools/ImageTools.java:8-9
    if (dw) *dw = a > 1 ? a : 1;
    if (dh) *dh = b > 1 ? b : 1;
}

int icb_submit_jpg_fit(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt, float init_q,
                       int64_t target, int has_cached, float cached_q, double cached_scale, uint8_t *out,
                       size_t out_cap, int64_t *ticket)
{
    int rc = check_jpg_args(h, px, w, hgt, pitch, fmt);
    if (rc) return rc;
    if (!ticket) return ICB_ERR_INVALID;
    Slot *s = acquire_slot(h);
    CK(cudaSetDevice(h->devs[s->devi].dev));
    const int dp = dev_pitch(w, 3);
    rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)dp * hgt + 16);
    if (!rc) rc = upload(h, s->stream, s->d_rgb, dp, px, w, hgt, pitch, 3);
    if (!rc) rc = start_fit(h, s, s->d_rgb, w, hgt, dp, fmt, init_q, target, has_cached, cached_q, cached_scale, out,
                            out_cap);
    if (rc) { cudaStreamSynchronize(s->stream); release_slot(h, s); return rc; }
    *ticket = make_ticket(s);
    return ICB_OK;
}

int icb_submit_jpg_fit_image(icb_handle *h, const icb_image *img, float init_q, int64_t target, int has_cached,
                             float cached_q, double cached_scale, uint8_t *out, size_t out_cap, int64_t *ticket)
{
    if (!h || !img || !ticket) return ICB_ERR_INVALID;
    if (img->fmt != ICB_FMT_RGB8 && img->fmt != ICB_FMT_BGR8) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    Slot *s = acquire_slot(h, img->devi);
    int rc = start_fit(h, s, img->d_px, img->w, img->h, img->pitch, img->fmt, init_q, target, has_cached, cached_q,
                       cached_scale, out, out_cap);
    if (rc) { cudaStreamSynchronize(s->stream); release_slot(h, s); return rc; }
    *ticket = make_ticket(s);
    return ICB_OK;
}

int icb_submit_resize_image(icb_handle *h, const icb_image *img, int dw, int dh, uint8_t *dst, int dpitch,
                            int64_t *ticket)
{
    if (!h || !img || !ticket || dw < 1 || dh < 1) return ICB_ERR_INVALID;
    const int bpp = bpp_of(img->fmt);
    if (dst && dpitch < dw * bpp) return fail(h, ICB_ERR_INVALID, "dpitch smaller than a row");
    Slot *s = acquire_slot(h, img->devi);
    CK(cudaSetDevice(h->devs[s->devi].dev));
    const int np = (int)round_up((size_t)dw * bpp, 16);
    int rc = ensure(h, s->d_rgb2, s->rgb2_cap, (size_t)np * dh + 16);
    if (!rc) rc = enqueue_resize(h, s, img->d_px, img->w, img->h, img->pitch, img->fmt, s->d_rgb2, dw, dh, np);
    if (!rc && dst) {                            // dst == NULL: the resized pixels stay in HBM (kernel-only timing)
        cudaError_t e = cudaMemcpy2DAsync(dst, dpitch, s->d_rgb2, np, (size_t)dw * bpp, dh, cudaMemcpyDeviceToHost,
                                          s->stream);
        if (e != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "D2H", e);
        h->d2h += (uint64_t)dw * bpp * dh;
    }
    if (rc) { cudaStreamSynchronize(s->stream); release_slot(h, s); return rc; }
    s->job = Job();
    s->job.phase = PH_RESIZE;
    *ticket = make_ticket(s);
    return ICB_OK;
}

int icb_wait(icb_handle *h, int64_t ticket, icb_result *res)
{
    if (!h) return ICB_ERR_INVALID;
    const size_t idx = (size_t)(ticket & 0xffff);
    if (idx >= h->slots.size()) return ICB_ERR_TICKET;
    Slot *s = &h->slots[idx];
    {
        std::lock_guard<std::mutex> lk(h->mu);
        if (!s->busy || (int64_t)((s->gen << 16) | (uint64_t)s->index) != ticket) return ICB_ERR_TICKET;
    }
    int status = ICB_OK;
    cudaSetDevice(h->devs[s->devi].dev);
    while (true) {
        cudaError_t e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) { status = fail(h, ICB_ERR_CUDA, "stream", e); break; }
        if (s->job.phase == PH_RESIZE) { s->job.res.status = ICB_OK; break; }
        int rc = fit_advance(h, s);
        if (rc == 0) { status = s->job.res.status; break; }
        if (rc < 0) { status = rc; cudaStreamSynchronize(s->stream); break; }
    }
    s->job.res.status = status;
    if (res) *res = s->job.res;
    h->images_done++;
    release_slot(h, s);
    return status;
}

int icb_jpg_fit(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt, float init_q, int64_t target,
                int has_cached, float cached_q, double cached_scale, uint8_t *out, size_t out_cap, size_t *out_len,
                float *best_q, double *scale, int *used_cache)
{
    int64_t t = 0;
    int rc = icb_submit_jpg_fit(h, px, w, hgt, pitch, fmt, init_q, target, has_cached, cached_q, cached_scale, out,
                                out_cap, &t);
    if (rc) return rc;
    icb_result r;
    rc = icb_wait(h, t, &r);
    if (out_len) *out_len = r.out_len;
    if (best_q) *best_q = r.best_q;
    if (scale) *scale = r.scale;
    if (used_cache) *used_cache = r.used_cache;
    return rc;
}

}  // extern "C"

// ======================================================================
//  C ABI, part 2: single encodes, hooks, resize, resident images, stats
// ======================================================================
namespace {

// upload + transform on a freshly acquired slot (synchronous helpers below use it)
int stage_and_transform(icb_handle *h, Slot *s, const uint8_t *px, int w, int hgt, int pitch, int fmt)
{
    CK(cudaSetDevice(h->devs[s->devi].dev));
    const int dp = dev_pitch(w, 3);
    int rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)dp * hgt + 16);
    if (!rc) rc = upload(h, s->stream, s->d_rgb, dp, px, w, hgt, pitch, 3);
    if (!rc) rc = enqueue_transform(h, s, s->d_rgb, w, hgt, dp, fmt);
    return rc;
}

int ensure_out(icb_handle *h, Slot *s, long long cap)
{
    return ensure(h, s->d_out, s->out_cap, round_up((size_t)cap, 16) + 32);
}

}  // namespace

extern "C" {

int icb_jpg_encode(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt, float q, uint8_t *out,
                   size_t out_cap, size_t *out_len)
{
    int rc = check_jpg_args(h, px, w, hgt, pitch, fmt);
    if (rc) return rc;
    if (!out_len) return ICB_ERR_INVALID;
    Slot *s = acquire_slot(h);
    rc = stage_and_transform(h, s, px, w, hgt, pitch, fmt);
    long long cap = 0;
    if (!rc) {
        cap = std::min<long long>(out ? (long long)out_cap : 0, worst_stream_bytes(s->geom));
        rc = ensure_out(h, s, cap);
    }
    if (!rc) rc = enqueue_passes(h, s, MODE_SINGLE, q, 0, 1, true, cap);
    if (!rc) {
        cudaError_t e = cudaMemcpyAsync(s->h_state, s->d_state, sizeof(FitState), cudaMemcpyDeviceToHost, s->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "encode", e);
    }
    if (!rc) {
        const long long size = s->h_state->best_size;
        *out_len = (size_t)size;
        h->passes += (uint64_t)s->h_state->passes;
        if (size > cap) {
            rc = ICB_ERR_CAPACITY;
        } else {
            cudaError_t e = cudaMemcpyAsync(out, s->d_out, (size_t)size, cudaMemcpyDeviceToHost, s->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
            if (e != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "encode D2H", e);
            h->d2h += (uint64_t)size;
        }
    } else {
        cudaStreamSynchronize(s->stream);
    }
    h->images_done++;
    release_slot(h, s);
    return rc;
}

int icb_jpg_sizes(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt, const float *qs, int n,
                  int64_t *sizes)
{
    int rc = check_jpg_args(h, px, w, hgt, pitch, fmt);
    if (rc) return rc;
    if (!qs || !sizes || n < 0) return ICB_ERR_INVALID;
    Slot *s = acquire_slot(h);
    rc = stage_and_transform(h, s, px, w, hgt, pitch, fmt);
    for (int i = 0; i < n && !rc; i++) {
        rc = enqueue_passes(h, s, MODE_SINGLE, qs[i], 0, 1, false, 0);
        if (rc) break;
        cudaError_t e = cudaMemcpyAsync(s->h_state, s->d_state, sizeof(FitState), cudaMemcpyDeviceToHost, s->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) { rc = fail(h, ICB_ERR_CUDA, "sizes", e); break; }
        sizes[i] = (int64_t)s->h_state->best_size;
        h->passes += (uint64_t)s->h_state->passes;
    }
    if (rc) cudaStreamSynchronize(s->stream);
    release_slot(h, s);
    return rc;
}

int icb_jpg_coefficients(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt, int16_t *out,
                         size_t out_elems)
{
    int rc = check_jpg_args(h, px, w, hgt, pitch, fmt);
    if (rc) return rc;
    if (!out) return ICB_ERR_INVALID;
    Slot *s = acquire_slot(h);
    rc = stage_and_transform(h, s, px, w, hgt, pitch, fmt);
    if (!rc) {
        const size_t need = (size_t)s->geom.nblocks * 64;
        if (out_elems < need) {
            rc = fail(h, ICB_ERR_CAPACITY, "coefficient buffer too small");
        } else {
            cudaError_t e = cudaMemcpyAsync(out, s->d_coef, need * 2, cudaMemcpyDeviceToHost, s->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
            if (e != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "coefficients", e);
        }
    }
    if (rc) cudaStreamSynchronize(s->stream);
    release_slot(h, s);
    return rc;
}

int icb_resize_bilinear(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt, uint8_t *dst, int dw,
                        int dh, int dpitch)
{
    if (!h) return ICB_ERR_INVALID;
    if (!src || !dst || w < 1 || hgt < 1 || dw < 1 || dh < 1) return fail(h, ICB_ERR_INVALID, "bad resize arguments");
    if (fmt < ICB_FMT_RGB8 || fmt > ICB_FMT_ABGR8) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    const int bpp = bpp_of(fmt);
    if (pitch < w * bpp || dpitch < dw * bpp) return fail(h, ICB_ERR_INVALID, "pitch smaller than a row");
    Slot *s = acquire_slot(h);
    int rc = ICB_OK;
    do {
        if (cudaSetDevice(h->devs[s->devi].dev) != cudaSuccess) { rc = ICB_ERR_CUDA; break; }
        const int sp = (int)round_up((size_t)w * bpp, 16), np = (int)round_up((size_t)dw * bpp, 16);
        rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)sp * hgt + 16);
        if (rc) break;
        rc = ensure(h, s->d_rgb2, s->rgb2_cap, (size_t)np * dh + 16);
        if (rc) break;
        rc = upload(h, s->stream, s->d_rgb, sp, src, w, hgt, pitch, bpp);
        if (rc) break;
        rc = enqueue_resize(h, s, s->d_rgb, w, hgt, sp, fmt, s->d_rgb2, dw, dh, np);
        if (rc) break;
        cudaError_t e = cudaMemcpy2DAsync(dst, dpitch, s->d_rgb2, np, (size_t)dw * bpp, dh, cudaMemcpyDeviceToHost,
                                          s->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "resize", e);
        h->d2h += (uint64_t)dw * bpp * dh;
    } while (0);
    if (rc) cudaStreamSynchronize(s->stream);
    h->images_done++;
    release_slot(h, s);
    return rc;
}

int icb_subsample_factor(int w, int hgt)
{
    const int max_dim = w > hgt ? w : hgt;                          // core/ImageCompression.java:142
    if (max_dim <= 4096) return 1;                                  // :143-144 PREFERRED_MAX_DIM
    const int f = (int)((double)max_dim / 4096);                    // :145 (int) Math.floor
    int s = 1;
    while (s * 2 <= f) s *= 2;                                      // :151 Integer.highestOneBit
    return s;
}

int icb_subsample(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt, int sfac, uint8_t *dst,
                  int dpitch)
{
    if (!h) return ICB_ERR_INVALID;
    if (!src || !dst || w < 1 || hgt < 1 || sfac < 1) return fail(h, ICB_ERR_INVALID, "bad subsample arguments");
    if (fmt < ICB_FMT_RGB8 || fmt > ICB_FMT_ABGR8) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    const int bpp = bpp_of(fmt);
    const int dw = (w + sfac - 1) / sfac, dh = (hgt + sfac - 1) / sfac;
    if (pitch < w * bpp || dpitch < dw * bpp) return fail(h, ICB_ERR_INVALID, "pitch smaller than a row");
    Slot *s = acquire_slot(h);
    int rc = ICB_OK;
    do {
        if (cudaSetDevice(h->devs[s->devi].dev) != cudaSuccess) { rc = ICB_ERR_CUDA; break; }
        const int sp = (int)round_up((size_t)w * bpp, 16), np = (int)round_up((size_t)dw * bpp, 16);
        rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)sp * hgt + 16);
        if (rc) break;
        rc = ensure(h, s->d_rgb2, s->rgb2_cap, (size_t)np * dh + 16);
        if (rc) break;
        rc = upload(h, s->stream, s->d_rgb, sp, src, w, hgt, pitch, bpp);
        if (rc) break;
        dim3 grid((dw + 31) / 32, (dh + 7) / 8);
        if (bpp == 4) k_subsample<4><<<grid, 256, 0, s->stream>>>(s->d_rgb, sp, sfac, s->d_rgb2, dw, dh, np);
        else k_subsample<3><<<grid, 256, 0, s->stream>>>(s->d_rgb, sp, sfac, s->d_rgb2, dw, dh, np);
        h->kernels++;
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess)
            e = cudaMemcpy2DAsync(dst, dpitch, s->d_rgb2, np, (size_t)dw * bpp, dh, cudaMemcpyDeviceToHost, s->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "subsample", e);
        h->d2h += (uint64_t)dw * bpp * dh;
    } while (0);
    if (rc) cudaStreamSynchronize(s->stream);
    h->images_done++;
    release_slot(h, s);
    return rc;
}

int icb_png_target_dims(int w, int hgt, int min_w, int min_h, int *dw, int *dh, double *scale)
{
    if (w < 1 || hgt < 1) return ICB_ERR_INVALID;
    if (w <= min_w && hgt <= min_h) return ICB_NO_FIT;              // :49-53 returns false
    const double wr = (double)min_w / w, hr = (double)min_h / hgt;  // :57-58
    const double sc = wr < hr ? wr : hr;                             // :61 Math.min
    if (scale) *scale = sc;
    icb_resize_dims(w, hgt, sc, dw, dh);                             // ImageTools.java:8-9
    return ICB_OK;
}

int icb_png_fit(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt, int min_w, int min_h,
                uint8_t *dst, size_t dst_cap, int *dw, int *dh)
{
    if (!h) return ICB_ERR_INVALID;
    if (!src || !dw || !dh || w < 1 || hgt < 1) return fail(h, ICB_ERR_INVALID, "null argument");   // :39-41 NPE
    const int rc0 = icb_png_target_dims(w, hgt, min_w, min_h, dw, dh, nullptr);
    if (rc0) return rc0;
    if (!dst) return ICB_OK;                                         // dimension query
    const int bpp = bpp_of(fmt);
    if (dst_cap < (size_t)*dw * *dh * bpp) return fail(h, ICB_ERR_CAPACITY, "dst too small");
    return icb_resize_bilinear(h, src, w, hgt, pitch, fmt, dst, *dw, *dh, *dw * bpp);
}

int icb_image_upload(icb_handle *h, int device_index, const uint8_t *px, int w, int hgt, int pitch, int fmt,
                     icb_image **img)
{
    if (!h || !px || !img || w < 1 || hgt < 1) return ICB_ERR_INVALID;
    if (device_index < 0 || device_index >= (int)h->devs.size()) return fail(h, ICB_ERR_INVALID, "device index");
    if (fmt < ICB_FMT_RGB8 || fmt > ICB_FMT_ABGR8) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    const int bpp = bpp_of(fmt);
    if (pitch < w * bpp) return fail(h, ICB_ERR_INVALID, "pitch smaller than a row");
    CK(cudaSetDevice(h->devs[device_index].dev));
    icb_image *im = new icb_image();
    im->devi = device_index;
    im->w = w; im->h = hgt; im->fmt = fmt;
    im->pitch = dev_pitch(w, bpp);
    cudaError_t e = cudaMalloc((void **)&im->d_px, (size_t)im->pitch * hgt + 16);
    if (e == cudaSuccess) e = cudaMemcpy2D(im->d_px, im->pitch, px, pitch, (size_t)w * bpp, hgt, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        if (im->d_px) cudaFree(im->d_px);
        delete im;
        return fail(h, ICB_ERR_CUDA, "image upload", e);
    }
    *img = im;
    return ICB_OK;
}

void icb_image_free(icb_handle *h, icb_image *img)
{
    if (!img) return;
    if (h && img->devi < (int)h->devs.size()) cudaSetDevice(h->devs[img->devi].dev);
    cudaFree(img->d_px);
    delete img;
}

int icb_stats(icb_handle *h, icb_stats_t *out)
{
    if (!h || !out) return ICB_ERR_INVALID;
    out->images_done = h->images_done.load();
    out->kernels_launched = h->kernels.load();
    out->entropy_passes = h->passes.load();
    out->h2d_bytes = h->h2d.load();
    out->d2h_bytes = h->d2h.load();
    out->last_transform_ms = h->last_xf_ms;
    out->last_entropy_ms = h->last_ent_ms;
    out->last_total_ms = h->last_total_ms;
    out->last_transform_launches = h->last_xf_launches;
    out->last_entropy_launches = h->last_ent_launches;
    return ICB_OK;
}

int icb_profile_transform_series(icb_handle *h, const icb_image *const *imgs, int n, int reps, float *ms_per_launch,
                                 int *launches)
{
    if (!h || !imgs || n < 1 || reps < 1) return ICB_ERR_INVALID;
    for (int i = 0; i < n; i++) {
        if (!imgs[i] || imgs[i]->devi != imgs[0]->devi) return ICB_ERR_INVALID;
        if (imgs[i]->fmt != ICB_FMT_RGB8 && imgs[i]->fmt != ICB_FMT_BGR8) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    }
    Slot *s = acquire_slot(h, imgs[0]->devi);
    int rc = ICB_OK;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    do {
        if (cudaSetDevice(h->devs[s->devi].dev) != cudaSuccess || cudaEventCreate(&e0) != cudaSuccess ||
            cudaEventCreate(&e1) != cudaSuccess) { rc = ICB_ERR_CUDA; break; }
        for (int r = -1; r < reps && !rc; r++) {            // round -1 is the untimed warm-up
            if (r == 0) cudaEventRecord(e0, s->stream);
            for (int i = 0; i < n && !rc; i++)
                rc = enqueue_transform(h, s, imgs[i]->d_px, imgs[i]->w, imgs[i]->h, imgs[i]->pitch, imgs[i]->fmt);
        }
        if (rc) break;
        cudaEventRecord(e1, s->stream);
        cudaError_t e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) { rc = fail(h, ICB_ERR_CUDA, "profile", e); break; }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms_per_launch) *ms_per_launch = ms / (float)(n * reps);
        if (launches) *launches = n * reps;
    } while (0);
    if (rc) cudaStreamSynchronize(s->stream);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    release_slot(h, s);
    return rc;
}

int icb_profile_jpg_fit_image(icb_handle *h, const icb_image *img, float init_q, int64_t target, int reps,
                              float *transform_ms, float *entropy_ms, float *total_ms, float *entropy_kernel_ms,
                              int *entropy_kernel_launches)
{
    if (!h || !img || reps < 1) return ICB_ERR_INVALID;
    if (img->fmt != ICB_FMT_RGB8 && img->fmt != ICB_FMT_BGR8) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    Slot *s = acquire_slot(h, img->devi);
    int rc = ICB_OK;
    cudaEvent_t e0 = nullptr, e1 = nullptr, e2 = nullptr, pe[16] = {nullptr};
    float xf = 0, ent = 0, entk = 0;
    int launches = 0;
    do {
        bool ok = cudaSetDevice(h->devs[s->devi].dev) == cudaSuccess && cudaEventCreate(&e0) == cudaSuccess &&
                  cudaEventCreate(&e1) == cudaSuccess && cudaEventCreate(&e2) == cudaSuccess;
        for (int i = 0; i < 16 && ok; i++) ok = cudaEventCreate(&pe[i]) == cudaSuccess;
        if (!ok) { rc = ICB_ERR_CUDA; break; }
        for (int r = 0; r < reps && !rc; r++) {
            const uint64_t k0 = h->kernels.load();
            cudaEventRecord(e0, s->stream);
            rc = enqueue_transform(h, s, img->d_px, img->w, img->h, img->pitch, img->fmt);
            if (rc) break;
            const uint64_t k1 = h->kernels.load();
            cudaEventRecord(e1, s->stream);
            const long long cap = std::min<long long>(target, worst_stream_bytes(s->geom));
            rc = ensure_out(h, s, cap);
            if (rc) break;
            const int npairs = max_search_iters(init_q);
            rc = enqueue_passes(h, s, MODE_SEARCH, init_q, target, npairs, true, cap, pe,
                                expect_first_pass_fails(h, img->w, img->h, init_q, target));
            if (rc) break;
            cudaEventRecord(e2, s->stream);
            cudaError_t e = cudaMemcpyAsync(s->h_state, s->d_state, sizeof(FitState), cudaMemcpyDeviceToHost,
                                            s->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
            if (e != cudaSuccess) { rc = fail(h, ICB_ERR_CUDA, "profile", e); break; }
            float a = 0, b = 0;
            cudaEventElapsedTime(&a, e0, e1);
            cudaEventElapsedTime(&b, e1, e2);
            xf += a;
            ent += b;
            launches = std::min(s->h_state->passes, npairs);       // passes that did work
            for (int i = 0; i < launches; i++) {
                float c = 0;
                cudaEventElapsedTime(&c, pe[2 * i], pe[2 * i + 1]);
                entk += c;
            }
#ifdef ICB_TIMING
            if (r == reps - 1) {
                const FitState &fs = *s->h_state;
                fprintf(stderr, "warp timeline (cycles from entry):");
                for (int i = 1; i < 24; i++)
                    fprintf(stderr, " %lld", fs.dbg_w[i] >= fs.dbg_w[0] ? (long long)(fs.dbg_w[i] - fs.dbg_w[0]) : -1ll);
                fprintf(stderr, "\n");
                for (int i = 0; i < fs.passes && i < 8; i++)
                    fprintf(stderr, "pass %d: tiles %.1f us, scan %.1f us, step %.1f us, gap to next %.1f us\n", i,
                            (fs.dbg_t[i][1] - fs.dbg_t[i][0]) * 1e-3, (fs.dbg_t[i][2] - fs.dbg_t[i][1]) * 1e-3,
                            (fs.dbg_t[i][3] - fs.dbg_t[i][2]) * 1e-3,
                            i + 1 < fs.passes ? ((double)fs.dbg_t[i + 1][0] - (double)fs.dbg_t[i][3]) * 1e-3 : 0.0);
            }
#endif
            h->passes += (uint64_t)s->h_state->passes;
            h->last_xf_launches = (int)(k1 - k0);
            h->last_ent_launches = (int)(h->kernels.load() - k1);
        }
    } while (0);
    if (rc) cudaStreamSynchronize(s->stream);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (e2) cudaEventDestroy(e2);
    for (int i = 0; i < 16; i++) if (pe[i]) cudaEventDestroy(pe[i]);
    xf /= (float)reps;
    ent /= (float)reps;
    entk /= (float)reps;
    h->last_xf_ms = xf;
    h->last_ent_ms = ent;
    h->last_total_ms = xf + ent;
    if (transform_ms) *transform_ms = xf;
    if (entropy_ms) *entropy_ms = ent;
    if (total_ms) *total_ms = xf + ent;
    if (entropy_kernel_ms) *entropy_kernel_ms = entk;
    if (entropy_kernel_launches) *entropy_kernel_launches = launches;
    release_slot(h, s);
    return rc;
}

// Device-side stopwatch over ALL slot streams of one GPU: mark(0) before the first
// submit, mark(1) after the last wait; elapsed = CUDA-event time between them.
int icb_mark(icb_handle *h, int device_index, int which)
{
    if (!h || device_index < 0 || device_index >= (int)h->devs.size() || which < 0 || which > 1) return ICB_ERR_INVALID;
    DeviceCtx &d = h->devs[device_index];
    CK(cudaSetDevice(d.dev));
    if (!d.timer_stream) {
        CK(cudaStreamCreateWithFlags(&d.timer_stream, cudaStreamNonBlocking));
        CK(cudaEventCreate(&d.mark[0]));
        CK(cudaEventCreate(&d.mark[1]));
        CK(cudaEventCreateWithFlags(&d.join, cudaEventDisableTiming));
    }
    for (Slot &s : h->slots) {
        if (s.devi != device_index) continue;
        CK(cudaEventRecord(d.join, s.stream));
        CK(cudaStreamWaitEvent(d.timer_stream, d.join, 0));
    }
    CK(cudaEventRecord(d.mark[which], d.timer_stream));
    return ICB_OK;
}

int icb_elapsed_ms(icb_handle *h, int device_index, float *ms)
{
    if (!h || !ms || device_index < 0 || device_index >= (int)h->devs.size()) return ICB_ERR_INVALID;
    DeviceCtx &d = h->devs[device_index];
    if (!d.timer_stream) return fail(h, ICB_ERR_INVALID, "icb_mark not called");
    CK(cudaSetDevice(d.dev));
    CK(cudaEventSynchronize(d.mark[1]));
    CK(cudaEventElapsedTime(ms, d.mark[0], d.mark[1]));
    return ICB_OK;
}

}  // extern "C"
