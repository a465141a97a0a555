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
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/imgcomp_b200.h"
#include "common.cuh"
#include "k_control.cuh"
#include "k_decode.cuh"
#include "jpeg_parse.h"
#include "k_entropy.cuh"
#include "k_png.cuh"
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

struct DecJob {                          // per-job block of the JPEG reader in device memory: parameters + the file's tables
    DecParams p;
    DecTables t;
};

enum Phase { PH_IDLE = 0, PH_CACHED, PH_SEARCH, PH_RESIZE, PH_PNG, PH_PNG_FETCH, PH_DONE };

struct Job {
    Phase phase = PH_IDLE;
    const unsigned char *d_src = nullptr;   // original pixels in HBM
    int w = 0, h = 0, pitch = 0, fmt = 0;
    float init_q = 0;
    int64_t target = 0;
    int has_cached = 0;
    float cached_q = 0;
    double cached_scale = 1.0;
    unsigned char *out = nullptr;           // caller's output buffer
    size_t out_cap = 0;
    bool out_pinned = false;                // the device can copy straight into it
    bool stage_stream = false;              // the JFIF stream of this job lands in the slot's h_out first
    size_t staged_out = 0;                  // bytes waiting in the slot's pinned h_out for the caller's pageable buffer
    int out_rows = 0, out_row_bytes = 0, out_pitch = 0;   // PH_RESIZE: shape of the staged rows
    double scale = 1.0;                     // scale of the phase in flight
    bool coef_is_scale1 = false;            // coefficients in HBM belong to the unscaled image
    long long dev_cap = 0;                  // capacity handed to the kernels
    int passes = 0;
    bool xf_this_phase = false;             // the phase in flight ran the transform (its kernel clock is fresh)
    bool from_file = false;                 // the pixels come out of the GPU JPEG reader: its verdict counts
    unsigned int dec_nblocks = 0, dec_nrst = 0;
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
    // page-locked staging for callers that bring pageable memory (a Java heap array copied by FFM,
    // a numpy array): pixels are copied in here slice by slice with the H2D of each slice queued behind
    // it, results land here and are copied out when the job is collected
    unsigned char *h_in = nullptr;    size_t h_in_cap = 0;
    unsigned char *h_out = nullptr;   size_t h_out_cap = 0;
    unsigned char *d_png = nullptr;   size_t png_cap = 0;       // PNG encoder work arena (k_png.cuh)
    unsigned char *d_file = nullptr;  size_t file_cap = 0;      // the PNG file as it is assembled
    unsigned long long *h_len = nullptr;                        // pinned: length of the PNG file of the job in flight
    // GPU JPEG reader (k_decode.cuh): the file, its unstuffed scan, per-subsequence states, planes, tables
    unsigned char *d_file_in = nullptr; size_t file_in_cap = 0;
    unsigned char *d_bits = nullptr;    size_t bits_cap = 0;
    unsigned char *d_sub = nullptr;     size_t sub_cap = 0;
    unsigned char *d_planes = nullptr;  size_t planes_cap = 0;
    unsigned char *d_dec = nullptr;     size_t dec_cap = 0;     // decoded raster when the decode rule subsamples it
    DecJob *d_decjob = nullptr;         DecJob *h_decjob = nullptr;         // h_*: pinned
    DecState *d_decst = nullptr;        DecState *h_decst = nullptr;
    size_t dec_scan_cap = 0, dec_blocks_cap = 0; int dec_w_cap = 0, dec_h_cap = 0;              // what the slot's decode graph is sized for
    cudaGraphExec_t dec_graph = nullptr; uint64_t dec_graph_key = 0; bool dec_graph_broken = false;
    Geom geom{};
    bool busy = false;
    bool lone = false;                      // the job in flight found no other slot of its GPU busy (enqueue_transform)
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
    std::unique_ptr<std::mutex[]> slot_mu;   // one per slot: serialises icb_poll / icb_wait on the same ticket
    int rr = 0;
    int table_mode = 0, jfif_minor = 2;
    PFN_encodeTiled encode_tiled = nullptr;
    std::string last_error;
    std::atomic<uint64_t> images_done{0}, kernels{0}, passes{0}, h2d{0}, d2h{0};
    // kernel clocks summed over finished jobs (FitState.kt): [0] transform, [1] entropy passes, [2] unused
    std::atomic<uint64_t> kclk_ns[3], kclk_n[3];
    float last_xf_ms = 0, last_ent_ms = 0, last_total_ms = 0;
    int last_xf_launches = 0, last_ent_launches = 0;
    // what the first midpoint of earlier searches of the same (w, h, q, target) did: key << 1 | failed
    std::atomic<uint64_t> first_pass_seen[256];
    icb_handle() { for (auto &e : first_pass_seen) e.store(0); for (int i = 0; i < 3; i++) { kclk_ns[i].store(0); kclk_n[i].store(0); } }
};

namespace {

// icb_last_error hands out the calling thread's own copy: another thread's failure cannot reassign the
// string while the caller reads it
thread_local std::string tl_last_error;

int fail(icb_handle *h, int code, const char *what, cudaError_t e = cudaSuccess)
{
    tl_last_error = std::string(what) + (e != cudaSuccess ? std::string(": ") + cudaGetErrorString(e) : "");
    if (h) {
        std::lock_guard<std::mutex> lk(h->mu);
        h->last_error = tl_last_error;
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
// ICB_PDL=1 / 0 forces it on / off; unset: on for a job that has its GPU to itself (latency: each kernel's
// set-up hides behind its predecessor's tail), off when other slots of the GPU are working -- early-resident
// CTAs that sit in griddepcontrol.wait hold registers and shared memory other jobs' kernels could use
// (measured r02, six slots busy: 10.3 k -> 11.0 k images/s with it off)
int pdl_policy()
{
    static const int v = [] { const char *e = getenv("ICB_PDL"); return !e ? -1 : (e[0] == '0' ? 0 : 1); }();
    return v;
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
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

inline int bpp_of(int fmt)
{
    switch (fmt) {
    case ICB_FMT_RGBA8: case ICB_FMT_ABGR8: case ICB_FMT_ARGB32: case ICB_FMT_XRGB32: return 4;
    case ICB_FMT_GRAY8: return 1;
    default: return 3;
    }
}
inline bool fmt_valid(int fmt) { return fmt >= ICB_FMT_RGB8 && fmt <= ICB_FMT_XRGB32; }
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
    unsigned huf[HUF_WORDS] = {0};                           // layout: common.cuh g_huf
    for (int t = 0; t < 2; t++) {
        memcpy(huf + 512 * t, ac_enc[t], sizeof(ac_enc[t]));
        for (int a = 0; a < 3; a++)
            for (int nb = 0; nb < 16; nb++) {
                const unsigned dc = dc_enc[t][nb];
                if (!(dc & 0xffu)) continue;                 // sizes 12-15 do not exist
                const unsigned eob = a ? ac_enc[a - 1][0] : 0u;     // symbol 0x00 of the previous block's AC table
                const unsigned len = (eob & 0xffu) + (dc & 0xffu);
                huf[512 * t + 256 + 16 * a + nb] = ((((eob >> 8) << (dc & 0xffu)) | (dc >> 8)) << 8) | len;
            }
    }
    CK(cudaMemcpyToSymbol(g_huf, huf, sizeof(huf)));
    {   // CRC-32 byte table and the "256 zero bytes" advance operator (k_png.cuh g_crc)
        static uint32_t crc[256 + 4 * 256];
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            crc[i] = c;
        }
        for (int j = 0; j < 4; j++)
            for (uint32_t b = 0; b < 256; b++) {
                uint32_t c = b << (8 * j);
                for (int k = 0; k < PNG_CRC_SEG; k++) c = crc[c & 0xffu] ^ (c >> 8);
                crc[256 + 256 * j + b] = c;
            }
        CK(cudaMemcpyToSymbol(g_crc, crc, sizeof(crc)));
    }
    unsigned char hdr[HEADER_BYTES + 16];
    build_header_template(hdr, dc_bits, ac_bits, ac_vals);
    CK(cudaMemcpyToSymbol(c_hdr, hdr, HEADER_BYTES));
    const int rs_max = resize_smem_bytes(RS_MAX_SEG, RS_MAX_TW, 3);
    CK(cudaFuncSetAttribute(k_resize<4, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, rs_max));
    CK(cudaFuncSetAttribute(k_resize<4, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, rs_max));
    CK(cudaFuncSetAttribute(k_resize<4, -1>, cudaFuncAttributeMaxDynamicSharedMemorySize, rs_max));
    CK(cudaFuncSetAttribute(k_resize<3, -1>, cudaFuncAttributeMaxDynamicSharedMemorySize, rs_max));
    CK(cudaFuncSetAttribute(k_resize<1, -1>, cudaFuncAttributeMaxDynamicSharedMemorySize, rs_max));
    CK(cudaFuncSetAttribute(k_transform<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, XF_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_transform<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, XF_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_transform<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, XF_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_transform<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, XF_SMEM_BYTES));
    // (function attributes belong to the device: every GPU of the handle needs them, not just the first one used)
    CK(cudaFuncSetAttribute(k_dec_write, cudaFuncAttributeMaxDynamicSharedMemorySize, DEC_WBUF_BYTES));
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

// icb_run_list's workers hold several tickets each; a worker that WAITED here for a slot (of one particular GPU, for a
// resident image) while holding tickets could wait for a slot only another waiting worker can free -- measured r02 on two
// GPUs: four workers, all blocked, none collecting.  Inside the executor acquire_slot therefore does not wait: the
// submit call comes back with ICB_BUSY_INTERNAL and the worker collects its oldest job before it tries again.
thread_local bool tl_acquire_nowait = false;
constexpr int ICB_BUSY_INTERNAL = 2;

// A free slot on the GPU with the fewest busy slots (devi >= 0: on that GPU); on a tie the GPU after the one
// taken last (so equally loaded GPUs take turns), and on a GPU its lowest free slot -- a caller that runs one job
// at a time keeps coming back to arenas that are sized for its images already instead of walking through every
// slot of the handle (each first visit costs the cudaMalloc / cudaFree of growing arenas: measured on cfg5,
// 3.4 ms per fit while the walk lasted against 1.3 ms).
Slot *acquire_slot(icb_handle *h, int devi = -1)
{
    std::unique_lock<std::mutex> lk(h->mu);
    const int n = (int)h->slots.size(), nd = (int)h->devs.size();
    while (true) {
        int busy[64] = {0};
        for (const Slot &s : h->slots) if (s.busy && s.devi < 64) busy[s.devi]++;
        Slot *best = nullptr;
        int best_key = 0;
        for (int k = 0; k < n; k++) {
            Slot &s = h->slots[k];
            if (s.busy || (devi >= 0 && s.devi != devi)) continue;
            const int turn = nd > 0 ? (s.devi - h->rr + nd) % nd : 0;       // 0 = the GPU whose turn it is
            const int key = (busy[s.devi & 63] * 64 + turn) * 65536 + k;
            if (!best || key < best_key) { best = &s; best_key = key; }
        }
        if (best) {
            best->busy = true;
            best->gen++;
            h->rr = nd > 0 ? (best->devi + 1) % nd : 0;
            return best;
        }
        if (tl_acquire_nowait) return nullptr;          // (icb_run_list: the worker collects one of its own jobs first)
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

// What a fit job asks of the entropy kernels: the mode, the quality it starts from, how many passes to
// queue, and the quality the first survivor lists are built for.
struct FitPlan {
    int mode;               // FitMode
    float q;                // initial quality (search) or the one quality to encode at
    long long target;
    int npasses;
    bool prune;             // search expected to overshoot at its first midpoint: lists for that midpoint, pruned as hi drops
    float list_q;
};

FitPlan make_plan(int mode, float q, long long target, int npasses, bool expect_fail)
{
    FitPlan pl;
    pl.mode = mode; pl.q = q; pl.target = target; pl.npasses = npasses;
    pl.prune = mode == MODE_SEARCH && expect_fail && npasses > 1;
    pl.list_q = pl.prune ? (0.0f + q) / 2.0f : q;               // = the first midpoint (:168)
    return pl;
}

// per-slot entropy arenas for geometry g: survivor lists, unit links, two pass arenas
int ensure_entropy_arenas(icb_handle *h, Slot *s, const Geom &g)
{
    const int nchunks = (g.ntiles + ENT_WARPS - 1) / ENT_WARPS;
    const size_t per = round_up(ent_pass_bytes(g.ntiles, nchunks, ent_scr_stride(g.unit_group)), 256);
    int rc = ensure(h, s->d_pass, s->pass_cap, 2 * per + 256);
    if (rc) return rc;
    return ensure(h, s->d_ent, s->ent_cap, ent_list_bytes(g.nunits));
}

void fill_ent_params(const Slot *s, EntParams &ep)
{
    const Geom &g = s->geom;
    const int nchunks = (g.ntiles + ENT_WARPS - 1) / ENT_WARPS;      // chunk = the 8 tiles one CTA takes per round
    const int stride_t = ent_scr_stride(g.unit_group);
    const size_t per = round_up(ent_pass_bytes(g.ntiles, nchunks, stride_t), 256);
    ep.coef = s->d_coef;
    ep.st = s->d_state;
    ep.ent = s->d_ent;
    ep.ent_cnt = s->d_ent + (size_t)g.nunits * UNIT_CAP;
    ep.unit_info = reinterpret_cast<unsigned short *>(ep.ent_cnt + round_up((size_t)g.nunits, 4));
    ep.unit_group = g.unit_group; ep.scr_stride = stride_t;
    ep.units_per_row = g.units_per_row; ep.mcu_cols = g.mcu_cols;
    ep.nblocks = g.nblocks;
    ep.nunits = g.nunits;
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
}

// The forward transform of one image; with a plan it also writes the survivor lists the plan's passes
// start from (the stand-alone list kernel k_entropy_prepare is then not needed).
int enqueue_transform(icb_handle *h, Slot *s, const unsigned char *d_px, int w, int hgt, int pitch, int fmt,
                      const FitPlan *plan)
{
    static const int forced_tile = [] { const char *e = getenv("ICB_ENT_TILE"); return e ? atoi(e) : 0; }();   // tests / A-B runs
    // a fit that has the GPU to itself (no other slot of this device in use) is a latency case: one unit
    // per tile gives its passes more warps; the tile size never changes a byte of the result
    int tile_pick = forced_tile;
    if (!tile_pick) {
        int busy = 0;
        {
            std::lock_guard<std::mutex> lk(h->mu);
            for (const Slot &o : h->slots) busy += (o.busy && o.devi == s->devi) ? 1 : 0;
        }
        if (busy <= 1) tile_pick = 128;
        s->lone = busy <= 1;
    }
    const Geom g = make_geom(w, hgt, pitch, fmt == ICB_FMT_BGR8, tile_pick);
    s->geom = g;
    int rc = ensure(h, s->d_coef, s->coef_cap, (size_t)g.nblocks * 128);
    if (rc) return rc;
    XfLists xl = {};
    if (plan) {
        rc = ensure_entropy_arenas(h, s, g);
        if (rc) return rc;
        EntParams ep;
        fill_ent_params(s, ep);
        xl.ent = ep.ent; xl.ent_cnt = ep.ent_cnt; xl.unit_info = ep.unit_info;
        xl.list_q = plan->list_q; xl.table_mode = h->table_mode;
        xl.st = s->d_state;
    }
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)(pitch / 4), (cuuint64_t)hgt};
    const cuuint64_t gstr[1] = {(cuuint64_t)pitch};
    const cuuint32_t box[2] = {TILE_ROW_BYTES / 4, 16};
    const cuuint32_t estr[2] = {1, 1};
    CUresult cr = h->encode_tiled(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, (void *)d_px, gdim, gstr, box, estr,
                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) return fail(h, ICB_ERR_CUDA, "cuTensorMapEncodeTiled failed");
    const int tiles_per_row = g.units_per_row;
    const int ntiles = g.nunits;
    static const int xf_cps = [] { const char *e = getenv("ICB_XF_CPS"); return e ? atoi(e) : XF_CTAS; }();   // tuning knob
    const int grid = std::min(ntiles, h->devs[s->devi].sms * std::max(1, std::min(XF_CTAS, xf_cps)));
    xl.st = s->d_state;                     // (kernel clock, with or without lists)
    if (plan) {
        if (g.bgr) k_transform<true, true><<<grid, XF_THREADS, XF_SMEM_BYTES, s->stream>>>(tmap, g, s->d_coef, tiles_per_row, ntiles, xl);
        else k_transform<false, true><<<grid, XF_THREADS, XF_SMEM_BYTES, s->stream>>>(tmap, g, s->d_coef, tiles_per_row, ntiles, xl);
    } else {
        if (g.bgr) k_transform<true, false><<<grid, XF_THREADS, XF_SMEM_BYTES, s->stream>>>(tmap, g, s->d_coef, tiles_per_row, ntiles, xl);
        else k_transform<false, false><<<grid, XF_THREADS, XF_SMEM_BYTES, s->stream>>>(tmap, g, s->d_coef, tiles_per_row, ntiles, xl);
    }
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

// The entropy side of a job: job start (with lists_ready the transform wrote the survivor lists and only
// the unit links are missing; otherwise the stand-alone list kernel builds them from the coefficients),
// the plan's passes, and optionally the emission of the best stream.
int enqueue_passes(icb_handle *h, Slot *s, const FitPlan &pl, bool lists_ready, bool emit, long long dev_cap,
                   cudaEvent_t *pass_ev = nullptr)
{
    const Geom &g = s->geom;
    int rc = ensure_entropy_arenas(h, s, g);
    if (rc) return rc;
    EntParams ep;
    fill_ent_params(s, ep);
    const int nctas = ep.nchunks;
    BeginArgs ba;
    ba.mode = pl.mode; ba.q = pl.q; ba.target = pl.target; ba.table_mode = h->table_mode; ba.jfif_minor = h->jfif_minor;
    ba.list_q = pl.list_q;
    ba.pad = 0;
    const int pneed = (g.nunits + PREP_WARPS - 1) / PREP_WARPS;
    const int pwave = h->devs[s->devi].sms * PREP_CTAS_PER_SM;
    const int prounds = (pneed + pwave - 1) / pwave;             // units a warp takes
    const dim3 prep_grid((pneed + prounds - 1) / prounds);
    const bool pdl0 = pdl_policy() < 0 ? s->lone : pdl_policy() == 1;
    if (lists_ready)
        CK(launch(k_job_begin, dim3((g.nunits + BEGIN_THREADS - 1) / BEGIN_THREADS), dim3(BEGIN_THREADS), 0, s->stream,
                  pdl0, ep, ba));
    else
        CK(launch(k_entropy_prepare<false>, prep_grid, dim3(PREP_THREADS), 0, s->stream, pdl0, ep, ba));
    h->kernels += 1;
    const int pgrid = pass_grid(g.ntiles, h->devs[s->devi].sms);
    const bool pdl = pass_ev == nullptr && (pdl_policy() < 0 ? s->lone : pdl_policy() == 1);
    for (int i = 0; i < pl.npasses; i++) {
        if (pass_ev) cudaEventRecord(pass_ev[2 * i], s->stream);
        CK(launch(k_entropy_pass, dim3(pgrid), dim3(ENT_THREADS), 0, s->stream, pdl, ep));
        if (pass_ev) cudaEventRecord(pass_ev[2 * i + 1], s->stream);
        h->kernels += 1;
        if (pl.prune && i + 1 < pl.npasses) {
            if (i == 0) {
                CK(launch(k_entropy_prepare<true>, prep_grid, dim3(PREP_THREADS), 0, s->stream, pdl, ep, ba));
                h->kernels += 1;
            }
            CK(launch(k_entropy_prune, dim3((g.nunits + ENT_WARPS - 1) / ENT_WARPS), dim3(ENT_THREADS), 0, s->stream, pdl,
                      ep, i + 1));
            h->kernels += 1;
        }
    }
    if (emit) {
        CK(launch(k_entropy_emit, dim3(nctas), dim3(ENT_THREADS), 0, s->stream, pdl, ep, s->d_out, dev_cap, g.w, g.h));
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
    // destination tile of the CTA: the widest power of two (<= 128 pixels) whose source row segment --
    // (tw - 1) source steps plus the two outer taps, rounded out to 16-byte vectors -- fits the staging rows
    const int bpp = bpp_of(fmt);
    const double step = (double)w / dw;
    static const int tl_max = [] { const char *e = getenv("ICB_RS_TL"); const int v = e ? atoi(e) : 7; return v >= 2 && v <= 7 ? v : 7; }();   // A/B knob
    static const int cps_max = [] { const char *e = getenv("ICB_RS_CPS"); const int v = e ? atoi(e) : 6; return v >= 1 && v <= 8 ? v : 6; }();
    int tl = tl_max;
    size_t seg = 0;
    for (; tl >= 2; tl--) {
        seg = round_up((size_t)(((1 << tl) - 1) * step + 3.0) * bpp + 32, 16);
        if (seg <= (size_t)RS_MAX_SEG) break;
    }
    if (tl < 2) {
        // more than ~200 source pixels per destination pixel (a 1 x 1 thumbnail of a large image): the taps of
        // neighbouring pixels share nothing, so there is nothing to stage -- plain gather, one thread per pixel
        rp.tw_log2 = 0; rp.seg_bytes = 0;
        const dim3 ggrid((dw + 31) / 32, (dh + 7) / 8);
        switch (fmt) {
        case ICB_FMT_RGBA8: case ICB_FMT_ARGB32: k_resize_gather<4, 3><<<ggrid, 256, 0, s->stream>>>(rp); break;
        case ICB_FMT_ABGR8: k_resize_gather<4, 0><<<ggrid, 256, 0, s->stream>>>(rp); break;
        case ICB_FMT_XRGB32: k_resize_gather<4, -1><<<ggrid, 256, 0, s->stream>>>(rp); break;
        case ICB_FMT_GRAY8: k_resize_gather<1, -1><<<ggrid, 256, 0, s->stream>>>(rp); break;
        default: k_resize_gather<3, -1><<<ggrid, 256, 0, s->stream>>>(rp); break;
        }
        CK(cudaGetLastError());
        h->kernels++;
        return ICB_OK;
    }
    rp.tw_log2 = tl;
    rp.seg_bytes = (int)seg;
    const int tw = 1 << tl;
    rp.tiles_x = (dw + tw - 1) / tw;
    rp.ntiles = rp.tiles_x * ((dh + RS_ROWS - 1) / RS_ROWS);
    const size_t smem = (size_t)resize_smem_bytes(rp.seg_bytes, tw, bpp);
    // persistent CTAs: as many as fit an SM by shared memory (at most 6), one wave
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>((size_t)cps_max, ((size_t)220 << 10) / smem));
    const dim3 grid((unsigned)std::min(rp.ntiles, h->devs[s->devi].sms * per_sm));
    switch (fmt) {
    case ICB_FMT_RGBA8: case ICB_FMT_ARGB32: k_resize<4, 3><<<grid, RS_THREADS, smem, s->stream>>>(rp); break;
    case ICB_FMT_ABGR8: k_resize<4, 0><<<grid, RS_THREADS, smem, s->stream>>>(rp); break;
    case ICB_FMT_XRGB32: k_resize<4, -1><<<grid, RS_THREADS, smem, s->stream>>>(rp); break;
    case ICB_FMT_GRAY8: k_resize<1, -1><<<grid, RS_THREADS, smem, s->stream>>>(rp); break;
    default: k_resize<3, -1><<<grid, RS_THREADS, smem, s->stream>>>(rp); break;
    }
    CK(cudaGetLastError());
    h->kernels++;
    return ICB_OK;
}

// Is p page-locked (or device-accessible) memory the copy engines can reach without a bounce buffer?
bool is_pinned(const void *p)
{
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged || at.type == cudaMemoryTypeDevice;
}

int ensure_pinned(icb_handle *h, unsigned char *&p, size_t &cap, size_t need)
{
    if (need <= cap) return ICB_OK;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    void *q = nullptr;
    cudaError_t e = cudaHostAlloc(&q, need, cudaHostAllocPortable | cudaHostAllocMapped);
    if (e != cudaSuccess) return fail(h, ICB_ERR_OOM, "cudaHostAlloc", e);
    p = (unsigned char *)q; cap = need;
    return ICB_OK;
}

// Host pixels -> the slot's device buffer.  Pinned sources copy asynchronously as they are; pageable
// ones go through the slot's pinned staging in slices of a few MB, the calling thread copying slice k + 1
// while the DMA engine moves slice k (several caller threads -- the reference's pool threads,
// CompressionBatch.java:64-88 -- stage their images side by side).
int upload(icb_handle *h, Slot *s, unsigned char *d, int dpitch, const unsigned char *px, int w, int hgt,
           int pitch, int bpp)
{
    cudaStream_t st = s->stream;
    const size_t row = (size_t)w * bpp;
    h->h2d += (uint64_t)row * hgt;
    if (is_pinned(px)) {
        if (pitch == dpitch)
            CK(cudaMemcpyAsync(d, px, (size_t)(hgt - 1) * pitch + row, cudaMemcpyHostToDevice, st));
        else
            CK(cudaMemcpy2DAsync(d, dpitch, px, pitch, row, hgt, cudaMemcpyHostToDevice, st));
        return ICB_OK;
    }
    int rc = ensure_pinned(h, s->h_in, s->h_in_cap, row * hgt + 16);
    if (rc) return rc;
    const int rows_per = (int)std::max<size_t>(1, ((size_t)4 << 20) / row);
    for (int y = 0; y < hgt; y += rows_per) {
        const int n = std::min(rows_per, hgt - y);
        unsigned char *stage = s->h_in + (size_t)y * row;
        if ((size_t)pitch == row) memcpy(stage, px + (size_t)y * pitch, row * n);
        else for (int r = 0; r < n; r++) memcpy(stage + (size_t)r * row, px + (size_t)(y + r) * pitch, row);
        CK(cudaMemcpy2DAsync(d + (size_t)y * dpitch, dpitch, stage, row, row, n, cudaMemcpyHostToDevice, st));
    }
    return ICB_OK;
}

// Queue the copy of the job's result rows (n_rows x row_bytes at device pitch spitch) to the caller:
// straight into a pinned destination, else into the slot's pinned h_out (collect_output finishes it).
int fetch_rows(icb_handle *h, Slot *s, const unsigned char *d_src, size_t spitch, size_t row_bytes, int n_rows,
               size_t out_pitch)
{
    Job &j = s->job;
    h->d2h += (uint64_t)row_bytes * n_rows;
    if (j.out_pinned) {
        CK(cudaMemcpy2DAsync(j.out, out_pitch, d_src, spitch, row_bytes, n_rows, cudaMemcpyDeviceToHost, s->stream));
        return ICB_OK;
    }
    int rc = ensure_pinned(h, s->h_out, s->h_out_cap, row_bytes * n_rows + 16);
    if (rc) return rc;
    CK(cudaMemcpy2DAsync(s->h_out, row_bytes, d_src, spitch, row_bytes, n_rows, cudaMemcpyDeviceToHost, s->stream));
    j.staged_out = row_bytes * n_rows;
    j.out_rows = n_rows; j.out_row_bytes = (int)row_bytes; j.out_pitch = (int)out_pitch;
    return ICB_OK;
}

void collect_output(Slot *s)                                // after the stream drained
{
    Job &j = s->job;
    if (!j.staged_out || !j.out) return;
    for (int r = 0; r < j.out_rows; r++)
        memcpy(j.out + (size_t)r * j.out_pitch, s->h_out + (size_t)r * j.out_row_bytes, (size_t)j.out_row_bytes);
    j.staged_out = 0;
}

}  // namespace

// ======================================================================
//  JPG fit state machine (core/ImageCompressionJpg.java:77-122)
// ======================================================================
namespace {

bool fetch_by_kernel()
{
    static const bool v = [] { const char *e = getenv("ICB_FETCH_KERNEL"); return e && e[0] == '1'; }();
    return v;
}

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
    const FitPlan pl = j.phase == PH_CACHED
        ? make_plan(MODE_SINGLE, j.cached_q, j.target, 1, false)
        : make_plan(MODE_SEARCH, j.init_q, j.target, max_search_iters(j.init_q),
                    scale < 1.0 || expect_first_pass_fails(h, j.w, j.h, j.init_q, j.target));
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
    j.xf_this_phase = need_xf;
    // A/B knob: ICB_FUSED_LISTS=0 makes the transform write coefficients only and the stand-alone list kernel
    // build the survivor lists from them (the round-1 structure); results are identical either way
    static const bool fused = [] { const char *e = getenv("ICB_FUSED_LISTS"); return !(e && e[0] == '0'); }();
    bool lists_ready = need_xf && fused;
    if (need_xf) {
        int rc = enqueue_transform(h, s, src, w, hgt, pitch, j.fmt, fused ? &pl : nullptr);
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
    int rc = enqueue_passes(h, s, pl, lists_ready, true, cap);
    if (rc) return rc;
    // the stream goes back by a kernel that knows best_size (exactly that many bytes, nothing on a phase
    // that failed): into the caller's buffer when it is page-locked and 16-byte aligned, else into the slot's
    // page-locked staging (copied out when the job is collected)
    // The stream goes back on the copy engine, speculatively: min(out_cap, cap) bytes are queued behind the emit
    // kernel now (the host learns best_size only after the phase drained; a second, exactly sized copy would cost
    // a host round trip per image).  A phase that fails copies dead bytes; the caller's buffer past out_len is
    // unspecified.  ICB_FETCH_KERNEL=1 selects the alternative measured in r02 -- a kernel that knows best_size
    // and writes exactly that many bytes into the page-locked destination: 12 % slower with six slots busy
    // (zero-copy stores over PCIe occupy SM store queues that the entropy passes of other jobs want).
    const bool fetch_kernel = fetch_by_kernel();
    if (j.out && j.out_cap) {
        unsigned char *dst = j.out;
        long long dcap = std::min<long long>((long long)j.out_cap, cap);
        if (!j.out_pinned || (fetch_kernel && ((uintptr_t)j.out & 15))) {
            rc = ensure_pinned(h, s->h_out, s->h_out_cap, round_up((size_t)cap, 16) + 32);
            if (rc) return rc;
            dst = s->h_out;
            j.stage_stream = true;
        }
        if (fetch_kernel) {
            CK(launch(k_fetch_stream, dim3(8), dim3(FETCH_THREADS), 0, s->stream, false, (const FitState *)s->d_state,
                      (const unsigned char *)s->d_out, cap, dst, dcap));
            h->kernels += 1;
        } else if (dcap > 0) {
            CK(cudaMemcpyAsync(dst, s->d_out, (size_t)dcap, cudaMemcpyDeviceToHost, s->stream));
            h->d2h += (uint64_t)dcap;
        }
    }
    CK(cudaMemcpyAsync(s->h_state, s->d_state, sizeof(FitState), cudaMemcpyDeviceToHost, s->stream));
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
    {   // kernel clocks of the phase that just drained
        auto add = [&](int k, const unsigned long long (&t)[2]) {
            if (t[1] > t[0] && t[0]) { h->kclk_ns[k] += t[1] - t[0]; h->kclk_n[k] += 1; }
        };
        if (j.xf_this_phase) add(0, st.kt[0]);
        for (int i = 0; i < st.passes && i < 8; i++) add(1, st.kt[2 + i]);
    }
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
        if (j.res.status == ICB_OK && j.out && st.best_size > 0) {
            if (fetch_by_kernel()) h->d2h += (uint64_t)st.best_size;     // k_fetch_stream moved exactly these bytes
            if (j.stage_stream) memcpy(j.out, s->h_out, (size_t)st.best_size);
        }
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
    j.out_pinned = out && is_pinned(out);
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

// ImageWriteParam.setCompressionQuality throws IllegalArgumentException outside [0, 1]; the reference would
// hit that as soon as a midpoint (or the cached quality) left the range.  NaN is refused as well.
bool bad_quality(float q) { return !(q >= 0.0f && q <= 1.0f); }

int check_fit_qualities(icb_handle *h, float init_q, int has_cached, float cached_q)
{
    if (init_q != init_q || init_q > 1.0f) return fail(h, ICB_ERR_INVALID, "initial quality outside [0, 1]");
    if (has_cached && bad_quality(cached_q)) return fail(h, ICB_ERR_INVALID, "cached quality outside [0, 1]");
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
    // The slot streams map onto the GPU's hardware work queues; with the default of 8, sixteen streams of short
    // dependent kernels share queues and wait for each other (measured r02, file path: 2 580 -> 3 090 images/s with
    // 16).  Only read when the CUDA context is created: a host that has used CUDA before this call keeps its own.
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
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
    h->slot_mu.reset(new std::mutex[h->slots.size()]);
    // interleave devices so round-robin acquisition spreads images over GPUs
    for (size_t i = 0; i < h->slots.size(); i++) {
        Slot &s = h->slots[i];
        s.index = (int)i;
        s.devi = (int)(i % devs.size());
        if (cudaSetDevice(h->devs[s.devi].dev) != cudaSuccess ||
            cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaMalloc((void **)&s.d_state, sizeof(FitState)) != cudaSuccess ||
            cudaMallocHost((void **)&s.h_state, sizeof(FitState)) != cudaSuccess ||
            cudaMallocHost((void **)&s.h_len, 64) != cudaSuccess) {
            icb_shutdown(h);
            return ICB_ERR_OOM;
        }
        if (max_w > 0 && max_h > 0) {
            const Geom g = make_geom(max_w, max_h, dev_pitch(max_w, 3), 0);
            // every arena a job of that size needs, so that no cudaMalloc / cudaFree (a device-wide
            // synchronisation) happens while other slots are working
            if (ensure(h, s.d_rgb, s.rgb_cap, (size_t)g.pitch * max_h + 16) ||
                ensure(h, s.d_coef, s.coef_cap, (size_t)g.nblocks * 128) ||
                ensure_entropy_arenas(h, &s, g) || ensure(h, s.d_out, s.out_cap, ((size_t)4 << 20) + 32)) {
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
        cudaFree(s.d_png); cudaFree(s.d_file);
        cudaFree(s.d_file_in); cudaFree(s.d_bits); cudaFree(s.d_sub); cudaFree(s.d_planes); cudaFree(s.d_dec);
        cudaFree(s.d_decjob); cudaFree(s.d_decst);
        if (s.dec_graph) cudaGraphExecDestroy(s.dec_graph);
        if (s.h_decjob) cudaFreeHost(s.h_decjob);
        if (s.h_decst) cudaFreeHost(s.h_decst);
        if (s.h_len) cudaFreeHost(s.h_len);
        if (s.h_in) cudaFreeHost(s.h_in);
        if (s.h_out) cudaFreeHost(s.h_out);
    }
    delete h;
}

int icb_device_count(const icb_handle *h) { return h ? (int)h->devs.size() : 0; }

const char *icb_last_error(const icb_handle *h) { return h ? tl_last_error.c_str() : "no handle"; }

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
    // portable + mapped: every GPU of the handle can DMA from it and write results straight into it
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable | cudaHostAllocMapped) != cudaSuccess) return nullptr;
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
    if (!rc) rc = check_fit_qualities(h, init_q, has_cached, cached_q);
    if (rc) return rc;
    if (!ticket) return ICB_ERR_INVALID;
    Slot *s = acquire_slot(h);
    if (!s) return ICB_BUSY_INTERNAL;
    const int dp = dev_pitch(w, 3);
    if (cudaSetDevice(h->devs[s->devi].dev) != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "cudaSetDevice");
    if (!rc) rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)dp * hgt + 16);
    if (!rc) rc = upload(h, s, s->d_rgb, dp, px, w, hgt, pitch, 3);
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
    int rc = check_fit_qualities(h, init_q, has_cached, cached_q);
    if (rc) return rc;
    Slot *s = acquire_slot(h, img->devi);
    if (!s) return ICB_BUSY_INTERNAL;
    rc = start_fit(h, s, img->d_px, img->w, img->h, img->pitch, img->fmt, init_q, target, has_cached, cached_q,
                   cached_scale, out, out_cap);
    if (rc) { cudaStreamSynchronize(s->stream); release_slot(h, s); return rc; }
    *ticket = make_ticket(s);
    return ICB_OK;
}

}  // extern "C"

namespace {

// resize of pixels already in HBM into the slot's d_rgb2, result rows queued for dst (NULL: they stay in HBM)
int start_resize(icb_handle *h, Slot *s, const unsigned char *d_src, int w, int hgt, int pitch, int fmt, int dw, int dh,
                 uint8_t *dst, int dpitch)
{
    const int bpp = bpp_of(fmt);
    const int np = (int)round_up((size_t)dw * bpp, 16);
    int rc = ensure(h, s->d_rgb2, s->rgb2_cap, (size_t)np * dh + 16);
    if (!rc) rc = enqueue_resize(h, s, d_src, w, hgt, pitch, fmt, s->d_rgb2, dw, dh, np);
    s->job = Job();
    s->job.phase = PH_RESIZE;
    s->job.out = dst;
    s->job.out_pinned = dst && is_pinned(dst);
    if (!rc && dst) rc = fetch_rows(h, s, s->d_rgb2, (size_t)np, (size_t)dw * bpp, dh, (size_t)dpitch);
    return rc;
}

// PNG channel layout of a raster format: bytes per PNG pixel and, per PNG channel (R G B A), its source byte
void png_layout(int fmt, int &pbpp, int &perm)
{
    switch (fmt) {
    case ICB_FMT_BGR8:   pbpp = 3; perm = 0x012; break;
    case ICB_FMT_RGBA8:  pbpp = 4; perm = 0x3210; break;
    case ICB_FMT_ABGR8:  pbpp = 4; perm = 0x0123; break;
    case ICB_FMT_ARGB32: pbpp = 4; perm = 0x3012; break;     // bytes B G R A
    case ICB_FMT_XRGB32: pbpp = 3; perm = 0x012; break;      // bytes B G R x
    case ICB_FMT_GRAY8:  pbpp = 1; perm = 0x0; break;
    default:             pbpp = 3; perm = 0x210; break;      // RGB8
    }
}

// core/ImageCompressionPng.java:70 on the device: pixels in HBM -> a PNG file in the slot's d_file, its length in
// h_len, the file bytes on their way to the caller (exactly as many as the file has).
int enqueue_png(icb_handle *h, Slot *s, const unsigned char *d_src, int w, int hgt, int pitch, int fmt)
{
    Job &j = s->job;
    PngParams pp = {};
    pp.src = d_src; pp.w = w; pp.h = hgt; pp.pitch = pitch; pp.sbpp = bpp_of(fmt);
    png_layout(fmt, pp.pbpp, pp.perm);
    pp.nfilt = (long long)hgt * (1 + (long long)w * pp.pbpp);
    pp.npieces = (int)std::max<long long>(1, pp.nfilt / PNG_PIECE);
    const size_t np = (size_t)pp.npieces;
    const size_t o_filt = 0, o_seg = round_up((size_t)pp.nfilt + 64, 256);
    const size_t o_len = o_seg + (np + 1) * PNG_PIECE_WORDS * 4, o_pos = round_up(o_len + (np + 1) * 4, 8);
    const size_t o_adl = o_pos + (np + 2) * 8, o_hist = o_adl + np * 8, o_codes = o_hist + 320 * 4;
    const size_t o_misc = o_codes + 320 * 4, total = o_misc + 64;
    int rc = ensure(h, s->d_png, s->png_cap, total);
    const size_t fcap = (size_t)pp.nfilt * 2 + (size_t)pp.nfilt / 512 + 65536;     // 15 bits per byte at the very worst
    if (!rc) rc = ensure(h, s->d_file, s->file_cap, fcap);
    if (rc) return rc;
    pp.filt = s->d_png + o_filt;
    pp.seg_bits = reinterpret_cast<uint32_t *>(s->d_png + o_seg);
    pp.seg_len = reinterpret_cast<unsigned int *>(s->d_png + o_len);
    pp.seg_pos = reinterpret_cast<unsigned long long *>(s->d_png + o_pos);
    pp.adler = reinterpret_cast<uint2 *>(s->d_png + o_adl);
    pp.hist = reinterpret_cast<unsigned int *>(s->d_png + o_hist);
    pp.codes = reinterpret_cast<uint32_t *>(s->d_png + o_codes);
    pp.out_len = reinterpret_cast<unsigned long long *>(s->d_png + o_misc);
    pp.adler_out = reinterpret_cast<unsigned int *>(s->d_png + o_misc + 8);
    pp.out = s->d_file;
    pp.out_cap = (long long)fcap;
    cudaStream_t st = s->stream;
    CK(cudaMemsetAsync(pp.hist, 0, 320 * 4, st));
    k_png_filter<<<hgt, 256, 0, st>>>(pp);
    k_png_hist<<<(pp.npieces + 7) / 8, 256, 0, st>>>(pp);
    k_png_tree<<<1, PNG_TREE_THREADS, 0, st>>>(pp);
    k_png_encode<<<(pp.npieces + 7) / 8, 256, 0, st>>>(pp);
    k_png_scan<<<1, 1024, 0, st>>>(pp);
    k_png_merge<<<(pp.npieces + 1 + 7) / 8, 256, 0, st>>>(pp);
    const long long max_idat = ((long long)fcap + PNG_IDAT - 1) / PNG_IDAT;
    k_png_finish<<<(unsigned)((std::min<long long>(max_idat, ((long long)pp.nfilt * 2) / PNG_IDAT + 2) + 7) / 8), 256, 0, st>>>(pp);
    CK(cudaGetLastError());
    h->kernels += 7;
    // the file's length comes back first; drive_job then fetches exactly that many bytes (PH_PNG_FETCH)
    CK(cudaMemcpyAsync(s->h_len, pp.out_len, 8, cudaMemcpyDeviceToHost, st));
    j.phase = PH_PNG;
    return ICB_OK;
}

int check_resize_args(icb_handle *h, const void *src, int w, int hgt, int pitch, int fmt, const void *dst, int dw, int dh,
                      int dpitch)
{
    if (!h) return ICB_ERR_INVALID;
    if (!src || !dst || w < 1 || hgt < 1 || dw < 1 || dh < 1) return fail(h, ICB_ERR_INVALID, "bad resize arguments");
    if (!fmt_valid(fmt)) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    const int bpp = bpp_of(fmt);
    if (pitch < w * bpp || dpitch < dw * bpp) return fail(h, ICB_ERR_INVALID, "pitch smaller than a row");
    return ICB_OK;
}

// After the stream drained: what the GPU JPEG reader (k_decode.cuh) says about the job's file.  The fixed point of
// the synchronisation rounds must have been reached, the stream must hold the image's blocks, and the last block
// must end inside the last byte (the padding) of the entropy-coded data.
int decode_verdict(icb_handle *h, Slot *s)
{
    const DecState &ds = *s->h_decst;
    static const bool dbg = getenv("ICB_DEC_DEBUG") != nullptr;
    if (dbg) {
        int r = 1;
        while (r < DEC_ROUNDS && ds.active[r]) r++;
        fprintf(stderr, "icb decode: %u bytes of scan, rounds until nothing changed %d, blocks %u / %u, end bit %u of %u\n",
                ds.total_bytes, r, ds.blocks_found, s->job.dec_nblocks, ds.end_bit, ds.total_bytes * 8u);
    }
    if (ds.active[DEC_ROUNDS - 1]) return fail(h, ICB_ERR_UNSUPPORTED, "JPEG reader: no fixed point within the synchronisation rounds");
    if (ds.blocks_found < s->job.dec_nblocks) return fail(h, ICB_ERR_INVALID, "JPEG reader: entropy-coded data ends early");
    if (ds.n_rst != s->job.dec_nrst) return fail(h, ICB_ERR_INVALID, "JPEG reader: restart markers do not match the restart interval");
    const unsigned int total_bits = ds.total_bytes * 8u;
    if (ds.end_bit > total_bits || total_bits - ds.end_bit >= 16u)
        return fail(h, ICB_ERR_INVALID, "JPEG reader: data left over after the last block");
    return ICB_OK;
}

// Drives the job that holds slot s: wait = false returns false as soon as the stream still has work.
// On true the job is finished and s->job.res is final.
bool drive_job(icb_handle *h, Slot *s, bool wait)
{
    std::lock_guard<std::mutex> lk(h->slot_mu[s->index]);
    if (s->job.phase == PH_DONE) return true;
    int status = ICB_OK;
    cudaSetDevice(h->devs[s->devi].dev);
    while (true) {
        cudaError_t e = wait ? cudaStreamSynchronize(s->stream) : cudaStreamQuery(s->stream);
        if (!wait && e == cudaErrorNotReady) return false;
        if (e != cudaSuccess) { status = fail(h, ICB_ERR_CUDA, "stream", e); break; }
        if (s->job.phase == PH_RESIZE) { collect_output(s); s->job.res.status = ICB_OK; break; }
        if (s->job.phase == PH_PNG) {                    // the length is here: fetch the file, exactly that many bytes
            Job &j = s->job;
            j.res.out_len = (size_t)*s->h_len;
            j.res.device = h->devs[s->devi].dev;
            if (j.res.out_len > j.out_cap || !j.out) { status = ICB_ERR_CAPACITY; break; }
            unsigned char *dst = j.out;
            if (!j.out_pinned) {
                if (ensure_pinned(h, s->h_out, s->h_out_cap, j.res.out_len + 16)) { status = ICB_ERR_OOM; break; }
                dst = s->h_out;
                j.stage_stream = true;
            }
            e = cudaMemcpyAsync(dst, s->d_file, j.res.out_len, cudaMemcpyDeviceToHost, s->stream);
            if (e != cudaSuccess) { status = fail(h, ICB_ERR_CUDA, "PNG D2H", e); break; }
            h->d2h += j.res.out_len + 8;
            j.phase = PH_PNG_FETCH;
            continue;
        }
        if (s->job.phase == PH_PNG_FETCH) {
            if (s->job.stage_stream) memcpy(s->job.out, s->h_out, s->job.res.out_len);
            status = ICB_OK;
            break;
        }
        int rc = fit_advance(h, s);
        if (rc == 0) {
            status = s->job.res.status;
            if (s->job.from_file) {                  // the pixels the fit ran on: did the GPU reader decode them?
                const int dv = decode_verdict(h, s);
                if (dv) { status = dv; s->job.res.out_len = 0; }
            }
            break;
        }
        if (rc < 0) { status = rc; cudaStreamSynchronize(s->stream); break; }
    }
    s->job.res.status = status;
    s->job.phase = PH_DONE;
    return true;
}

Slot *slot_of_ticket(icb_handle *h, int64_t ticket)
{
    const size_t idx = (size_t)(ticket & 0xffff);
    if (idx >= h->slots.size()) return nullptr;
    Slot *s = &h->slots[idx];
    std::lock_guard<std::mutex> lk(h->mu);
    if (!s->busy || make_ticket(s) != ticket) return nullptr;
    return s;
}

}  // namespace

extern "C" {

int icb_submit_resize_image(icb_handle *h, const icb_image *img, int dw, int dh, uint8_t *dst, int dpitch,
                            int64_t *ticket)
{
    if (!h || !img || !ticket || dw < 1 || dh < 1) return ICB_ERR_INVALID;
    const int bpp = bpp_of(img->fmt);
    if (dst && dpitch < dw * bpp) return fail(h, ICB_ERR_INVALID, "dpitch smaller than a row");
    Slot *s = acquire_slot(h, img->devi);
    if (!s) return ICB_BUSY_INTERNAL;
    int rc = cudaSetDevice(h->devs[s->devi].dev) == cudaSuccess ? ICB_OK : fail(h, ICB_ERR_CUDA, "cudaSetDevice");
    if (!rc) rc = start_resize(h, s, img->d_px, img->w, img->h, img->pitch, img->fmt, dw, dh, dst, dpitch);
    if (rc) { cudaStreamSynchronize(s->stream); release_slot(h, s); return rc; }
    *ticket = make_ticket(s);
    return ICB_OK;
}

int icb_submit_resize(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt, uint8_t *dst, int dw, int dh,
                      int dpitch, int64_t *ticket)
{
    int rc = check_resize_args(h, src, w, hgt, pitch, fmt, dst, dw, dh, dpitch);
    if (rc) return rc;
    if (!ticket) return ICB_ERR_INVALID;
    const int bpp = bpp_of(fmt);
    Slot *s = acquire_slot(h);
    if (!s) return ICB_BUSY_INTERNAL;
    const int sp = (int)round_up((size_t)w * bpp, 16);
    if (cudaSetDevice(h->devs[s->devi].dev) != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "cudaSetDevice");
    if (!rc) rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)sp * hgt + 16);
    if (!rc) rc = upload(h, s, s->d_rgb, sp, src, w, hgt, pitch, bpp);
    if (!rc) rc = start_resize(h, s, s->d_rgb, w, hgt, sp, fmt, dw, dh, dst, dpitch);
    if (rc) { cudaStreamSynchronize(s->stream); release_slot(h, s); return rc; }
    *ticket = make_ticket(s);
    return ICB_OK;
}

int icb_submit_png_fit(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt, int min_w, int min_h,
                       uint8_t *dst, size_t dst_cap, int *dw, int *dh, int64_t *ticket)
{
    if (!h) return ICB_ERR_INVALID;
    if (!src || !dst || !dw || !dh || !ticket || w < 1 || hgt < 1) return fail(h, ICB_ERR_INVALID, "null argument");   // :39-41 NPE
    if (!fmt_valid(fmt)) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    const int rc0 = icb_png_target_dims(w, hgt, min_w, min_h, dw, dh, nullptr);
    if (rc0) return rc0;                                            // ICB_NO_FIT: nothing queued, no ticket
    const int bpp = bpp_of(fmt);
    if (dst_cap < (size_t)*dw * *dh * bpp) return fail(h, ICB_ERR_CAPACITY, "dst too small");
    return icb_submit_resize(h, src, w, hgt, pitch, fmt, dst, *dw, *dh, *dw * bpp, ticket);
}

int icb_submit_png_encode(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt, uint8_t *out,
                          size_t out_cap, int64_t *ticket)
{
    if (!h) return ICB_ERR_INVALID;
    if (!px || !out || !ticket || w < 1 || hgt < 1) return fail(h, ICB_ERR_INVALID, "null argument");
    if (!fmt_valid(fmt)) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    const int bpp = bpp_of(fmt);
    if (pitch < w * bpp) return fail(h, ICB_ERR_INVALID, "pitch smaller than a row");
    Slot *s = acquire_slot(h);
    if (!s) return ICB_BUSY_INTERNAL;
    const int sp = (int)round_up((size_t)w * bpp, 16);
    int rc = cudaSetDevice(h->devs[s->devi].dev) == cudaSuccess ? ICB_OK : fail(h, ICB_ERR_CUDA, "cudaSetDevice");
    if (!rc) rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)sp * hgt + 16);
    if (!rc) rc = upload(h, s, s->d_rgb, sp, px, w, hgt, pitch, bpp);
    if (!rc) {
        s->job = Job();
        s->job.out = out; s->job.out_cap = out_cap; s->job.out_pinned = is_pinned(out);
        rc = enqueue_png(h, s, s->d_rgb, w, hgt, sp, fmt);
    }
    if (rc) { cudaStreamSynchronize(s->stream); release_slot(h, s); return rc; }
    *ticket = make_ticket(s);
    return ICB_OK;
}

int icb_png_encode(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt, uint8_t *out, size_t out_cap,
                   size_t *out_len)
{
    int64_t t = 0;
    int rc = icb_submit_png_encode(h, px, w, hgt, pitch, fmt, out, out_cap, &t);
    if (rc) return rc;
    icb_result r;
    rc = icb_wait(h, t, &r);
    if (out_len) *out_len = r.out_len;
    return rc;
}

int icb_submit_png_compress(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt, int min_w, int min_h,
                            uint8_t *out, size_t out_cap, int *dw, int *dh, int64_t *ticket)
{
    if (!h) return ICB_ERR_INVALID;
    if (!src || !out || !dw || !dh || !ticket || w < 1 || hgt < 1) return fail(h, ICB_ERR_INVALID, "null argument");   // :39-41 NPE
    if (!fmt_valid(fmt)) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    const int bpp = bpp_of(fmt);
    if (pitch < w * bpp) return fail(h, ICB_ERR_INVALID, "pitch smaller than a row");
    const int rc0 = icb_png_target_dims(w, hgt, min_w, min_h, dw, dh, nullptr);
    if (rc0) return rc0;                                            // ICB_NO_FIT: the seam returns false, nothing queued
    Slot *s = acquire_slot(h);
    if (!s) return ICB_BUSY_INTERNAL;
    const int sp = (int)round_up((size_t)w * bpp, 16), np = (int)round_up((size_t)*dw * bpp, 16);
    int rc = cudaSetDevice(h->devs[s->devi].dev) == cudaSuccess ? ICB_OK : fail(h, ICB_ERR_CUDA, "cudaSetDevice");
    if (!rc) rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)sp * hgt + 16);
    if (!rc) rc = ensure(h, s->d_rgb2, s->rgb2_cap, (size_t)np * *dh + 16);
    if (!rc) rc = upload(h, s, s->d_rgb, sp, src, w, hgt, pitch, bpp);
    if (!rc) rc = enqueue_resize(h, s, s->d_rgb, w, hgt, sp, fmt, s->d_rgb2, *dw, *dh, np);      // :66, pixels stay in HBM
    if (!rc) {
        s->job = Job();
        s->job.out = out; s->job.out_cap = out_cap; s->job.out_pinned = is_pinned(out);
        rc = enqueue_png(h, s, s->d_rgb2, *dw, *dh, np, fmt);                                    // :70
    }
    if (rc) { cudaStreamSynchronize(s->stream); release_slot(h, s); return rc; }
    *ticket = make_ticket(s);
    return ICB_OK;
}

int icb_png_compress(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt, int min_w, int min_h,
                     uint8_t *out, size_t out_cap, size_t *out_len, int *dw, int *dh)
{
    int64_t t = 0;
    int rc = icb_submit_png_compress(h, src, w, hgt, pitch, fmt, min_w, min_h, out, out_cap, dw, dh, &t);
    if (rc) return rc;
    icb_result r;
    rc = icb_wait(h, t, &r);
    if (out_len) *out_len = r.out_len;
    return rc;
}

int icb_poll(icb_handle *h, int64_t ticket, int *done)
{
    if (!h || !done) return ICB_ERR_INVALID;
    Slot *s = slot_of_ticket(h, ticket);
    if (!s) return ICB_ERR_TICKET;
    *done = drive_job(h, s, false) ? 1 : 0;
    return ICB_OK;
}

int icb_wait(icb_handle *h, int64_t ticket, icb_result *res)
{
    if (!h) return ICB_ERR_INVALID;
    Slot *s = slot_of_ticket(h, ticket);
    if (!s) return ICB_ERR_TICKET;
    drive_job(h, s, true);
    const int status = s->job.res.status;
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
int stage_and_transform(icb_handle *h, Slot *s, const uint8_t *px, int w, int hgt, int pitch, int fmt,
                        const FitPlan *plan)
{
    CK(cudaSetDevice(h->devs[s->devi].dev));
    const int dp = dev_pitch(w, 3);
    int rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)dp * hgt + 16);
    if (!rc) rc = upload(h, s, s->d_rgb, dp, px, w, hgt, pitch, 3);
    if (!rc) rc = enqueue_transform(h, s, s->d_rgb, w, hgt, dp, fmt, plan);
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
    if (bad_quality(q)) return fail(h, ICB_ERR_INVALID, "quality outside [0, 1]");     // ImageWriteParam.setCompressionQuality
    const FitPlan pl = make_plan(MODE_SINGLE, q, 0, 1, false);
    Slot *s = acquire_slot(h);
    if (!s) return ICB_BUSY_INTERNAL;
    rc = stage_and_transform(h, s, px, w, hgt, pitch, fmt, &pl);
    long long cap = 0;
    if (!rc) {
        cap = std::min<long long>(out ? (long long)out_cap : 0, worst_stream_bytes(s->geom));
        rc = ensure_out(h, s, cap);
    }
    if (!rc) rc = enqueue_passes(h, s, pl, true, true, cap);
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
    for (int i = 0; i < n; i++) if (bad_quality(qs[i])) return fail(h, ICB_ERR_INVALID, "quality outside [0, 1]");
    Slot *s = acquire_slot(h);
    if (!s) return ICB_BUSY_INTERNAL;
    rc = stage_and_transform(h, s, px, w, hgt, pitch, fmt, nullptr);
    for (int i = 0; i < n && !rc; i++) {        // one transform; every size from the coefficients in HBM
        rc = enqueue_passes(h, s, make_plan(MODE_SINGLE, qs[i], 0, 1, false), false, false, 0);
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
    if (!s) return ICB_BUSY_INTERNAL;
    rc = stage_and_transform(h, s, px, w, hgt, pitch, fmt, nullptr);
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
    int64_t t = 0;
    int rc = icb_submit_resize(h, src, w, hgt, pitch, fmt, dst, dw, dh, dpitch, &t);
    if (rc) return rc;
    return icb_wait(h, t, nullptr);
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
    if (!fmt_valid(fmt)) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    const int bpp = bpp_of(fmt);
    const int dw = (w + sfac - 1) / sfac, dh = (hgt + sfac - 1) / sfac;
    if (pitch < w * bpp || dpitch < dw * bpp) return fail(h, ICB_ERR_INVALID, "pitch smaller than a row");
    Slot *s = acquire_slot(h);
    if (!s) return ICB_BUSY_INTERNAL;
    int rc = ICB_OK;
    do {
        if (cudaSetDevice(h->devs[s->devi].dev) != cudaSuccess) { rc = ICB_ERR_CUDA; break; }
        const int sp = (int)round_up((size_t)w * bpp, 16), np = (int)round_up((size_t)dw * bpp, 16);
        rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)sp * hgt + 16);
        if (rc) break;
        rc = ensure(h, s->d_rgb2, s->rgb2_cap, (size_t)np * dh + 16);
        if (rc) break;
        rc = upload(h, s, s->d_rgb, sp, src, w, hgt, pitch, bpp);
        if (rc) break;
        dim3 grid((dw + 127) / 128, (dh + 7) / 8);
        if (bpp == 4) k_subsample<4><<<grid, 256, 0, s->stream>>>(s->d_rgb, sp, sfac, s->d_rgb2, dw, dh, np);
        else if (bpp == 1) k_subsample<1><<<grid, 256, 0, s->stream>>>(s->d_rgb, sp, sfac, s->d_rgb2, dw, dh, np);
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
    if (!fmt_valid(fmt)) return fail(h, ICB_ERR_UNSUPPORTED, "format");
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
    for (int i = 0; i < 3; i++) { out->kernel_clock_ns[i] = h->kclk_ns[i].load(); out->kernel_clock_launches[i] = h->kclk_n[i].load(); }
    return ICB_OK;
}

int icb_profile_transform_series(icb_handle *h, const icb_image *const *imgs, int n, int reps, float list_q,
                                 float *ms_per_launch, int *launches)
{
    const FitPlan pl = make_plan(MODE_SEARCH, list_q, 0, 1, false);
    const FitPlan *plan = list_q > 0.0f ? &pl : nullptr;
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
                rc = enqueue_transform(h, s, imgs[i]->d_px, imgs[i]->w, imgs[i]->h, imgs[i]->pitch, imgs[i]->fmt, plan);
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

int icb_profile_resize_series(icb_handle *h, const icb_image *img, int dw, int dh, int reps, int with_png,
                              float *ms_per_launch)
{
    if (!h || !img || dw < 1 || dh < 1 || reps < 1 || !ms_per_launch) return ICB_ERR_INVALID;
    Slot *s = acquire_slot(h, img->devi);
    if (!s) return ICB_BUSY_INTERNAL;
    int rc = ICB_OK;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    do {
        if (cudaSetDevice(h->devs[s->devi].dev) != cudaSuccess || cudaEventCreate(&e0) != cudaSuccess ||
            cudaEventCreate(&e1) != cudaSuccess) { rc = ICB_ERR_CUDA; break; }
        const int bpp = bpp_of(img->fmt);
        const int np = (int)round_up((size_t)dw * bpp, 16);
        rc = ensure(h, s->d_rgb2, s->rgb2_cap, (size_t)np * dh + 16);
        s->job = Job();
        for (int r = -2; r < reps && !rc; r++) {            // two untimed warm-up rounds
            if (r == 0) cudaEventRecord(e0, s->stream);
            rc = enqueue_resize(h, s, img->d_px, img->w, img->h, img->pitch, img->fmt, s->d_rgb2, dw, dh, np);
            if (!rc && with_png) rc = enqueue_png(h, s, s->d_rgb2, dw, dh, np, img->fmt);
        }
        if (rc) break;
        cudaEventRecord(e1, s->stream);
        cudaError_t e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) { rc = fail(h, ICB_ERR_CUDA, "profile", e); break; }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        *ms_per_launch = ms / (float)reps;
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
    if (!s) return ICB_BUSY_INTERNAL;
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
            const int npairs = max_search_iters(init_q);
            const FitPlan pl = make_plan(MODE_SEARCH, init_q, target, npairs,
                                         expect_first_pass_fails(h, img->w, img->h, init_q, target));
            cudaEventRecord(e0, s->stream);
            rc = enqueue_transform(h, s, img->d_px, img->w, img->h, img->pitch, img->fmt, &pl);
            if (rc) break;
            const uint64_t k1 = h->kernels.load();
            cudaEventRecord(e1, s->stream);
            const long long cap = std::min<long long>(target, worst_stream_bytes(s->geom));
            rc = ensure_out(h, s, cap);
            if (rc) break;
            rc = enqueue_passes(h, s, pl, true, true, cap, pe);
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

// ======================================================================
//  C ABI, part 3: the batch executor (CompressionBatch.java:64-108) in native code
// ======================================================================
// Executors.newFixedThreadPool(n) + one task per file-list line becomes: `threads` worker threads of this
// library pulling items from ONE atomic cursor (work granularity = one image, so 12 MP / 24 MP / PNG
// mixes balance over the GPUs of the handle), each keeping `inflight` jobs in the per-GPU stream work
// queues -- upload of item k+1 overlaps the kernels of item k and the download of item k-1 on other slots.
#include <deque>
#include <thread>

namespace {

int submit_item(icb_handle *h, icb_item *it, int64_t *ticket)
{
    switch (it->kind) {
    case ICB_ITEM_JPG_FIT:
        return icb_submit_jpg_fit(h, it->px, it->w, it->h, it->pitch, it->fmt, it->init_q, it->target, it->has_cached,
                                  it->cached_q, it->cached_scale, it->out, it->out_cap, ticket);
    case ICB_ITEM_JPG_FIT_IMAGE:
        return icb_submit_jpg_fit_image(h, it->img, it->init_q, it->target, it->has_cached, it->cached_q,
                                        it->cached_scale, it->out, it->out_cap, ticket);
    case ICB_ITEM_PNG_FIT:
        return icb_submit_png_fit(h, it->px, it->w, it->h, it->pitch, it->fmt, it->min_w, it->min_h, it->out, it->out_cap,
                                  &it->dw, &it->dh, ticket);
    case ICB_ITEM_PNG_COMPRESS:
        return icb_submit_png_compress(h, it->px, it->w, it->h, it->pitch, it->fmt, it->min_w, it->min_h, it->out,
                                       it->out_cap, &it->dw, &it->dh, ticket);
    case ICB_ITEM_JPG_FILE_FIT:
        return icb_submit_jpg_file_fit(h, it->px, (size_t)(it->pitch > 0 ? it->pitch : 0), it->init_q, it->target,
                                       it->has_cached, it->cached_q, it->cached_scale, it->out, it->out_cap, &it->dw,
                                       &it->dh, ticket);
    case ICB_ITEM_RESIZE_IMAGE:
        if (!it->img) return ICB_ERR_INVALID;
        return icb_submit_resize_image(h, it->img, it->dw, it->dh, it->out, it->out ? it->dw * bpp_of(it->img->fmt) : 0,
                                       ticket);
    default:
        return ICB_ERR_INVALID;
    }
}

}  // namespace

extern "C" int icb_run_list(icb_handle *h, icb_item *items, size_t n, int threads, int inflight)
{
    if (!h || (!items && n)) return ICB_ERR_INVALID;
    threads = std::max(1, std::min(threads, 256));
    // a thread that holds every slot while it asks for one more would wait for itself: cap what is in flight
    const int slots = (int)h->slots.size();
    inflight = std::max(1, std::min(inflight, std::max(1, slots / threads)));
    std::atomic<size_t> cursor{0};
    auto work = [&]() {
        std::deque<std::pair<icb_item *, int64_t>> pending;
        auto finish = [&]() {
            icb_item *it = pending.front().first;
            const int64_t t = pending.front().second;
            pending.pop_front();
            icb_wait(h, t, &it->res);
        };
        while (true) {
            const size_t k = cursor.fetch_add(1, std::memory_order_relaxed);
            if (k >= n) break;
            icb_item *it = &items[k];
            if ((int)pending.size() == inflight) finish();
            int64_t t = 0;
            int rc;
            while (true) {
                tl_acquire_nowait = !pending.empty();       // holding tickets: never wait for a slot
                rc = submit_item(h, it, &t);
                tl_acquire_nowait = false;
                if (rc != ICB_BUSY_INTERNAL) break;
                finish();                                   // frees a slot (ours), then try again
            }
            if (rc != ICB_OK) {                     // ICB_NO_FIT from the PNG rule, or an error: nothing was queued
                it->res = icb_result{};
                it->res.status = rc;
                continue;
            }
            pending.emplace_back(it, t);
        }
        while (!pending.empty()) finish();
    };
    if (threads == 1) {
        work();
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; t++) pool.emplace_back(work);
        for (auto &t : pool) t.join();
    }
    return ICB_OK;
}

// ======================================================================
//  C ABI, part 4: a PNG encode that allocates nothing
// ======================================================================
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels and
// the same work-arena layout as enqueue_png above, but the arena is the file-scope array below instead of a
// slot's cudaMalloc'ed d_png, pixels and file are the caller's DEVICE buffers, and the kernels go on the
// caller's stream.  Every GPU that loads the library holds the array until exit.
namespace {

constexpr long long NOALLOC_MAX_FILT = 1 << 20;        // filtered bytes h * (1 + w * png_bpp): e.g. 512 x 511 RGBA
constexpr size_t noalloc_arena_bytes()
{
    const size_t np = (size_t)(NOALLOC_MAX_FILT / PNG_PIECE);
    size_t o = ((size_t)NOALLOC_MAX_FILT + 64 + 255) / 256 * 256;      // filt
    o += (np + 1) * PNG_PIECE_WORDS * 4;                                // seg_bits
    o = (o + (np + 1) * 4 + 7) / 8 * 8;                                 // seg_len
    o += (np + 2) * 8 + np * 8 + 320 * 4 + 320 * 4 + 64;                // seg_pos, adler, hist, codes, misc
    return (o + 255) / 256 * 256;
}
constexpr size_t NOALLOC_ARENA_BYTES = noalloc_arena_bytes();           // 5.3 MB

}  // namespace

// In host code this name is only a stand-in for the symbol, not a device address: the address the kernels
// get comes from cudaGetSymbolAddress, once per device (noalloc_arena below).
static __device__ __align__(256) unsigned char g_noalloc_arena[NOALLOC_ARENA_BYTES];

namespace {

std::mutex g_noalloc_mu;
unsigned char *g_noalloc_ptr[64] = {nullptr};           // per device ordinal: device address of g_noalloc_arena

// set-up, once per device, with that device current: the arena's device address and the CRC tables
// k_png_finish reads (the same values setup_device stores; a handle need not exist)
int noalloc_arena(icb_handle *h, int dev, unsigned char **out)
{
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_arena);
        if (e != cudaSuccess || !ptr) return fail(h, ICB_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_arena)", e);
        static uint32_t crc[256 + 4 * 256];
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            crc[i] = c;
        }
        for (int j = 0; j < 4; j++)
            for (uint32_t b = 0; b < 256; b++) {
                uint32_t c = b << (8 * j);
                for (int k = 0; k < PNG_CRC_SEG; k++) c = crc[c & 0xffu] ^ (c >> 8);
                crc[256 + 256 * j + b] = c;
            }
        e = cudaMemcpyToSymbol(g_crc, crc, sizeof(crc));
        if (e != cudaSuccess) return fail(h, ICB_ERR_CUDA, "cudaMemcpyToSymbol(g_crc)", e);
        g_noalloc_ptr[dev] = (unsigned char *)ptr;
    }
    *out = g_noalloc_ptr[dev];
    return ICB_OK;
}

}  // namespace

extern "C" int icb_png_encode_device_noalloc(icb_handle *h, const uint8_t *d_px, int w, int hgt, int pitch, int fmt,
                                             uint8_t *d_out, size_t out_cap, unsigned long long *d_out_len,
                                             void *stream)
{
    if (!d_px || !d_out || !d_out_len || w < 1 || hgt < 1) return fail(h, ICB_ERR_INVALID, "null argument");
    if (!fmt_valid(fmt)) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    const int bpp = bpp_of(fmt);
    if (pitch < w * bpp) return fail(h, ICB_ERR_INVALID, "pitch smaller than a row");
    PngParams pp = {};
    pp.src = d_px; pp.w = w; pp.h = hgt; pp.pitch = pitch; pp.sbpp = bpp;
    png_layout(fmt, pp.pbpp, pp.perm);
    pp.nfilt = (long long)hgt * (1 + (long long)w * pp.pbpp);
    if (pp.nfilt > NOALLOC_MAX_FILT) {
        char msg[160];
        snprintf(msg, sizeof msg, "icb_png_encode_device_noalloc: h * (1 + w * png_bpp) = %lld exceeds the static arena's "
                 "limit of %lld bytes", pp.nfilt, NOALLOC_MAX_FILT);
        return fail(h, ICB_ERR_CAPACITY, msg);
    }
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return fail(h, ICB_ERR_NO_DEVICE, "no current device");
    unsigned char *arena = nullptr;
    int rc = noalloc_arena(h, dev, &arena);
    if (rc) return rc;
    pp.npieces = (int)std::max<long long>(1, pp.nfilt / PNG_PIECE);
    const size_t np = (size_t)pp.npieces;
    const size_t o_seg = round_up((size_t)pp.nfilt + 64, 256);
    const size_t o_len = o_seg + (np + 1) * PNG_PIECE_WORDS * 4, o_pos = round_up(o_len + (np + 1) * 4, 8);
    const size_t o_adl = o_pos + (np + 2) * 8, o_hist = o_adl + np * 8, o_codes = o_hist + 320 * 4;
    const size_t o_misc = o_codes + 320 * 4;
    if (o_misc + 64 > NOALLOC_ARENA_BYTES) return fail(h, ICB_ERR_CAPACITY, "static arena layout");   // (cannot happen)
    pp.filt = arena;
    pp.seg_bits = reinterpret_cast<uint32_t *>(arena + o_seg);
    pp.seg_len = reinterpret_cast<unsigned int *>(arena + o_len);
    pp.seg_pos = reinterpret_cast<unsigned long long *>(arena + o_pos);
    pp.adler = reinterpret_cast<uint2 *>(arena + o_adl);
    pp.hist = reinterpret_cast<unsigned int *>(arena + o_hist);
    pp.codes = reinterpret_cast<uint32_t *>(arena + o_codes);
    pp.out_len = d_out_len;
    pp.adler_out = reinterpret_cast<unsigned int *>(arena + o_misc + 8);
    pp.out = d_out;
    pp.out_cap = (long long)out_cap;
    cudaStream_t st = (cudaStream_t)stream;
    // the arena keeps the last call's contents: the histogram must start at zero (the kernels write every
    // other field they read: k_png_tree / k_png_encode zero their bit-stream segments themselves)
    CK(cudaMemsetAsync(pp.hist, 0, 320 * 4, st));
    k_png_filter<<<hgt, 256, 0, st>>>(pp);
    k_png_hist<<<(pp.npieces + 7) / 8, 256, 0, st>>>(pp);
    k_png_tree<<<1, PNG_TREE_THREADS, 0, st>>>(pp);
    k_png_encode<<<(pp.npieces + 7) / 8, 256, 0, st>>>(pp);
    k_png_scan<<<1, 1024, 0, st>>>(pp);
    k_png_merge<<<(pp.npieces + 1 + 7) / 8, 256, 0, st>>>(pp);
    k_png_finish<<<(unsigned)((((long long)pp.nfilt * 2) / PNG_IDAT + 2 + 7) / 8), 256, 0, st>>>(pp);
    CK(cudaGetLastError());
    if (h) h->kernels += 7;
    return ICB_OK;
}

// ======================================================================
//  C ABI, part 5: the JPEG reader on the GPU (core/ImageCompression.java:119-155)
// ======================================================================
namespace {

static std::atomic<uint64_t> g_submit_ns{0}, g_submit_n{0}, g_acquire_ns{0}, g_graph_ns{0}, g_fit_ns{0}, g_dec_ns{0}, g_t1{0}, g_t2{0}, g_t3{0}, g_t4{0};
static inline uint64_t now_ns() { return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int map_parse(icb_handle *h, int jp)
{
    if (jp == JP_OK) return ICB_OK;
    if (jp == JP_UNSUPPORTED)
        return fail(h, ICB_ERR_UNSUPPORTED, "JPEG reader takes baseline 8-bit YCbCr 4:2:0 / 4:2:2 / 4:4:4 in one scan");
    return fail(h, ICB_ERR_INVALID, "not a JPEG file / truncated header");
}

// file bytes (host) -> decoded raster in d_out (pitch out_pitch, RGB8 or BGR8), all queued on the slot's stream
// The decode chain of one slot as a CUDA graph.  Every kernel reads its parameters from a block in device memory
// (DecJob: uploaded per job together with the file's tables), grids and memset sizes follow the slot's CAPACITIES
// (longest scan / largest image seen so far, rounded up; surplus CTAs leave at once), so the graph is captured once
// per slot and replayed: a job costs one upload of the block, one of the file, one graph launch -- 3 driver calls
// instead of 30 (measured r02: ~200 us of serialised driver time per file job with plain launches).
int dec_rounds()
{
    static const int v = [] { const char *e = getenv("ICB_DEC_ROUNDS"); const int x = e ? atoi(e) : DEC_ROUNDS; return x >= 2 && x <= DEC_ROUNDS ? x : DEC_ROUNDS; }();
    return v;
}

void launch_decode_chain(Slot *s, cudaStream_t st, size_t sub_bytes_cap, size_t nblocks_cap, int nchunks_cap, unsigned int nsub_cap,
                         int w_cap, int h_cap, int hs, int vs)
{
    const DecParams *pp = &s->d_decjob->p;
    // A/B knob (timing only, results wrong): leave kernels out -- 1 zero, 2 scan/unstuff, 4 sync rounds, 8 write, 16 DC, 32 IDCT, 64 colour
    static const int skip = [] { const char *e = getenv("ICB_DEC_SKIP"); return e ? atoi(e) : 0; }();
    // state: everything zero, the scan ends at the end of the file unless a marker turns up earlier; no record yet
    cudaMemsetAsync(s->d_decst, 0, sizeof(DecState), st);
    cudaMemsetAsync(&s->d_decst->scan_end, 0xff, 4, st);
    static const bool zero_by_memset = [] { const char *e = getenv("ICB_DEC_MEMSET"); return e && e[0] == '1'; }();     // A/B knob
    static const bool zero_up_front = [] { const char *e = getenv("ICB_DEC_ZERO"); return e && e[0] == '1'; }();      // A/B: the round-2 first form
    if ((skip & 1) || !zero_up_front) { }
    else if (zero_by_memset) cudaMemsetAsync(s->d_coef, 0, nblocks_cap * 128, st);
    else k_dec_zero<<<1184, 256, 0, st>>>(reinterpret_cast<uint4 *>(s->d_coef), nblocks_cap * 8);
    cudaMemsetAsync(s->d_sub, 0xff, sub_bytes_cap, st);
    if (!(skip & 2)) {
        k_dec_scan<<<nchunks_cap, 256, 0, st>>>(pp);
        k_dec_offsets<<<1, 1024, 0, st>>>(pp);
        k_dec_unstuff<<<nchunks_cap, 256, 0, st>>>(pp);
    }
    const unsigned int sgrid = (nsub_cap + DEC_THREADS - 1) / DEC_THREADS;
    const int nrounds = dec_rounds();
    // A/B knob: unused dynamic shared memory per CTA of the two symbol-walking kernels, to cap how many of them an SM takes
    static const size_t pad = [] { const char *e = getenv("ICB_DEC_PAD_KB"); const int v = e ? atoi(e) : 0; return (size_t)(v > 0 && v <= 150 ? v : 0) << 10; }();
    static const bool pad_set = [&] {
        if (pad) {
            cudaFuncSetAttribute(k_dec_sync, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pad);
        }
        return true;
    }();
    (void)pad_set;
    for (int r = 0; r < nrounds && !(skip & 4); r++) k_dec_sync<<<sgrid, DEC_THREADS, pad, st>>>(pp, r == 0 ? 0 : DEC_ROUNDS - nrounds + r);
    k_dec_blocks<<<1, 1024, 0, st>>>(pp);
    if (!(skip & 8)) k_dec_write<<<(nsub_cap + DEC_WTHREADS - 1) / DEC_WTHREADS, DEC_WTHREADS, DEC_WBUF_BYTES, st>>>(pp);
    const size_t ndc = (nblocks_cap / 3 + DEC_DC_THREADS - 1) / DEC_DC_THREADS;          // (an MCU has at least 3 blocks)
    if (!(skip & 16)) {
        k_dec_dc_local<<<(unsigned int)ndc, DEC_DC_THREADS, 0, st>>>(pp);
        k_dec_dc_cta<<<1, 1024, 0, st>>>(pp);
    }
    if (!(skip & 32)) k_dec_idct<<<(unsigned int)((nblocks_cap + 127) / 128), 128, 0, st>>>(pp);
    const dim3 cgrid((w_cap + 255) / 256, (h_cap + 7) / 8);
    if (skip & 64) { }
    else if (hs == 2 && vs == 2) k_dec_color<2, 2><<<cgrid, 256, 0, st>>>(pp);
    else if (hs == 2) k_dec_color<2, 1><<<cgrid, 256, 0, st>>>(pp);
    else k_dec_color<1, 1><<<cgrid, 256, 0, st>>>(pp);
}

// file bytes (host) -> decoded raster in d_out (pitch out_pitch, RGB8 or BGR8), all queued on the slot's stream
int enqueue_decode(icb_handle *h, Slot *s, const uint8_t *file, size_t n, const JpegHeader &hd, unsigned char *d_out,
                   int out_pitch, bool bgr)
{
    cudaStream_t st = s->stream;
    const int nbm = hd.hs * hd.vs + 2;                                          // blocks per MCU
    const int mcu_cols = (hd.w + 8 * hd.hs - 1) / (8 * hd.hs), mcu_rows = (hd.h + 8 * hd.vs - 1) / (8 * hd.vs);
    const size_t nblocks = (size_t)mcu_cols * mcu_rows * nbm;
    const size_t scan_bytes = n - hd.scan_off;
    static const unsigned int sub_bits = [] {
        const char *e = getenv("ICB_DEC_SUB");
        const int v = e ? atoi(e) : DEC_SUB_BITS;
        return (unsigned int)(v >= 256 && v <= 8192 && v % 32 == 0 ? v : DEC_SUB_BITS);   // (12 bits of block count: >= 4 bits a block)
    }();
    static const unsigned int carry = [] { const char *e = getenv("ICB_DEC_CARRY"); const int v = e ? atoi(e) : DEC_CARRY; return (unsigned int)(v >= 1 && v <= 64 ? v : DEC_CARRY); }();
    static const bool use_graph = [] { const char *e = getenv("ICB_DEC_GRAPH"); return !(e && e[0] == '0'); }();
    if ((scan_bytes * 8 + sub_bits - 1) / sub_bits >= (1u << 17))
        return fail(h, ICB_ERR_UNSUPPORTED, "JPEG reader: scan too long for the subsequence records");
    // capacities of the slot: they only grow, in steps (1 MB of scan, 256 blocks of 16 pixels), so that slots settle
    const size_t scan_cap = std::max(s->dec_scan_cap, round_up(scan_bytes, (size_t)1 << 20));
    const int wc = std::max(s->dec_w_cap, (int)round_up((size_t)hd.w, 256)), hc = std::max(s->dec_h_cap, (int)round_up((size_t)hd.h, 256));
    const size_t nblocks_cap = std::max(s->dec_blocks_cap,
                                        (size_t)((wc + 8 * hd.hs - 1) / (8 * hd.hs)) * ((hc + 8 * hd.vs - 1) / (8 * hd.vs)) * nbm);
    const unsigned int nsub_cap = (unsigned int)std::min<size_t>((scan_cap * 8 + sub_bits - 1) / sub_bits, (1u << 17) - 1);
    const int nchunks_cap = (int)((scan_cap + DEC_CHUNK - 1) / DEC_CHUNK) + 1;     // (+1: chunks start at scan_off rounded down to 16)
    if (!s->d_decjob) {
        if (cudaMalloc((void **)&s->d_decjob, sizeof(DecJob)) != cudaSuccess ||
            cudaMalloc((void **)&s->d_decst, sizeof(DecState)) != cudaSuccess ||
            cudaMallocHost((void **)&s->h_decjob, sizeof(DecJob)) != cudaSuccess ||
            cudaMallocHost((void **)&s->h_decst, sizeof(DecState)) != cudaSuccess)
            return fail(h, ICB_ERR_OOM, "decode tables");
    }
    int rc = ensure(h, s->d_file_in, s->file_in_cap, round_up(n, (size_t)1 << 20) + 256);
    if (!rc) rc = ensure(h, s->d_bits, s->bits_cap, scan_cap + 512);
    const size_t sub_bytes = round_up((size_t)nsub_cap * 8, 256);
    const size_t nmcu_cap = nblocks_cap / 3, ndc_cap = (nmcu_cap + DEC_DC_THREADS - 1) / DEC_DC_THREADS;
    const size_t o_base = sub_bytes, o_chunk = o_base + round_up((size_t)nsub_cap * 4, 256);
    const size_t o_mcu = o_chunk + round_up((size_t)(nchunks_cap + 1) * 4, 256), o_cta = o_mcu + round_up(nmcu_cap * 12, 256);
    const size_t o_crst = o_cta + round_up(ndc_cap * 12, 256), o_rst = o_crst + round_up((size_t)(nchunks_cap + 1) * 4, 256);
    if (!rc) rc = ensure(h, s->d_sub, s->sub_cap, o_rst + round_up(nmcu_cap * 4, 256));
    if (!rc) rc = ensure(h, s->d_planes, s->planes_cap, nblocks_cap * 64 + 256);
    if (!rc) rc = ensure(h, s->d_coef, s->coef_cap, nblocks_cap * 128);
    if (rc) return rc;
    s->dec_scan_cap = scan_cap; s->dec_w_cap = wc; s->dec_h_cap = hc; s->dec_blocks_cap = nblocks_cap;

    const uint64_t ta = now_ns();
    DecJob &job = *s->h_decjob;
    jpeg_build_tables(hd, job.t);
    g_t1 += now_ns() - ta;
    DecParams &dp = job.p;
    dp = DecParams{};
    dp.file = s->d_file_in; dp.file_len = (unsigned int)n; dp.scan_off = (unsigned int)hd.scan_off;
    dp.bits = s->d_bits;
    dp.rec = reinterpret_cast<unsigned long long *>(s->d_sub);
    dp.sub_bits = sub_bits;
    dp.carry = carry;
    static const int wbuf = [] { const char *e = getenv("ICB_DEC_WBUF"); return e && e[0] == '0' ? 0 : 1; }();
    dp.wbuf = wbuf;
    dp.blk_base = reinterpret_cast<unsigned int *>(s->d_sub + o_base);
    dp.chunk_drop = reinterpret_cast<unsigned int *>(s->d_sub + o_chunk);
    dp.mcu_pref = reinterpret_cast<int *>(s->d_sub + o_mcu);
    dp.cta_pref = reinterpret_cast<int *>(s->d_sub + o_cta);
    dp.chunk_rst = reinterpret_cast<unsigned int *>(s->d_sub + o_crst);
    dp.rst_pos = reinterpret_cast<unsigned int *>(s->d_sub + o_rst);
    dp.rst_cap = (unsigned int)nmcu_cap;
    dp.ri = hd.ri;
    dp.coef = s->d_coef;
    dp.planes = s->d_planes;
    dp.rgb = d_out; dp.rgb_pitch = out_pitch; dp.bgr = bgr ? 1 : 0;
    dp.w = hd.w; dp.h = hd.h; dp.mcu_cols = mcu_cols; dp.mcu_rows = mcu_rows;
    dp.hs = hd.hs; dp.vs = hd.vs; dp.nb = nbm;
    dp.nblocks = (unsigned int)nblocks;
    dp.nsub_max = (unsigned int)((scan_bytes * 8 + sub_bits - 1) / sub_bits);
    dp.nchunks = (int)((scan_bytes + (hd.scan_off & 15) + DEC_CHUNK - 1) / DEC_CHUNK);
    dp.tab = &s->d_decjob->t;
    dp.st = s->d_decst;
    const uint64_t tb = now_ns();
    CK(cudaMemcpyAsync(s->d_decjob, s->h_decjob, sizeof(DecJob), cudaMemcpyHostToDevice, st));
    const uint64_t tc = now_ns();
    g_t2 += tc - tb;
    rc = upload(h, s, s->d_file_in, (int)std::min<size_t>(n, 0x7fffffff), file, (int)n, 1, (int)n, 1);
    if (rc) return rc;
    g_t3 += now_ns() - tc;

    if (use_graph && !s->dec_graph_broken) {
        // the graph stands for these buffers and capacities; anything else: capture again
        const uint64_t key = ((uint64_t)(uintptr_t)s->d_coef * 0x9E3779B97F4A7C15ull) ^ ((uint64_t)(uintptr_t)s->d_sub * 0xC2B2AE3D27D4EB4Full) ^
                             ((uint64_t)(uintptr_t)s->d_decst << 7) ^ ((uint64_t)scan_cap << 3) ^ ((uint64_t)wc << 40) ^ ((uint64_t)hc << 20) ^
                             (uint64_t)dec_rounds() ^ ((uint64_t)(hd.hs * 4 + hd.vs) << 60) ^ ((uint64_t)nblocks_cap << 9);
        if (!s->dec_graph || s->dec_graph_key != key) {
            if (s->dec_graph) { cudaGraphExecDestroy(s->dec_graph); s->dec_graph = nullptr; }
            cudaGraph_t g = nullptr;
            CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            launch_decode_chain(s, st, sub_bytes, nblocks_cap, nchunks_cap, nsub_cap, wc, hc, hd.hs, hd.vs);
            cudaError_t e = cudaStreamEndCapture(st, &g);
            if (e == cudaSuccess && g) {
                e = cudaGraphInstantiate(&s->dec_graph, g, 0);
                cudaGraphDestroy(g);
            }
            if (e != cudaSuccess || !s->dec_graph) {         // a driver that cannot capture this: plain launches from now on
                cudaGetLastError();
                s->dec_graph = nullptr;
                s->dec_graph_broken = true;
            }
            s->dec_graph_key = key;
        }
    }
    if (use_graph && !s->dec_graph_broken && s->dec_graph) {
        const uint64_t tg = now_ns();
        CK(cudaGraphLaunch(s->dec_graph, st));
        g_graph_ns += now_ns() - tg;
    } else {
        launch_decode_chain(s, st, sub_bytes, nblocks_cap, nchunks_cap, nsub_cap, wc, hc, hd.hs, hd.vs);
        CK(cudaGetLastError());
    }
    h->kernels += 9 + dec_rounds();
    const uint64_t te = now_ns();
    CK(cudaMemcpyAsync(s->h_decst, s->d_decst, sizeof(DecState), cudaMemcpyDeviceToHost, st));
    g_t4 += now_ns() - te;
    s->job.dec_nblocks = (unsigned int)nblocks;
    s->job.dec_nrst = hd.ri > 0 ? (unsigned int)(((size_t)mcu_cols * mcu_rows + hd.ri - 1) / hd.ri - 1) : 0u;
    return ICB_OK;
}

}  // namespace

extern "C" {

int icb_jpg_file_info(const uint8_t *file, size_t n, int *w, int *hgt)
{
    JpegHeader hd;
    const int jp = jpeg_parse(file, n, hd);
    if (jp) return jp == JP_UNSUPPORTED ? ICB_ERR_UNSUPPORTED : ICB_ERR_INVALID;
    if (w) *w = hd.w;
    if (hgt) *hgt = hd.h;
    return ICB_OK;
}

int icb_jpg_decode(icb_handle *h, const uint8_t *file, size_t n, int fmt, uint8_t *dst, size_t dst_cap, int dpitch,
                   int *w, int *hgt)
{
    if (!h) return ICB_ERR_INVALID;
    if (!file || !dst || n > 0x7ffffff0u) return fail(h, ICB_ERR_INVALID, "bad arguments");
    if (fmt != ICB_FMT_RGB8 && fmt != ICB_FMT_BGR8) return fail(h, ICB_ERR_UNSUPPORTED, "format");
    JpegHeader hd;
    int rc = map_parse(h, jpeg_parse(file, n, hd));
    if (rc) return rc;
    if (w) *w = hd.w;
    if (hgt) *hgt = hd.h;
    if (dpitch < hd.w * 3 || dst_cap < (size_t)dpitch * (hd.h - 1) + (size_t)hd.w * 3) return fail(h, ICB_ERR_CAPACITY, "dst too small");
    Slot *s = acquire_slot(h);
    if (!s) return ICB_BUSY_INTERNAL;
    const int dp = dev_pitch(hd.w, 3);
    if (cudaSetDevice(h->devs[s->devi].dev) != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "cudaSetDevice");
    s->job = Job();
    if (!rc) rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)dp * hd.h + 16);
    if (!rc) rc = enqueue_decode(h, s, file, n, hd, s->d_rgb, dp, fmt == ICB_FMT_BGR8);
    if (!rc) {
        cudaError_t e = cudaMemcpy2DAsync(dst, dpitch, s->d_rgb, dp, (size_t)hd.w * 3, hd.h, cudaMemcpyDeviceToHost, s->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "decode", e);
        h->d2h += (uint64_t)hd.w * 3 * hd.h;
    } else {
        cudaStreamSynchronize(s->stream);
    }
    if (!rc) rc = decode_verdict(h, s);
    h->images_done++;
    release_slot(h, s);
    return rc;
}

struct SubmitTimer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    ~SubmitTimer()
    {
        g_submit_ns += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
        const uint64_t k = ++g_submit_n;
        if (getenv("ICB_DEC_DEBUG") && k % 256 == 0) {
            fprintf(stderr, "icb submit_jpg_file_fit: %.1f us per call on the host (%.1f us of it waiting for a slot; decode enqueue %.1f of which "
                    "graph launch %.1f; fit enqueue %.1f)\n", g_submit_ns.load() / 1e3 / k, g_acquire_ns.load() / 1e3 / k,
                    g_dec_ns.load() / 1e3 / k, g_graph_ns.load() / 1e3 / k, g_fit_ns.load() / 1e3 / k);
            fprintf(stderr, "   tables %.1f, block upload %.1f, file upload %.1f, state download %.1f us\n", g_t1.load() / 1e3 / k,
                    g_t2.load() / 1e3 / k, g_t3.load() / 1e3 / k, g_t4.load() / 1e3 / k);
        }
    }
};

int icb_submit_jpg_file_fit(icb_handle *h, const uint8_t *file, size_t n, float init_q, int64_t target, int has_cached,
                            float cached_q, double cached_scale, uint8_t *out, size_t out_cap, int *w, int *hgt,
                            int64_t *ticket)
{
    SubmitTimer timer_;
    if (!h) return ICB_ERR_INVALID;
    if (!file || !ticket || n > 0x7ffffff0u) return fail(h, ICB_ERR_INVALID, "bad arguments");
    int rc = check_fit_qualities(h, init_q, has_cached, cached_q);
    if (rc) return rc;
    JpegHeader hd;
    rc = map_parse(h, jpeg_parse(file, n, hd));
    if (rc) return rc;
    // core/ImageCompression.java:140-153: the decode rule -- what the seam sees is the subsampled raster
    const int sf = icb_subsample_factor(hd.w, hd.h);
    const int fw = (hd.w + sf - 1) / sf, fh = (hd.h + sf - 1) / sf;
    if (w) *w = fw;
    if (hgt) *hgt = fh;
    const auto ta = std::chrono::steady_clock::now();
    Slot *s = acquire_slot(h);
    if (!s) return ICB_BUSY_INTERNAL;
    g_acquire_ns += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - ta).count();
    if (cudaSetDevice(h->devs[s->devi].dev) != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "cudaSetDevice");
    s->job = Job();
    const int fp = dev_pitch(fw, 3);
    if (!rc) rc = ensure(h, s->d_rgb, s->rgb_cap, (size_t)fp * fh + 16);
    unsigned int nblk = 0;
    const uint64_t td = now_ns();
    if (!rc && sf == 1) {
        rc = enqueue_decode(h, s, file, n, hd, s->d_rgb, fp, false);
    } else if (!rc) {
        const int dp = dev_pitch(hd.w, 3);
        rc = ensure(h, s->d_dec, s->dec_cap, (size_t)dp * hd.h + 16);
        if (!rc) rc = enqueue_decode(h, s, file, n, hd, s->d_dec, dp, false);
        if (!rc) {
            k_subsample<3><<<dim3((fw + 127) / 128, (fh + 7) / 8), 256, 0, s->stream>>>(s->d_dec, dp, sf, s->d_rgb, fw, fh, fp);
            h->kernels++;
            if (cudaGetLastError() != cudaSuccess) rc = fail(h, ICB_ERR_CUDA, "subsample");
        }
    }
    nblk = s->job.dec_nblocks;
    const unsigned int nrst = s->job.dec_nrst;
    const uint64_t tf = now_ns();
    g_dec_ns += tf - td;
    static const bool dec_only = [] { const char *e = getenv("ICB_DEC_ONLY"); return e && e[0] == '1'; }();     // A/B: the reader alone
    if (dec_only && !rc) {
        s->job.phase = PH_RESIZE;                   // (a job with nothing to collect: drive_job just waits for the stream)
        s->job.from_file = false;
        *ticket = make_ticket(s);
        return ICB_OK;
    }
    if (!rc) rc = start_fit(h, s, s->d_rgb, fw, fh, fp, ICB_FMT_RGB8, init_q, target, has_cached, cached_q, cached_scale, out, out_cap);
    g_fit_ns += now_ns() - tf;
    if (rc) { cudaStreamSynchronize(s->stream); release_slot(h, s); return rc; }
    s->job.from_file = true;
    s->job.dec_nblocks = nblk;
    s->job.dec_nrst = nrst;
    *ticket = make_ticket(s);
    return ICB_OK;
}

int icb_jpg_file_fit(icb_handle *h, const uint8_t *file, size_t n, float init_q, int64_t target, int has_cached,
                     float cached_q, double cached_scale, uint8_t *out, size_t out_cap, size_t *out_len, float *best_q,
                     double *scale, int *used_cache, int *w, int *hgt)
{
    int64_t t = 0;
    int rc = icb_submit_jpg_file_fit(h, file, n, init_q, target, has_cached, cached_q, cached_scale, out, out_cap, w, hgt, &t);
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
