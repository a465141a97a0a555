// common.cuh -- shared constants and device-side state for the sm_100a JPEG path.
//
// Arithmetic follows IJG libjpeg 6b as the JDK's JPEGImageWriter runs it
// (reference call site: core/ImageCompressionJpg.java:137-143); section tags
// a2.x refer to SURVEY.md section 8(a).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace icb {

__device__ __forceinline__ unsigned long long gtime()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ---- programmatic dependent launch (sm_90+): the kernels of one fit form a chain on one stream;
// each lets its successor start early (launch latency and the successor's table set-up overlap
// the predecessor's tail) and waits for its predecessor's results only where it first needs them
__device__ __forceinline__ void pdl_launch_dependents()
{
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait()
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// ---- transform kernel tiling ----
#ifndef ICB_TILE_MCUS
#define ICB_TILE_MCUS 21
#endif
// MCUs per tile (one MCU row); 48 * TILE_MCUS <= 1024 (TMA box of <= 256 words).  21 MCUs = 126 blocks: the DCT stage
// (one thread per block) keeps 126 of 128 threads busy, and a tile's blocks make a survivor-list unit of the size the
// entropy pass likes (measured r02: transform 23.4 -> 22.7 us at 12 MP against 16-MCU tiles)
constexpr int TILE_MCUS = ICB_TILE_MCUS;
constexpr int TILE_ROW_BYTES = TILE_MCUS * 48;      // bytes of RGB per pixel row of a tile
constexpr int TILE_IN_BYTES = TILE_ROW_BYTES * 16;  // one TMA box
constexpr int TILE_BLOCKS = TILE_MCUS * 6;          // blocks per tile
constexpr int XF_THREADS = 128;
#ifndef ICB_XF_STAGES
#define ICB_XF_STAGES 2
#endif
#ifndef ICB_XF_CTAS
#define ICB_XF_CTAS 5
#endif
constexpr int XF_STAGES = ICB_XF_STAGES;            // TMA boxes in flight per CTA
constexpr int XF_CTAS = ICB_XF_CTAS;                // resident CTAs per SM the kernel is compiled for (register cap)

// ---- survivor-list units and entropy tiles ----
// A UNIT is the run of blocks one transform tile produces (TILE_MCUS MCUs of one MCU row, fewer at the
// right edge): consecutive in scan order, and the transform kernel writes its survivor list while the
// coefficients are still on chip.  An entropy TILE (what one warp codes per pass) is unit_group
// consecutive units.  Every tile costs a fixed overhead per pass (window set-up, final flush, 0xFF
// fold, tile record), so tiles should be large; but a pass needs a couple of thousand tiles to fill
// 148 SMs: two units per tile once that still leaves about 2000 tiles.  The per-tile 0xFF counters
// are 16 bits wide: 2 * TILE_BLOCKS * 208 bytes must stay below 65536.
constexpr int UNIT_BLOCKS = TILE_BLOCKS;
constexpr int UNIT_CAP = UNIT_BLOCKS * 64;          // survivor entries per unit, worst case (a multiple of 64)
constexpr int MAX_UNIT_GROUP = 2;
static_assert(MAX_UNIT_GROUP * UNIT_BLOCKS * 208 < 65536, "16-bit 0xFF counters per tile");

// forced (ICB_ENT_TILE, tests): <= 128 -> one unit per tile, larger -> two
__host__ __device__ inline int ent_unit_group(int nunits, int forced)
{
    if (forced > 0) return forced <= 128 ? 1 : 2;
    return nunits >= 2 * 2000 ? 2 : 1;
}
// worst case per block: DC 9+11 bits, 63 x (16+10) bits = 1658 bits -> 208 B
__host__ __device__ inline int ent_scr_stride(int unit_group) { return unit_group * UNIT_BLOCKS * 52 + 4; }   // scratch words per tile

// ---- geometry (IJG jcmaster.c initial_setup, 3 components 2x2,1x1,1x1) ----
struct Geom {
    int w, h;
    int pitch;              // device pitch in bytes (multiple of 16)
    int bgr;                // 1 = BGR byte order
    int mcu_cols, mcu_rows;
    int yblk_w, yblk_h;     // real luma block counts
    int crows_real;         // ceil(h/2): downsampled rows that come from pixels
    int nblocks;            // 6 * mcu_cols * mcu_rows
    int units_per_row;      // transform tiles per MCU row
    int nunits;             // survivor-list units = transform tiles
    int unit_group;         // units per entropy tile (1 or 2)
    int ntiles;             // entropy tiles = ceil(nunits / unit_group)
};

__host__ __device__ inline Geom make_geom(int w, int h, int pitch, int bgr, int forced_tile = 0)
{
    Geom g;
    g.w = w; g.h = h; g.pitch = pitch; g.bgr = bgr;
    g.mcu_cols = (w + 15) >> 4;
    g.mcu_rows = (h + 15) >> 4;
    g.yblk_w = (w + 7) >> 3;
    g.yblk_h = (h + 7) >> 3;
    g.crows_real = (h + 1) >> 1;
    g.nblocks = 6 * g.mcu_cols * g.mcu_rows;
    g.units_per_row = (g.mcu_cols + TILE_MCUS - 1) / TILE_MCUS;
    g.nunits = g.units_per_row * g.mcu_rows;
    g.unit_group = ent_unit_group(g.nunits, forced_tile);
    g.ntiles = (g.nunits + g.unit_group - 1) / g.unit_group;
    return g;
}

// first block (scan order) and block count of unit u
__host__ __device__ inline void unit_span(int u, int units_per_row, int mcu_cols, int &b0, int &nb)
{
    const int my = u / units_per_row, tx = u - my * units_per_row;
    const int m0 = tx * TILE_MCUS;
    b0 = (my * mcu_cols + m0) * 6;
    nb = 6 * (mcu_cols - m0 < TILE_MCUS ? mcu_cols - m0 : TILE_MCUS);
}

// ---- entropy pass tiling ----
constexpr int ENT_WARPS = 8;
constexpr int ENT_THREADS = ENT_WARPS * 32;
constexpr int ENT_WIN = 256;                        // shared-memory bit window per warp (words)
constexpr int ENT_WSTRIDE = ENT_WIN + 4;
constexpr int ENT_RING = 16;                        // CTA rounds whose tile records fit in shared memory at once

constexpr int HEADER_BYTES = 623;                   // a2.8, 3-component JFIF

// ---- device state of one fit job (lives in HBM, one per slot) ----
enum FitMode { MODE_SEARCH = 0, MODE_SINGLE = 1 };

// quantisation tables of one quality, indexed by key = 2 * zigzag pos + table (0 luma, 1 chroma)
struct QTables {
    uint2 rh[128];                  // x = reciprocal: floor(t / (8 q)) == umulhi(t, x) for t < 2^17;  y = 4 q
    unsigned char qtab[2][64];      // the table itself, zigzag order (DQT payload)
};

struct FitState {
    // search control (core/ImageCompressionJpg.java:158-200), float32 as in Java
    float lo, hi, best;
    float cur_q;
    int iter;               // iterations started
    int done;
    int cur_valid;          // the coming pass has work
    int mode;
    int cur_buf, best_buf;  // ping-pong output buffers
    int tsel;               // table slot (0..2) of the coming pass
    int n_trace;
    int table_mode, jfif_minor;
    int passes;             // entropy passes actually run
    int pad0;
    long long target;
    long long cur_size, best_size;
    float trace_q[8];
    long long trace_size[8];
    long long pad_a;        // keeps tab 16-byte aligned (the pass loads table entries in pairs)
    // three table slots: the coming pass reads tab[tsel]; while it runs, block 0 fills the other two
    // with the tables of both possible next midpoints ((tsel+1)%3: this pass fits, (tsel+2)%3: it
    // does not), so the bisection step at the end of the pass only flips tsel
    QTables tab[3];
    unsigned char cand_qtab[2][2][64];  // [ping-pong buffer] tables of the pass whose stream sits in that buffer (DQT payload)
    // pass scratch
    unsigned int done_a;            // finished CTAs of k_entropy_pass
    int pad1, pad2, pad3;
    // kernel clock (always on, a handful of timer reads per job): GPU global timer when block 0 of a kernel of
    // this job started its work and when the kernel finished -- the duration a kernel really has inside a busy
    // step, with other jobs' kernels sharing the GPU (bench.py's in-step roofline).  [0] transform (end = the
    // job-start kernel's griddepcontrol.wait returning), [1] job start / list kernel (start only), [2..9] entropy
    // passes by iteration (end = the last CTA's bisection step), [10] emit (start only)
    unsigned long long kt[11][2];
    unsigned int pad_kt[2];
    unsigned long long dbg_k[4];    // ICB_TIMING: globaltimer at survivor-kernel start / emit start / emit end (block 0)
    unsigned long long dbg_w[24];   // ICB_TIMING: clock64 timeline of one warp of the last pass
    unsigned long long dbg_t[8][4]; // per pass: globaltimer at kernel start / tail start / step start / tail end
    unsigned long long ff_total;    // stuffed 0x00 bytes (= number of 0xFF data bytes)
    unsigned long long total_bits;  // entropy-coded bits before padding
    // quality whose tables bound the survivor lists as they stand after the k-th pass of a search
    // (k_entropy_prune number k reads [k] and writes [k + 1])
    float list_q[12];
};

static_assert(offsetof(FitState, iter) == 16 && offsetof(FitState, cur_buf) == 32, "k_entropy_pass reads these as int4");
static_assert(offsetof(FitState, tab) % 16 == 0 && sizeof(QTables) % 16 == 0 && offsetof(FitState, trace_q) == 88, "control block = first 6 int4");
static_assert(offsetof(FitState, tab) % 8 == 0 && offsetof(FitState, cand_qtab) % 4 == 0, "vector copies of the tables");

// ---- Annex K tables (T.81), natural order; IJG jcparam.c / JDK JPEGQTable ----
__constant__ unsigned char c_base_lum[64] = {
    16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55,
    14, 13, 16, 24, 40, 57, 69, 56, 14, 17, 22, 29, 51, 87, 80, 62,
    18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92,
    49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
__constant__ unsigned char c_base_chr[64] = {
    17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99,
    24, 26, 56, 99, 99, 99, 99, 99, 47, 66, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
// zigzag index -> natural index
__constant__ unsigned char c_zigzag[64] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5,
    12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
    58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// Huffman tables as they go into the DHT segments
__constant__ unsigned char c_dc_bits[2][16] = {
    {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0},
    {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0}};
__constant__ unsigned char c_ac_bits[2][16] = {
    {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d},
    {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77}};
__constant__ unsigned char c_ac_vals[2][162] = {
    {0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07,
     0x22, 0x71, 0x14, 0x32, 0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0,
     0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0a, 0x16, 0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28,
     0x29, 0x2a, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49,
     0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69,
     0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89,
     0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7,
     0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5,
     0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe1, 0xe2,
     0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
     0xf9, 0xfa},
    {0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71,
     0x13, 0x22, 0x32, 0x81, 0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0,
     0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16, 0x24, 0x34, 0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26,
     0x27, 0x28, 0x29, 0x2a, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48,
     0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68,
     0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x82, 0x83, 0x84, 0x85, 0x86, 0x87,
     0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5,
     0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3,
     0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda,
     0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
     0xf9, 0xfa}};

// Derived encode tables (jchuff.c jpeg_make_c_derived_tbl), filled by the host at
// init: entry = (code << 8) | length, index = plane * 256 + a * 16 + size with plane = 2 * table + isdc
// (bits 16-17 of a survivor entry).  AC planes (0, 2): a = zero run, i.e. a * 16 + size = the symbol.  DC planes
// (1, 3): a = 0 the plain DC code; a = 1 / 2 the EOB of the luminance / chrominance AC table joined in front of
// it (what a DC entry emits when the block before it ended early) -- at most 4 + 11 bits.
// Global memory, not __constant__: every CTA of the entropy pass copies the table into shared
// memory with one coalesced load per thread (an indexed constant load replays 32 times per warp).
constexpr int HUF_WORDS = 1024;
__device__ __align__(16) unsigned int g_huf[HUF_WORDS];

}  // namespace icb
