/*
 * imgcomp_b200.h -- C ABI of the B200-native JPG target-size / bilinear-downscale
 * hot path of PolloChang/image-compression.
 *
 * The reference has no FFI of its own: the path sits behind four static Java
 * methods.  Each entry point below names the reference
 * seam it replaces; INTEGRATION.md shows the java.lang.foreign (Panama FFM)
 * downcall a maintainer binds to it.  Plain pointers and sizes only; the library
 * never throws or aborts -- every call returns an icb_status.
 *
 * All citations are into /root/reference/src/main/java/work/pollochang/compression/image/.
 */
#ifndef IMGCOMP_B200_H
#define IMGCOMP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Status codes.  Mapping to core/CompressionResult.java:3-16 (done by the Java
 * caller, core/ImageCompression.java:81-104):
 *   ICB_OK          -> seam returns true  -> COMPRESSED_SUCCESS
 *   ICB_NO_FIT      -> seam returns false -> FAILED_COMPRESSION (output deleted)
 *   ICB_ERR_OOM     -> OutOfMemoryError   -> FAILED_OUT_OF_MEMORY
 *   ICB_ERR_INVALID -> NullPointerException / IllegalArgumentException
 *   ICB_ERR_UNSUPPORTED -> fall back to the stock Java path (gray / CMYK / ICC)
 *   other negatives -> FAILED_UNKNOWN */
typedef enum {
    ICB_OK = 0,
    ICB_NO_FIT = 1,
    ICB_ERR_INVALID = -1,
    ICB_ERR_OOM = -2,
    ICB_ERR_CUDA = -3,
    ICB_ERR_UNSUPPORTED = -4,
    ICB_ERR_CAPACITY = -5,   /* caller's output buffer too small; *out_len = needed */
    ICB_ERR_NO_DEVICE = -6,
    ICB_ERR_TICKET = -7
} icb_status;

/* Pixel formats = byte order in memory of the decoded BufferedImage raster. */
typedef enum {
    ICB_FMT_RGB8 = 0,    /* what JPEGImageWriter hands libjpeg                       */
    ICB_FMT_BGR8 = 1,    /* BufferedImage.TYPE_3BYTE_BGR (ImageIO JPEG reader)       */
    ICB_FMT_RGBA8 = 2,   /* resize only                                              */
    ICB_FMT_ABGR8 = 3,   /* BufferedImage.TYPE_4BYTE_ABGR (ImageIO PNG reader); resize only */
    ICB_FMT_GRAY8 = 4,   /* BufferedImage.TYPE_BYTE_GRAY; resize / subsample only (1-component JPG: stock writer) */
    ICB_FMT_ARGB32 = 5,  /* BufferedImage.TYPE_INT_ARGB raster as bytes (little-endian ints: B G R A) --
                            tools/ImageTools.java:12-15 falls back to it for type-0 / custom images; resize only */
    ICB_FMT_XRGB32 = 6   /* BufferedImage.TYPE_INT_RGB raster as bytes (B G R x; x is interpolated like a
                            colour byte, Java leaves it 0); resize only */
} icb_fmt;

typedef struct icb_handle icb_handle;
typedef struct icb_image icb_image;      /* a decoded image resident in one GPU's HBM */

/* Pool creation.  Replaces Executors.newFixedThreadPool(availableProcessors)
 * (CompressionBatch.java:64-68): every listed GPU gets `slots_per_gpu` work
 * slots, each a CUDA stream + device arenas sized for max_w x max_h (grown on
 * demand).  devices == NULL / n_devices == 0 means "every visible GPU". */
int icb_init(const int *devices, int n_devices, int slots_per_gpu, int max_w, int max_h,
             icb_handle **out_handle);
void icb_shutdown(icb_handle *h);
int icb_device_count(const icb_handle *h);
const char *icb_last_error(const icb_handle *h);

/* [JDK-mem] switches kept in one place (SURVEY.md section 7 hard part 3):
 * table_mode 0 = JDK float32 JPEG.convertToLinearQuality + getScaledInstance,
 * 1 = IJG integer jpeg_quality_scaling; jfif_minor = APP0 version byte (JDK 2). */
int icb_set_profile(icb_handle *h, int table_mode, int jfif_minor);

/* Page-locked host staging for FFM MemorySegments (decode straight into these). */
void *icb_alloc_pinned(size_t bytes);
void icb_free_pinned(void *p);

/* core/ImageCompressionJpg.java:77-122 compressJpgWithTargetSize, including
 * :216-238 tryCachedParams, :158-200 findBestQualityByBinarySearch and :255-260
 * saveCompressedImage.  File write (:258) and cache.put (:111) stay in Java: the
 * JFIF bytes and the learned (best_q, scale) come back instead.
 *   px/pitch/fmt : decoded pixels (RGB8 or BGR8), caller-owned
 *   init_q       : CompressionParams.quality()      target: targetMaxSizeBytes()
 *   has_cached / cached_q / cached_scale : cache.get(key) (LearnedParams.java:8)
 *   out/out_cap  : caller-owned output buffer; out_len <= target on ICB_OK
 * init_q > 1 (or NaN) and a cached quality outside [0, 1] return ICB_ERR_INVALID: the reference's
 * ImageWriteParam.setCompressionQuality throws IllegalArgumentException there. */
int icb_jpg_fit(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt,
                float init_q, int64_t target,
                int has_cached, float cached_q, double cached_scale,
                uint8_t *out, size_t out_cap, size_t *out_len,
                float *best_q, double *scale, int *used_cache);

/* core/ImageCompressionJpg.java:136-147 compressJpgToStream(image, os, quality):
 * one baseline JFIF encode at an explicit quality. */
int icb_jpg_encode(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt,
                   float q, uint8_t *out, size_t out_cap, size_t *out_len);

/* Test / bench hook (no reference seam): exact stuffed stream size of the same
 * image at each of n qualities, one forward transform. */
int icb_jpg_sizes(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt,
                  const float *qs, int n, int64_t *sizes);

/* Test hook: the unquantised islow coefficients the transform kernel produced
 * (MCU-interleaved block order, 64 int16 per block, zigzag order). */
int icb_jpg_coefficients(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt,
                         int16_t *out, size_t out_elems);

/* tools/ImageTools.java:7-26 resizeImage: bilinear drawImage into dw x dh.
 * icb_resize_dims is lines :8-9 (newW = max(1,(int)(w*scale))). */
void icb_resize_dims(int w, int hgt, double scale, int *dw, int *dh);
int icb_resize_bilinear(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt,
                        uint8_t *dst, int dw, int dh, int dpitch);

/* core/ImageCompressionPng.java:37-75 compressPngWithTargetSize up to (not
 * including) ImageIO.write :70, which stays on the host: dimension rule :49-61,
 * then the resize.  ICB_NO_FIT = "already within min_w x min_h" (:49-53 returns
 * false).  dst must hold dw*dh*bpp bytes (query with dst == NULL first). */
int icb_png_fit(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt,
                int min_w, int min_h, uint8_t *dst, size_t dst_cap, int *dw, int *dh);
/* The dimension rule alone (core/ImageCompressionPng.java:49-61 + tools/ImageTools.java:8-9),
 * host arithmetic, no device needed: ICB_NO_FIT when w <= min_w && h <= min_h, else ICB_OK
 * with the target size and the double scale = min(min_w / w, min_h / h). */
int icb_png_target_dims(int w, int hgt, int min_w, int min_h, int *dw, int *dh, double *scale);

/* Input staging.  core/ImageCompression.java:140-153: when the longer side
 * exceeds 4096 the reference decodes with setSourceSubsampling(s, s, 0, 0),
 * s = highestOneBit(floor(maxDim / 4096)).  icb_subsample_factor is that rule (host
 * arithmetic); icb_subsample is the decimation itself for callers that decode at full
 * resolution: dst gets source pixels (x*s, y*s), ceil(w/s) x ceil(h/s) of them. */
int icb_subsample_factor(int w, int hgt);
int icb_subsample(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt, int s,
                  uint8_t *dst, int dpitch);

/* Asynchronous form = the per-GPU stream work queue that replaces the executor
 * (CompressionBatch.java:77-88 submit, :97-107 await).  submit blocks only while
 * every slot is busy; wait returns that job's status and releases the slot.
 * Host buffers must stay valid until wait returns (pinned ones copy async). */
typedef struct {
    int status;
    size_t out_len;
    float best_q;
    double scale;
    int used_cache;
    int n_passes;       /* entropy passes run on the device for this image */
    int device;
} icb_result;

int icb_submit_jpg_fit(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt,
                       float init_q, int64_t target,
                       int has_cached, float cached_q, double cached_scale,
                       uint8_t *out, size_t out_cap, int64_t *ticket);
int icb_wait(icb_handle *h, int64_t ticket, icb_result *res);
/* Non-blocking: *done = 1 once the job has finished (a later icb_wait returns at once), 0 while its
 * stream still has work.  A fit that needs another phase (cached miss, next scale of the 0.85 loop) is
 * moved on from here, so a caller that only polls still makes progress.  The slot stays taken until
 * icb_wait collects the result. */
int icb_poll(icb_handle *h, int64_t ticket, int *done);
/* A submitted job holds its work slot until icb_wait returns.  Every entry point takes a slot, so a
 * thread that has slots_per_gpu * n_devices tickets outstanding must wait on one of them before it
 * calls anything else (the call would block for a slot only that thread can free).
 *
 * Host buffers may be pageable (a Java heap array copied into a native segment, a numpy array): the
 * submitting thread then copies the pixels into the slot's page-locked staging slice by slice while
 * the copy engine moves the previous slice, and results are staged the same way on the way back.
 * Buffers from icb_alloc_pinned skip that copy. */

/* tools/ImageTools.java:7-26 / core/ImageCompressionPng.java:37-66 in the queue: host pixels in,
 * resized pixels out (dst filled when icb_wait returns).  icb_submit_png_fit applies the dimension
 * rule first: ICB_NO_FIT means nothing was queued and no ticket was issued. */
int icb_submit_resize(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt,
                      uint8_t *dst, int dw, int dh, int dpitch, int64_t *ticket);
int icb_submit_png_fit(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt,
                       int min_w, int min_h, uint8_t *dst, size_t dst_cap, int *dw, int *dh,
                       int64_t *ticket);

/* Device-resident inputs (bench `value`: pixels already in HBM when timing starts). */
int icb_image_upload(icb_handle *h, int device_index, const uint8_t *px, int w, int hgt, int pitch,
                     int fmt, icb_image **img);
void icb_image_free(icb_handle *h, icb_image *img);
int icb_submit_jpg_fit_image(icb_handle *h, const icb_image *img, float init_q, int64_t target,
                             int has_cached, float cached_q, double cached_scale,
                             uint8_t *out, size_t out_cap, int64_t *ticket);
/* dst == NULL leaves the resized pixels in the slot's HBM buffer (no D2H: kernel-only timing) */
int icb_submit_resize_image(icb_handle *h, const icb_image *img, int dw, int dh,
                            uint8_t *dst, int dpitch, int64_t *ticket);

/* The batch executor itself (CompressionBatch.java:64-108: newFixedThreadPool + one task per list line +
 * awaitTermination) in native code: `threads` worker threads of the library pull items from ONE atomic
 * cursor and each keeps `inflight` jobs in the per-GPU stream work queues of the handle (all its GPUs).
 * Returns when every item has a result in item.res (status as from icb_wait; ICB_NO_FIT for a PNG item
 * already inside the box).  The Java host calls this once per chunk of decoded lines, or keeps its own
 * pool on icb_submit_* / icb_wait -- both drive the same queues. */
typedef enum {
    ICB_ITEM_JPG_FIT = 0,        /* host pixels -> icb_submit_jpg_fit      */
    ICB_ITEM_JPG_FIT_IMAGE = 1,  /* resident image -> icb_submit_jpg_fit_image */
    ICB_ITEM_PNG_FIT = 2,        /* host pixels -> icb_submit_png_fit (resized pixels out) */
    ICB_ITEM_PNG_COMPRESS = 3,   /* host pixels -> icb_submit_png_compress (PNG file out)  */
    ICB_ITEM_RESIZE_IMAGE = 4,   /* resident image -> icb_submit_resize_image (dw, dh given; out may be NULL) */
    ICB_ITEM_JPG_FILE_FIT = 5    /* JPEG file bytes -> icb_submit_jpg_file_fit: px = the file, pitch = its length in
                                    bytes; dw, dh receive the dimensions of the decoded (subsampled) raster */
} icb_item_kind;

typedef struct {
    int kind;
    const uint8_t *px;           /* host pixels (kinds 0, 2, 3) */
    const icb_image *img;        /* resident image (kinds 1, 4) */
    int w, h, pitch, fmt;
    float init_q;                /* CompressionParams.quality()            */
    int64_t target;              /* CompressionParams.targetMaxSizeBytes() */
    int has_cached;
    float cached_q;
    double cached_scale;
    int min_w, min_h;            /* CompressionParams.minWidth() / minHeight() (PNG) */
    uint8_t *out;
    size_t out_cap;
    int dw, dh;                  /* out: PNG / resize dimensions (in for kind 4) */
    icb_result res;              /* out */
} icb_item;

int icb_run_list(icb_handle *h, icb_item *items, size_t n, int threads, int inflight);

/* SURVEY 8 f3 -- core/ImageCompressionPng.java:70 ImageIO.write(resized, "png", file) on the GPU: the
 * PNG FILE (signature, IHDR, 8 KB IDAT chunks, IEND) of the given pixels; per-row filter choice by the
 * minimum sum of absolute values, one dynamic-Huffman deflate block.  A zlib stream is not unique: the
 * file decodes to exactly the input pixels, its size is close to (not equal to) the JDK writer's.
 * icb_png_compress is the whole seam compressPngWithTargetSize (:37-75): dimension rule, resize with the
 * pixels staying in HBM, PNG file out; ICB_NO_FIT = already inside min_w x min_h (returns false, :49-53).
 * ICB_ERR_CAPACITY: out too small, *out_len (icb_result.out_len) = bytes needed. */
int icb_png_encode(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt,
                   uint8_t *out, size_t out_cap, size_t *out_len);
int icb_submit_png_encode(icb_handle *h, const uint8_t *px, int w, int hgt, int pitch, int fmt,
                          uint8_t *out, size_t out_cap, int64_t *ticket);
int icb_png_compress(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt,
                     int min_w, int min_h, uint8_t *out, size_t out_cap, size_t *out_len,
                     int *dw, int *dh);
int icb_submit_png_compress(icb_handle *h, const uint8_t *src, int w, int hgt, int pitch, int fmt,
                            int min_w, int min_h, uint8_t *out, size_t out_cap, int *dw, int *dh,
                            int64_t *ticket);

/* SURVEY 8 f2 (input staging) -- core/ImageCompression.java:119-155 decodeImageWithSubsampling on the GPU:
 * reader.read(0, param) of the ImageIO JPEG reader (libjavajpeg = IJG 6b: sequential Huffman decode, islow inverse
 * DCT, fancy upsampling (h2v2 / h2v1), YCbCr -> RGB) for baseline 8-bit YCbCr 4:2:0, 4:2:2 and 4:4:4 files with one scan (restart
 * intervals included), so that the caller uploads the FILE instead of the decoded raster (12 MP: ~6 MB instead of 36 MB over
 * PCIe).  ICB_ERR_UNSUPPORTED = any other kind of JPEG (progressive, gray, CMYK, 12 bit, other sampling factors) or a stream the parallel decoder did not resolve: the caller decodes on the host as before and calls the
 * pixel entry points.  ICB_ERR_INVALID = not a JPEG / truncated.
 *   icb_jpg_file_info   header parse on the host (:128-129 reader.getWidth / getHeight), no device needed
 *   icb_jpg_decode      the decoded raster (RGB8 or BGR8 = TYPE_3BYTE_BGR) back on the host: test / parity entry
 *   icb_jpg_file_fit    :119-155 + core/ImageCompressionJpg.java:77-122 in one call: decode, the source-subsampling
 *                       rule of :140-153 (icb_subsample_factor), then exactly icb_jpg_fit on the pixels, which never
 *                       leave the GPU.  *w, *hgt = dimensions of the raster the seam saw (for CacheTools.createKey). */
int icb_jpg_file_info(const uint8_t *file, size_t n, int *w, int *hgt);
int icb_jpg_decode(icb_handle *h, const uint8_t *file, size_t n, int fmt, uint8_t *dst, size_t dst_cap, int dpitch,
                   int *w, int *hgt);
int icb_jpg_file_fit(icb_handle *h, const uint8_t *file, size_t n, float init_q, int64_t target,
                     int has_cached, float cached_q, double cached_scale,
                     uint8_t *out, size_t out_cap, size_t *out_len,
                     float *best_q, double *scale, int *used_cache, int *w, int *hgt);
int icb_submit_jpg_file_fit(icb_handle *h, const uint8_t *file, size_t n, float init_q, int64_t target,
                            int has_cached, float cached_q, double cached_scale,
                            uint8_t *out, size_t out_cap, int *w, int *hgt, int64_t *ticket);

/* No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The PNG
 * encoder of icb_png_encode on DEVICE buffers, allocating nothing: no cudaMalloc / cudaFree / cudaHostAlloc and
 * no allocating library routine, on any call.  Its work arena (filtered rows, piece bit streams, histogram,
 * code tables) is a file-scope `static __device__` array of 5.3 MB that every GPU which loads the library
 * holds until exit; limit: h * (1 + w * png_bytes_per_pixel) <= 1 048 576 filtered bytes (e.g. 512 x 511 RGBA,
 * 590 x 590 RGB) -- larger images return ICB_ERR_CAPACITY with the sizes in icb_last_error.
 *   d_px / pitch / fmt : pixels in device memory of the CURRENT device (any icb_fmt), caller-owned
 *   d_out / out_cap    : device buffer for the PNG file; bytes past out_cap are not written
 *   d_out_len          : device word that receives the file's length (compare with out_cap after the stream
 *                        has drained: a larger value means the file did not fit and d_out is incomplete)
 *   stream             : a cudaStream_t (NULL = the default stream); the seven kernels are queued on it and the
 *                        call returns without synchronising
 *   h                  : optional (may be NULL): only receives the error text and the launch counter
 * The first call on each device looks up the arena's device address (cudaGetSymbolAddress) and uploads the
 * 5 KB CRC table; make that call outside a graph capture.  Every call shares the one arena, which keeps the last
 * call's contents (the call clears what must start at zero): calls that overlap on the GPU -- two streams, or
 * two threads -- race; serialise them on one stream. */
int icb_png_encode_device_noalloc(icb_handle *h, const uint8_t *d_px, int w, int hgt, int pitch, int fmt,
                                  uint8_t *d_out, size_t out_cap, unsigned long long *d_out_len, void *stream);

/* Counters (CompressionBatch.java:56-62 are summed by the Java caller; these are
 * the device-side ones) and per-stage CUDA-event times of the last profiled job. */
typedef struct {
    uint64_t images_done;
    uint64_t kernels_launched;     /* every kernel this library launched */
    uint64_t entropy_passes;
    uint64_t h2d_bytes, d2h_bytes;
    float last_transform_ms;       /* set by icb_profile_jpg_fit_image only */
    float last_entropy_ms;
    float last_total_ms;
    int last_transform_launches, last_entropy_launches;
    /* Kernel clocks, always on: for every finished fit, the GPU-global-timer span from the first CTA's start
     * to the last CTA's end of [0] the transform, [1] each entropy pass, [2] the emit kernel, summed, and how
     * many launches went into each sum.  Unlike the profile calls below these are taken INSIDE the running
     * batch, with other jobs' kernels sharing the GPU: sum / launches is a kernel's in-step duration. */
    uint64_t kernel_clock_ns[3];
    uint64_t kernel_clock_launches[3];
} icb_stats_t;
int icb_stats(icb_handle *h, icb_stats_t *out);

/* Runs one fit on a resident image with CUDA events, on the slot's own stream,
 * around the transform kernel, around the whole search (step + entropy + publish
 * kernels) and around every entropy-pass launch that did work (bench roofline). */
int icb_profile_jpg_fit_image(icb_handle *h, const icb_image *img, float init_q, int64_t target,
                              int reps, float *transform_ms, float *entropy_ms, float *total_ms,
                              float *entropy_kernel_ms, int *entropy_kernel_launches);

/* Average duration of the transform kernel alone: the transforms of n resident images (all on one
 * GPU) are enqueued back to back on one slot's stream, reps rounds, between two CUDA events after
 * one untimed round (so neither the launch gap after an idle stream nor the event pair of a single
 * launch is in the figure).  list_q > 0: the kernel also writes the survivor lists for that quality, as
 * it does inside a fit (the production form); 0: coefficients only.  *launches = kernels timed. */
int icb_profile_transform_series(icb_handle *h, const icb_image *const *imgs, int n, int reps, float list_q,
                                 float *ms_per_launch, int *launches);

/* The resize kernel alone (with_png = 0) or resize + PNG encoder (1) on a resident image: reps runs back to
 * back on one slot's stream between two CUDA events after two untimed ones; nothing leaves the GPU. */
int icb_profile_resize_series(icb_handle *h, const icb_image *img, int dw, int dh, int reps, int with_png,
                              float *ms_per_launch);

/* Device-side stopwatch over all slot streams of one GPU (bench timing with CUDA
 * events although the work runs on library-owned streams): icb_mark(.., 0) before
 * the first submit, icb_mark(.., 1) after the last wait, then icb_elapsed_ms. */
int icb_mark(icb_handle *h, int device_index, int which);
int icb_elapsed_ms(icb_handle *h, int device_index, float *ms);

#ifdef __cplusplus
}
#endif
#endif
