/*
 * oracle.h -- CPU oracle for the JPG target-size / bilinear-downscale hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it, and there only as the checker / the timed CPU baseline.
 *
 * What it restates (reference = /root/reference, PolloChang/image-compression):
 *   - control flow: src/main/java/work/pollochang/compression/image/core/
 *       ImageCompressionJpg.java:77-122 (scale loop), :136-147 (encode),
 *       :158-200 (float32 quality bisection), :216-238 (cached params),
 *       :255-260 (final save); ImageCompressionPng.java:43-66 (dimension rule);
 *       tools/ImageTools.java:7-26 (resize dims + bilinear drawImage).
 *   - arithmetic: NOT in the reference tree.  ImageCompressionJpg.java:137-143
 *     calls javax.imageio's JPEGImageWriter = OpenJDK 21 java.desktop
 *     libjavajpeg (IJG libjpeg 6b) reached through TwelveMonkeys imageio-jpeg
 *     3.12.0 (build.gradle:53).  The published IJG-6b baseline algorithm
 *     (jccolor/jcsample/jcprepct/jfdctint/jcdctmgr/jccoefct/jchuff/jcmarker) and
 *     the JDK's JPEG.convertToLinearQuality + JPEGQTable.getScaledInstance +
 *     JFIFMarkerSegment defaults are restated here from their published
 *     behaviour ([JDK-mem] in SURVEY.md).
 *
 * PARITY PIN: the reference has no JPG tests and cannot run here (no JVM), so
 * the IJG arithmetic + marker layout is pinned against an independent
 * implementation available in this container, Pillow/libjpeg-turbo, fed the
 * same quantisation tables (tests/golden/, tools/make_golden.py).  The two
 * JDK-only facts (float32 table formula, JFIF version bytes 01 02) and the
 * Java2D fixed-point bilinear remain "parity unpinned" against a real JDK.
 */
#ifndef IMGCOMP_ORACLE_H
#define IMGCOMP_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* pixel formats (byte order in memory) */
enum { ORC_FMT_RGB8 = 0, ORC_FMT_BGR8 = 1, ORC_FMT_RGBA8 = 2, ORC_FMT_ABGR8 = 3,
       ORC_FMT_GRAY8 = 4, ORC_FMT_ARGB32 = 5, ORC_FMT_XRGB32 = 6 };    /* 4-6: resize only */

/* "JDK profile" switches: every [JDK-mem] fact lives behind one of these. */
typedef struct {
    int table_mode;  /* 0 = JDK float32 (JPEG.convertToLinearQuality + getScaledInstance)
                        1 = IJG integer jpeg_quality_scaling: (base*S+50)/100            */
    int jfif_minor;  /* 2 = JDK JFIFMarkerSegment default (01 02); 1 = libjpeg (01 01)   */
} orc_profile;

/* a2.1: quality -> two 64-entry tables, natural order, clamped to [1,255]. */
void orc_build_qtables(float q, const orc_profile *prof, uint8_t lum[64], uint8_t chr[64]);

/* Geometry helpers shared with the tests (IJG jcmaster.c initial_setup). */
typedef struct {
    int w, h;
    int mcu_cols, mcu_rows;      /* ceil(w/16), ceil(h/16)                  */
    int yblk_w, yblk_h;          /* ceil(w/8), ceil(h/8)                     */
    int cblk_w, cblk_h;          /* ceil(ceil(w/2)/8), ceil(ceil(h/2)/8)     */
    long nblocks;                /* 6 * mcu_cols * mcu_rows                  */
} orc_geom;
void orc_geometry(int w, int h, orc_geom *g);

/* a2.2-a2.4 + a2.6: unquantised islow coefficients (8x true DCT), MCU-interleaved
 * block order Y00 Y01 Y10 Y11 Cb Cr, each block 64 int16 in ZIGZAG order; dummy
 * blocks carry AC=0 and the DC of the block IJG copies it from.  out has
 * 64*nblocks entries. */
int orc_forward_coefficients(const uint8_t *px, int w, int h, int pitch, int fmt, int16_t *out);

/* a2 with explicit tables (natural order).  Writes at most cap bytes; *out_len
 * is always the full stream length (so cap=0 sizes without storing). */
int orc_jpeg_encode_tables(const uint8_t *px, int w, int h, int pitch, int fmt,
                           const uint8_t lum[64], const uint8_t chr[64], int jfif_minor,
                           uint8_t *out, size_t cap, size_t *out_len);

/* a2: compressJpgToStream(img, os, q) under a profile. */
int orc_jpeg_encode(const uint8_t *px, int w, int h, int pitch, int fmt, float q,
                    const orc_profile *prof, uint8_t *out, size_t cap, size_t *out_len);

/* a1: findBestQualityByBinarySearch.  trace_q/trace_size (8 entries each, may
 * be NULL) receive the visited midpoints and their sizes; *n_trace the count. */
float orc_find_best_quality(const uint8_t *px, int w, int h, int pitch, int fmt,
                            long target, float initial_q, const orc_profile *prof,
                            float *trace_q, long *trace_size, int *n_trace);

/* a5+a4+a1+a3: compressJpgWithTargetSize.  Returns 0 = fits (bytes in out),
 * 1 = no quality/scale fits, <0 error.  used_cache: 1 if tryCachedParams hit. */
int orc_jpg_fit(const uint8_t *px, int w, int h, int pitch, int fmt,
                float initial_q, long target,
                int has_cached, float cached_q, double cached_scale,
                const orc_profile *prof,
                uint8_t *out, size_t cap, size_t *out_len,
                float *best_q, double *scale, int *used_cache, int *n_encodes);

/* a6: ImageTools.resizeImage dims; then Java2D bilinear ([JDK-mem] 32.32 fixed
 * point source walk, 8-bit fractions) and a float64 pixel-centre bilinear. */
void orc_resize_dims(int w, int h, double scale, int *nw, int *nh);
int orc_resize_bilinear(const uint8_t *src, int w, int h, int pitch, int fmt,
                        uint8_t *dst, int dw, int dh, int dpitch);
int orc_resize_bilinear_f64(const uint8_t *src, int w, int h, int pitch, int fmt,
                            uint8_t *dst, int dw, int dh, int dpitch);

/* f2 (input staging): the decode rule of core/ImageCompression.java:140-153 and what
 * javax.imageio's ImageReadParam.setSourceSubsampling(s, s, 0, 0) hands back: source
 * pixels (x*s, y*s), ceil(w/s) x ceil(h/s) of them -- point sampling, no averaging
 * [JDK-mem: IIOParam javadoc]. */
int orc_subsample_factor(int w, int h);
int orc_subsample(const uint8_t *src, int w, int h, int pitch, int bpp, int s,
                  uint8_t *dst, int dpitch);

/* f2 (input staging): the JPEG reader of core/ImageCompression.java:119-155 for the files the CUDA decoder
 * takes (baseline SOF0, 8 bit, YCbCr 4:2:0 / 4:2:2 / 4:4:4, one scan, no restart interval) -- jpeg_decode_oracle.c. */
enum { ORC_DEC_OK = 0, ORC_DEC_UNSUPPORTED = -4, ORC_DEC_CORRUPT = -1 };
int orc_jpeg_info(const uint8_t *file, size_t n, int *w, int *h);
int orc_jpeg_decode_coefficients(const uint8_t *file, size_t n, int16_t *out);
int orc_jpeg_sampling(const uint8_t *file, size_t n, int *hs, int *vs);
int orc_jpeg_decode(const uint8_t *file, size_t n, int fmt, uint8_t *dst, int dpitch);

/* a7: ImageCompressionPng.java:49-61.  Returns 0 and the new dims when a resize
 * is due, 1 when the image already fits (reference returns false). */
int orc_png_target_dims(int w, int h, int min_w, int min_h, int *nw, int *nh, double *scale);

#ifdef __cplusplus
}
#endif
#endif
