/*
 * resize_oracle.c -- CPU restatement of ImageTools.resizeImage
 * (tools/ImageTools.java:7-26) and the PNG dimension rule
 * (core/ImageCompressionPng.java:43-66).  TEST INFRASTRUCTURE ONLY (oracle.h).
 *
 * The pixel arithmetic is not in the reference tree: Graphics2D.drawImage with
 * VALUE_INTERPOLATION_BILINEAR (ImageTools.java:20-23) runs OpenJDK 21 Java2D's
 * software loop (libawt TransformHelper.c Transform + BilinearInterp, LoopMacros.h
 * bilinear helper, AlphaMath.c mul8/div8 tables).  Its published behaviour is
 * restated below from memory ([JDK-mem], parity unpinned against a real JDK):
 *   - inverse transform scale = 1.0 / ((double)newW / w), source walk in 32.32
 *     fixed point from the destination pixel centre, minus one half;
 *   - 2x2 taps, coordinates clamped to the image, 8-bit fractions (top byte of
 *     the 32-bit fraction), result ((v<<16 terms) + (1<<15)) >> 16;
 *   - sources with alpha are interpolated premultiplied (mul8table) and written
 *     back through SrcOver onto the cleared destination (div8table).
 * orc_resize_bilinear_f64 is the plain float64 pixel-centre bilinear the stated
 * +-1 LSB tolerance (BASELINE.json north_star) is about.
 */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ImageTools.java:8-9 */
void orc_resize_dims(int w, int h, double scale, int *nw, int *nh)
{
    int a = (int)(w * scale), b = (int)(h * scale);
    *nw = a > 1 ? a : 1;
    *nh = b > 1 ? b : 1;
}

/* core/ImageCompression.java:140-153: subsampling = highestOneBit(floor(maxDim / 4096)) when
 * the longer side exceeds PREFERRED_MAX_DIM = 4096, else 1 */
int orc_subsample_factor(int w, int h)
{
    int max_dim = w > h ? w : h, s = 1;
    if (max_dim > 4096) {
        int f = (int)floor((double)max_dim / 4096);
        if (f > 1) { s = 1; while (s * 2 <= f) s *= 2; }      /* Integer.highestOneBit */
    }
    return s;
}

/* ImageReadParam.setSourceSubsampling(s, s, 0, 0): every s-th pixel of every s-th row */
int orc_subsample(const uint8_t *src, int w, int h, int pitch, int bpp, int s, uint8_t *dst, int dpitch)
{
    if (!src || !dst || s < 1 || bpp < 1) return -1;
    int dw = (w + s - 1) / s, dh = (h + s - 1) / s;
    for (int y = 0; y < dh; y++)
        for (int x = 0; x < dw; x++)
            memcpy(dst + (size_t)y * dpitch + (size_t)x * bpp, src + (size_t)(y * s) * pitch + (size_t)(x * s) * bpp,
                   (size_t)bpp);
    return 0;
}

/* ImageCompressionPng.java:49-61 */
int orc_png_target_dims(int w, int h, int min_w, int min_h, int *nw, int *nh, double *scale)
{
    if (w <= min_w && h <= min_h) return 1;
    double wr = (double)min_w / w, hr = (double)min_h / h;
    double s = wr < hr ? wr : hr;
    if (scale) *scale = s;
    orc_resize_dims(w, h, s, nw, nh);
    return 0;
}

static uint8_t mul8[256][256], div8[256][256];
static int alpha_ready;

/* AlphaMath.c initAlphaTables [JDK-mem] */
static void init_alpha(void)
{
    if (alpha_ready) return;
    memset(mul8, 0, sizeof(mul8));
    memset(div8, 0, sizeof(div8));
    for (unsigned i = 1; i < 256; i++) {
        unsigned inc = 0x10101u * i;          /* JDK uses int; the byte kept is the same */
        unsigned val = inc + (1u << 23);
        for (unsigned j = 1; j < 256; j++) {
            mul8[i][j] = (uint8_t)(val >> 24);
            val += inc;
        }
    }
    for (unsigned i = 1; i < 256; i++) {
        unsigned inc = ((0xffu << 24) + i / 2) / i;
        unsigned val = 1u << 23;
        unsigned j;
        for (j = 0; j < i; j++) {
            div8[i][j] = (uint8_t)(val >> 24);
            val += inc;
        }
        for (; j < 256; j++) div8[i][j] = 255;
    }
    alpha_ready = 1;
}

static void fmt_layout(int fmt, int *bpp, int *alpha_off)
{
    switch (fmt) {
    case ORC_FMT_RGBA8: *bpp = 4; *alpha_off = 3; break;
    case ORC_FMT_ABGR8: *bpp = 4; *alpha_off = 0; break;
    case ORC_FMT_ARGB32: *bpp = 4; *alpha_off = 3; break;      /* TYPE_INT_ARGB as little-endian bytes B G R A */
    case ORC_FMT_XRGB32: *bpp = 4; *alpha_off = -1; break;     /* TYPE_INT_RGB: no alpha, four plain bytes */
    case ORC_FMT_GRAY8: *bpp = 1; *alpha_off = -1; break;      /* TYPE_BYTE_GRAY */
    default:            *bpp = 3; *alpha_off = -1; break;
    }
}

/* TransformHelper.c DblToLong: truncation of d * 2^32 */
static int64_t dbl_to_long(double d) { return (int64_t)(d * 4294967296.0); }

int orc_resize_bilinear(const uint8_t *src, int w, int h, int pitch, int fmt,
                        uint8_t *dst, int dw, int dh, int dpitch)
{
    if (!src || !dst || w < 1 || h < 1 || dw < 1 || dh < 1) return -1;
    int bpp, aoff;
    fmt_layout(fmt, &bpp, &aoff);
    if (aoff >= 0) init_alpha();
    double m00 = (double)dw / w, m11 = (double)dh / h;      /* DrawImage.scaleImage */
    double ixx = 1.0 / m00, iyy = 1.0 / m11;                /* AffineTransform.createInverse */
    int64_t dxl = dbl_to_long(ixx), dyl = dbl_to_long(iyy);
    int64_t xbase = dbl_to_long(0.5 * ixx), ybase = dbl_to_long(0.5 * iyy);
    const int64_t half = (int64_t)1 << 31;
    for (int dy = 0; dy < dh; dy++) {
        int64_t yl = ybase + (int64_t)dy * dyl - half;
        int yw = (int)(yl >> 32);
        int yf = (int)(((uint32_t)yl) >> 24);
        int y0 = yw < 0 ? 0 : yw;
        int y1 = (yw + 1 < h) ? (yw < 0 ? 0 : yw + 1) : y0;
        if (y0 > h - 1) y0 = y1 = h - 1;
        const uint8_t *r0 = src + (size_t)y0 * pitch, *r1 = src + (size_t)y1 * pitch;
        uint8_t *o = dst + (size_t)dy * dpitch;
        for (int dx = 0; dx < dw; dx++) {
            int64_t xl = xbase + (int64_t)dx * dxl - half;
            int xw = (int)(xl >> 32);
            int xf = (int)(((uint32_t)xl) >> 24);
            int x0 = xw < 0 ? 0 : xw;
            int x1 = (xw + 1 < w) ? (xw < 0 ? 0 : xw + 1) : x0;
            if (x0 > w - 1) x0 = x1 = w - 1;
            const uint8_t *p00 = r0 + (size_t)x0 * bpp, *p01 = r0 + (size_t)x1 * bpp;
            const uint8_t *p10 = r1 + (size_t)x0 * bpp, *p11 = r1 + (size_t)x1 * bpp;
            int res[4];
            for (int c = 0; c < bpp; c++) {
                int c00 = p00[c], c01 = p01[c], c10 = p10[c], c11 = p11[c];
                if (aoff >= 0 && c != aoff) {               /* IntArgb -> IntArgbPre */
                    c00 = mul8[p00[aoff]][c00];
                    c01 = mul8[p01[aoff]][c01];
                    c10 = mul8[p10[aoff]][c10];
                    c11 = mul8[p11[aoff]][c11];
                }
                int top = (c00 << 8) + (c01 - c00) * xf;     /* BL_INTERP */
                int bot = (c10 << 8) + (c11 - c10) * xf;
                int v = (top << 8) + (bot - top) * yf;
                res[c] = (v + (1 << 15)) >> 16;
            }
            if (aoff >= 0) {                                 /* SrcOver onto cleared dst */
                int a = res[aoff];
                for (int c = 0; c < bpp; c++) {
                    if (c == aoff) o[dx * bpp + c] = (uint8_t)a;
                    else if (a == 0) o[dx * bpp + c] = 0;
                    else if (a == 255) o[dx * bpp + c] = (uint8_t)res[c];
                    else o[dx * bpp + c] = div8[a][res[c]];
                }
            } else {
                for (int c = 0; c < bpp; c++) o[dx * bpp + c] = (uint8_t)res[c];
            }
        }
    }
    return 0;
}

int orc_resize_bilinear_f64(const uint8_t *src, int w, int h, int pitch, int fmt,
                            uint8_t *dst, int dw, int dh, int dpitch)
{
    if (!src || !dst || w < 1 || h < 1 || dw < 1 || dh < 1) return -1;
    int bpp, aoff;
    fmt_layout(fmt, &bpp, &aoff);
    for (int dy = 0; dy < dh; dy++) {
        double sy = (dy + 0.5) * (double)h / dh - 0.5;
        double fy0 = floor(sy);
        double fy = sy - fy0;
        int y0 = (int)fy0, y1 = y0 + 1;
        if (y0 < 0) y0 = 0;
        if (y1 < 0) y1 = 0;
        if (y0 > h - 1) y0 = h - 1;
        if (y1 > h - 1) y1 = h - 1;
        for (int dx = 0; dx < dw; dx++) {
            double sx = (dx + 0.5) * (double)w / dw - 0.5;
            double fx0 = floor(sx);
            double fx = sx - fx0;
            int x0 = (int)fx0, x1 = x0 + 1;
            if (x0 < 0) x0 = 0;
            if (x1 < 0) x1 = 0;
            if (x0 > w - 1) x0 = w - 1;
            if (x1 > w - 1) x1 = w - 1;
            for (int c = 0; c < bpp; c++) {
                double c00 = src[(size_t)y0 * pitch + x0 * bpp + c];
                double c01 = src[(size_t)y0 * pitch + x1 * bpp + c];
                double c10 = src[(size_t)y1 * pitch + x0 * bpp + c];
                double c11 = src[(size_t)y1 * pitch + x1 * bpp + c];
                double top = c00 + (c01 - c00) * fx, bot = c10 + (c11 - c10) * fx;
                double v = top + (bot - top) * fy;
                int iv = (int)(v + 0.5);
                dst[(size_t)dy * dpitch + dx * bpp + c] = (uint8_t)(iv < 0 ? 0 : iv > 255 ? 255 : iv);
            }
        }
    }
    return 0;
}
