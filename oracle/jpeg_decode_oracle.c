/*
 * jpeg_decode_oracle.c -- CPU oracle for SURVEY section 8 row f2 (input staging): what
 * core/ImageCompression.java:119-155 gets back from javax.imageio's JPEG reader for a baseline
 * 4:2:0 JFIF file -- the decoded RGB raster the JPG seam (ImageCompressionJpg.java:77) is handed.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  The arithmetic is not in /root/reference: the JDK reader is
 * libjavajpeg = IJG libjpeg 6b with its defaults (dct_method JDCT_ISLOW, do_fancy_upsampling TRUE,
 * out_color_space JCS_RGB) [JDK-mem].  Restated from the published IJG-6b decoder:
 *   jdmarker.c  marker parse                      -> parse_headers
 *   jdhuff.c    decode_mcu (sequential baseline)   -> decode_block
 *   jddctmgr.c / jidctint.c  dequantise + islow inverse DCT, range limit with the & RANGE_MASK wrap -> idct_islow
 *   jdsample.c  h2v2_fancy_upsample (triangle filter) with jdmainct.c's context rows: the row above the
 *               first and below the last REAL chroma row is that row itself      -> chroma_at
 *   jdcolor.c   ycc_rgb_convert (16-bit fixed-point tables)                       -> ycc_to_rgb
 * PARITY PIN: Pillow / libjpeg-turbo decodes (bit-exact with IJG 6b for islow + fancy upsampling) of
 * files written at several sizes and qualities, tests/golden/jpeg_decode.npz (tests/golden/make_golden_decode.py).
 * Against a real JDK the decoded pixels are "parity unpinned" like every other [JDK-mem] fact.
 *
 * Scope = what the CUDA decoder takes: SOF0, 8 bits, 3 components, luma sampled 2x2 (4:2:0), 2x1 (4:2:2) or 1x1
 * (4:4:4) with 1x1 chroma, restart intervals (DRI / RSTn), one
 * interleaved scan.  Everything else returns ORC_DEC_UNSUPPORTED: the caller keeps
 * the stock host decode for such files.
 */
#include <stdlib.h>
#include <string.h>

#include "oracle.h"

typedef struct {
    int w, h;
    int mcu_cols, mcu_rows;
    int ri;                             /* restart interval in MCUs (DRI), 0 = none */
    int hs, vs;                         /* luma sampling factors: 2x2 (4:2:0), 2x1 (4:2:2), 1x1 (4:4:4) */
    unsigned short qt[4][64];           /* natural order */
    int qt_set[4];
    /* jdhuff.c derived tables */
    int maxcode[4][18], valoff[4][18];
    unsigned char huffval[4][256];
    int huf_set[4];                     /* index = class * 2 + id (class 0 DC, 1 AC) */
    int comp_tq[3], comp_td[3], comp_ta[3];
    size_t scan_off;                    /* first byte of entropy-coded data */
} dec_t;

static const unsigned char k_zz[64] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

/* jdhuff.c jpeg_make_d_derived_tbl (the maxcode / valoffset part) */
static int build_huff(dec_t *d, int idx, const unsigned char *bits, const unsigned char *vals, int nvals)
{
    int code = 0, k = 0;
    for (int l = 1; l <= 16; l++) {
        d->valoff[idx][l] = k - code;
        k += bits[l - 1];
        code += bits[l - 1];
        d->maxcode[idx][l] = bits[l - 1] ? code - 1 : -1;
        code <<= 1;
    }
    d->maxcode[idx][17] = 0xFFFFF;
    if (k != nvals || k > 256) return -1;
    memcpy(d->huffval[idx], vals, (size_t)k);
    d->huf_set[idx] = 1;
    return 0;
}

static int parse_headers(const uint8_t *f, size_t n, dec_t *d)
{
    memset(d, 0, sizeof(*d));
    if (n < 4 || f[0] != 0xFF || f[1] != 0xD8) return ORC_DEC_CORRUPT;
    size_t p = 2;
    int have_sof = 0, comp_id[3] = {0, 0, 0};
    while (p + 4 <= n) {
        if (f[p] != 0xFF) return ORC_DEC_CORRUPT;
        while (p < n && f[p] == 0xFF) p++;                   /* fill bytes */
        if (p >= n) return ORC_DEC_CORRUPT;
        const int m = f[p++];
        if (m == 0xD8 || (m >= 0xD0 && m <= 0xD7) || m == 0x01) continue;
        if (m == 0xD9) return ORC_DEC_CORRUPT;
        if (p + 2 > n) return ORC_DEC_CORRUPT;
        const size_t len = ((size_t)f[p] << 8) | f[p + 1];
        if (len < 2 || p + len > n) return ORC_DEC_CORRUPT;
        const uint8_t *s = f + p + 2;
        const size_t sl = len - 2;
        if (m == 0xDB) {                                      /* DQT */
            size_t o = 0;
            while (o < sl) {
                const int pq = s[o] >> 4, tq = s[o] & 15;
                if (tq > 3) return ORC_DEC_CORRUPT;
                if (pq != 0) return ORC_DEC_UNSUPPORTED;
                if (o + 65 > sl) return ORC_DEC_CORRUPT;
                for (int i = 0; i < 64; i++) d->qt[tq][k_zz[i]] = s[o + 1 + i];
                d->qt_set[tq] = 1;
                o += 65;
            }
        } else if (m == 0xC4) {                               /* DHT */
            size_t o = 0;
            while (o < sl) {
                if (o + 17 > sl) return ORC_DEC_CORRUPT;
                const int tc = s[o] >> 4, th = s[o] & 15;
                if (tc > 1 || th > 1) return ORC_DEC_UNSUPPORTED;     /* baseline: two tables per class */
                int cnt = 0;
                for (int i = 0; i < 16; i++) cnt += s[o + 1 + i];
                if (o + 17 + (size_t)cnt > sl) return ORC_DEC_CORRUPT;
                if (build_huff(d, tc * 2 + th, s + o + 1, s + o + 17, cnt)) return ORC_DEC_CORRUPT;
                o += 17 + (size_t)cnt;
            }
        } else if (m == 0xC0) {                               /* SOF0 */
            if (sl < 6 + 9 || s[0] != 8 || s[5] != 3) return ORC_DEC_UNSUPPORTED;
            d->h = (s[1] << 8) | s[2];
            d->w = (s[3] << 8) | s[4];
            if (d->w < 1 || d->h < 1) return ORC_DEC_UNSUPPORTED;
            for (int c = 0; c < 3; c++) {
                comp_id[c] = s[6 + 3 * c];
                const int hv = s[7 + 3 * c];
                if (c == 0) {
                    if (hv != 0x22 && hv != 0x21 && hv != 0x11) return ORC_DEC_UNSUPPORTED;
                    d->hs = hv >> 4; d->vs = hv & 15;
                } else if (hv != 0x11) return ORC_DEC_UNSUPPORTED;
                d->comp_tq[c] = s[8 + 3 * c] & 3;
            }
            have_sof = 1;
        } else if (m >= 0xC1 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
            return ORC_DEC_UNSUPPORTED;                       /* extended, progressive, lossless, arithmetic */
        } else if (m == 0xDD) {                               /* DRI */
            if (sl < 2) return ORC_DEC_CORRUPT;
            d->ri = (s[0] << 8) | s[1];
        } else if (m == 0xDA) {                               /* SOS */
            if (!have_sof) return ORC_DEC_CORRUPT;
            if (sl < 1 + 6 + 3 || s[0] != 3) return ORC_DEC_UNSUPPORTED;
            for (int c = 0; c < 3; c++) {
                if (s[1 + 2 * c] != comp_id[c]) return ORC_DEC_UNSUPPORTED;      /* scan order = frame order */
                d->comp_td[c] = s[2 + 2 * c] >> 4;
                d->comp_ta[c] = s[2 + 2 * c] & 15;
                if (d->comp_td[c] > 1 || d->comp_ta[c] > 1) return ORC_DEC_UNSUPPORTED;
                if (!d->huf_set[d->comp_td[c]] || !d->huf_set[2 + d->comp_ta[c]] || !d->qt_set[d->comp_tq[c]])
                    return ORC_DEC_CORRUPT;
            }
            if (s[7] != 0 || s[8] != 63 || s[9] != 0) return ORC_DEC_UNSUPPORTED;
            d->scan_off = p + len;
            d->mcu_cols = (d->w + 8 * d->hs - 1) / (8 * d->hs);
            d->mcu_rows = (d->h + 8 * d->vs - 1) / (8 * d->vs);
            return ORC_DEC_OK;
        }
        p += len;
    }
    return ORC_DEC_CORRUPT;
}

/* ---- jdhuff.c bit reader: 0xFF 0x00 -> 0xFF, any other marker ends the data (zeros are fed) ---- */
typedef struct {
    const uint8_t *f;
    size_t n, p;
    unsigned long long acc;
    int nbits, hit_marker;
} bits_t;

static void fill(bits_t *b)
{
    while (b->nbits <= 56) {
        int c = 0;
        if (!b->hit_marker && b->p < b->n) {
            c = b->f[b->p++];
            if (c == 0xFF) {
                int c2 = b->p < b->n ? b->f[b->p] : 0xD9;
                if (c2 == 0) b->p++;
                else { b->hit_marker = 1; c = 0; }
            }
        }
        b->acc |= (unsigned long long)c << (56 - b->nbits);
        b->nbits += 8;
    }
}

static int getbits(bits_t *b, int k)
{
    if (!k) return 0;
    if (b->nbits < k) fill(b);
    const int v = (int)(b->acc >> (64 - k));
    b->acc <<= k;
    b->nbits -= k;
    return v;
}

static int decode_sym(bits_t *b, const dec_t *d, int idx)
{
    if (b->nbits < 17) fill(b);
    int code = 0;
    for (int l = 1; l <= 16; l++) {
        code = (code << 1) | (int)(b->acc >> 63);
        b->acc <<= 1;
        b->nbits--;
        if (code <= d->maxcode[idx][l]) return d->huffval[idx][(code + d->valoff[idx][l]) & 255];
    }
    return 0;                                                 /* jdhuff.c: "corrupt data: fake a zero" */
}

#define EXTEND(r, s) ((r) < (1 << ((s) - 1)) ? (r) - (1 << (s)) + 1 : (r))

/* jdhuff.c decode_mcu, one block: coefficients in natural order */
static void decode_block(bits_t *b, const dec_t *d, int comp, int *last_dc, short blk[64])
{
    memset(blk, 0, 64 * sizeof(short));
    int s = decode_sym(b, d, d->comp_td[comp]);
    if (s) {
        const int r = getbits(b, s);
        s = EXTEND(r, s);
    }
    *last_dc += s;
    blk[0] = (short)*last_dc;
    for (int k = 1; k < 64; k++) {
        const int rs = decode_sym(b, d, 2 + d->comp_ta[comp]);
        const int r = rs >> 4;
        s = rs & 15;
        if (s) {
            k += r;
            const int v = getbits(b, s);
            if (k < 64) blk[k_zz[k]] = (short)EXTEND(v, s);
        } else {
            if (r != 15) break;
            k += 15;
        }
    }
}

/* ---- jidctint.c jpeg_idct_islow (dequantisation folded in, jddctmgr.c) ---- */
#define CB 13
#define P1 2
#define DESCALE(x, n) (((x) + (1 << ((n) - 1))) >> (n))

static unsigned char range_limit_idct(int x)          /* jdmaster.c prepare_range_limit_table, index x & RANGE_MASK */
{
    const int v = x & 1023;
    return (unsigned char)(v < 128 ? v + 128 : (v < 512 ? 255 : (v < 896 ? 0 : v - 896)));
}

static void idct_1d(const int in[8], int out[8], int shift, int pass1)
{
    int z1, z2, z3, z4, z5, tmp0, tmp1, tmp2, tmp3, tmp10, tmp11, tmp12, tmp13;
    z2 = in[2]; z3 = in[6];
    z1 = (z2 + z3) * 4433;
    tmp2 = z1 + z3 * (-15137);
    tmp3 = z1 + z2 * 6270;
    z2 = in[0]; z3 = in[4];
    tmp0 = (z2 + z3) << CB;
    tmp1 = (z2 - z3) << CB;
    tmp10 = tmp0 + tmp3; tmp13 = tmp0 - tmp3; tmp11 = tmp1 + tmp2; tmp12 = tmp1 - tmp2;
    tmp0 = in[7]; tmp1 = in[5]; tmp2 = in[3]; tmp3 = in[1];
    z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2; z4 = tmp1 + tmp3;
    z5 = (z3 + z4) * 9633;
    tmp0 *= 2446; tmp1 *= 16819; tmp2 *= 25172; tmp3 *= 12299;
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
    z3 += z5; z4 += z5;
    tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
    (void)pass1;
    out[0] = DESCALE(tmp10 + tmp3, shift); out[7] = DESCALE(tmp10 - tmp3, shift);
    out[1] = DESCALE(tmp11 + tmp2, shift); out[6] = DESCALE(tmp11 - tmp2, shift);
    out[2] = DESCALE(tmp12 + tmp1, shift); out[5] = DESCALE(tmp12 - tmp1, shift);
    out[3] = DESCALE(tmp13 + tmp0, shift); out[4] = DESCALE(tmp13 - tmp0, shift);
}

/* (the zero-column / zero-row short cuts of jidctint.c give the same values as the full butterflies:
 * DESCALE(dc << 13, 11) = dc << 2 and DESCALE(ws0 << 13, 18) = DESCALE(ws0, 5); they are left out) */
static void idct_islow(const short blk[64], const unsigned short qt[64], unsigned char *out, int pitch)
{
    int ws[64], in[8], o[8];
    for (int c = 0; c < 8; c++) {
        for (int r = 0; r < 8; r++) in[r] = (int)blk[8 * r + c] * (int)(short)qt[8 * r + c];
        idct_1d(in, o, CB - P1, 1);
        for (int r = 0; r < 8; r++) ws[8 * r + c] = o[r];
    }
    for (int r = 0; r < 8; r++) {
        idct_1d(ws + 8 * r, o, CB + P1 + 3, 0);
        for (int c = 0; c < 8; c++) out[r * pitch + c] = range_limit_idct(o[c]);
    }
}

int orc_jpeg_info(const uint8_t *file, size_t n, int *w, int *h)
{
    dec_t d;
    const int rc = parse_headers(file, n, &d);
    if (rc) return rc;
    if (w) *w = d.w;
    if (h) *h = d.h;
    return ORC_DEC_OK;
}

/* Coefficients as the Huffman decoder leaves them (DC prediction undone, NOT dequantised): MCU-interleaved
 * block order (4:2:0: Y00 Y01 Y10 Y11 Cb Cr; 4:2:2: Y0 Y1 Cb Cr; 4:4:4: Y Cb Cr), 64 int16 per block in NATURAL
 * order.  out holds 64 * blocks-per-MCU * mcus entries. */
int orc_jpeg_decode_coefficients(const uint8_t *file, size_t n, int16_t *out)
{
    dec_t d;
    const int rc = parse_headers(file, n, &d);
    if (rc) return rc;
    bits_t b = {file, n, d.scan_off, 0, 0, 0};
    int last_dc[3] = {0, 0, 0};
    const long nm = (long)d.mcu_cols * d.mcu_rows;
    const int ny = d.hs * d.vs, nb = ny + 2;
    for (long m = 0; m < nm; m++) {
        if (d.ri && m && m % d.ri == 0) {
            /* jdhuff.c process_restart: drop the padding bits, read the RSTn marker, reset the DC predictions */
            b.acc = 0; b.nbits = 0;
            if (b.hit_marker) b.p++;                          /* fill() stopped on the 0xFF: p stands on the marker code */
            else {
                while (b.p + 1 < b.n && !(b.f[b.p] == 0xFF && b.f[b.p + 1] >= 0xD0 && b.f[b.p + 1] <= 0xD7)) b.p++;
                b.p += 2;
            }
            b.hit_marker = 0;
            last_dc[0] = last_dc[1] = last_dc[2] = 0;
        }
        for (int k = 0; k < nb; k++) {
            const int comp = k < ny ? 0 : k - ny + 1;
            decode_block(&b, &d, comp, &last_dc[comp], out + (m * nb + k) * 64);
        }
    }
    return ORC_DEC_OK;
}

int orc_jpeg_sampling(const uint8_t *file, size_t n, int *hs, int *vs)
{
    dec_t d;
    const int rc = parse_headers(file, n, &d);
    if (rc) return rc;
    if (hs) *hs = d.hs;
    if (vs) *vs = d.vs;
    return ORC_DEC_OK;
}

int orc_jpeg_decode(const uint8_t *file, size_t n, int fmt, uint8_t *dst, int dpitch)
{
    dec_t d;
    int rc = parse_headers(file, n, &d);
    if (rc) return rc;
    if (fmt != ORC_FMT_RGB8 && fmt != ORC_FMT_BGR8) return ORC_DEC_UNSUPPORTED;
    const int yw = d.mcu_cols * 8 * d.hs, yh = d.mcu_rows * 8 * d.vs, cw = d.mcu_cols * 8, ch = d.mcu_rows * 8;
    const long nm = (long)d.mcu_cols * d.mcu_rows;
    const int ny = d.hs * d.vs, nb = ny + 2;
    int16_t *coef = (int16_t *)malloc((size_t)nm * nb * 64 * sizeof(int16_t));
    unsigned char *Y = (unsigned char *)malloc((size_t)yw * yh), *Cb = (unsigned char *)malloc((size_t)cw * ch),
                  *Cr = (unsigned char *)malloc((size_t)cw * ch);
    if (!coef || !Y || !Cb || !Cr) { free(coef); free(Y); free(Cb); free(Cr); return ORC_DEC_CORRUPT; }
    rc = orc_jpeg_decode_coefficients(file, n, coef);
    for (long m = 0; m < nm && !rc; m++) {
        const int mx = (int)(m % d.mcu_cols), my = (int)(m / d.mcu_cols);
        for (int k = 0; k < ny; k++)
            idct_islow(coef + (m * nb + k) * 64, d.qt[d.comp_tq[0]],
                       Y + (size_t)(my * 8 * d.vs + (k / d.hs) * 8) * yw + mx * 8 * d.hs + (k % d.hs) * 8, yw);
        idct_islow(coef + (m * nb + ny) * 64, d.qt[d.comp_tq[1]], Cb + (size_t)(my * 8) * cw + mx * 8, cw);
        idct_islow(coef + (m * nb + ny + 1) * 64, d.qt[d.comp_tq[2]], Cr + (size_t)(my * 8) * cw + mx * 8, cw);
    }
    /* jdsample.c jinit_upsampler: h2v2 / h2v1 fancy only when the downsampled width exceeds 2, else pixel
     * replication (h2v2_upsample / h2v1_upsample); 1x1: fullsize_upsample */
    const int dsw = (d.w + d.hs - 1) / d.hs, dsh = (d.h + d.vs - 1) / d.vs;
    const int fancy = dsw > 2;
    for (int y = 0; y < d.h && !rc; y++) {
        const int cy = d.vs == 2 ? y >> 1 : y;
        /* h2v2: jdmainct.c context rows: above the first / below the last real row = that row */
        int cn = (y & 1) ? cy + 1 : cy - 1;
        if (cn < 0) cn = 0;
        if (cn > dsh - 1) cn = dsh - 1;
        for (int x = 0; x < d.w; x++) {
            const int cx = d.hs == 2 ? x >> 1 : x;
            int cbv, crv;
            if (d.hs == 1 || !fancy) {
                cbv = Cb[(size_t)cy * cw + cx];
                crv = Cr[(size_t)cy * cw + cx];
            } else if (d.vs == 1) {
                /* h2v1_fancy_upsample: 3/4 nearer + 1/4 further sample, rounding 1 (left neighbour) / 2 (right);
                 * the first and the last output column copy their sample */
                const unsigned char *planes[2] = {Cb, Cr};
                int res[2];
                for (int pl = 0; pl < 2; pl++) {
                    const unsigned char *r0 = planes[pl] + (size_t)cy * cw;
                    if (!(x & 1)) res[pl] = cx == 0 ? r0[0] : (3 * r0[cx] + r0[cx - 1] + 1) >> 2;
                    else res[pl] = cx == dsw - 1 ? r0[cx] : (3 * r0[cx] + r0[cx + 1] + 2) >> 2;
                }
                cbv = res[0]; crv = res[1];
            } else {
                const unsigned char *planes[2] = {Cb, Cr};
                int res[2];
                for (int pl = 0; pl < 2; pl++) {
                    const unsigned char *r0 = planes[pl] + (size_t)cy * cw, *r1 = planes[pl] + (size_t)cn * cw;
                    const int cur = 3 * r0[cx] + r1[cx];
                    if (!(x & 1)) {
                        if (cx == 0) res[pl] = (cur * 4 + 8) >> 4;
                        else res[pl] = (cur * 3 + (3 * r0[cx - 1] + r1[cx - 1]) + 8) >> 4;
                    } else {
                        if (cx == dsw - 1) res[pl] = (cur * 4 + 7) >> 4;
                        else res[pl] = (cur * 3 + (3 * r0[cx + 1] + r1[cx + 1]) + 7) >> 4;
                    }
                }
                cbv = res[0]; crv = res[1];
            }
            /* jdcolor.c build_ycc_rgb_table + ycc_rgb_convert */
            const int yy = Y[(size_t)y * yw + x], cbx = cbv - 128, crx = crv - 128;
            int r = yy + ((91881 * crx + 32768) >> 16);
            int bl = yy + ((116130 * cbx + 32768) >> 16);
            int g = yy + ((-22554 * cbx + 32768 + -46802 * crx) >> 16);
            r = r < 0 ? 0 : (r > 255 ? 255 : r);
            g = g < 0 ? 0 : (g > 255 ? 255 : g);
            bl = bl < 0 ? 0 : (bl > 255 ? 255 : bl);
            unsigned char *o = dst + (size_t)y * dpitch + 3 * x;
            if (fmt == ORC_FMT_BGR8) { o[0] = (unsigned char)bl; o[1] = (unsigned char)g; o[2] = (unsigned char)r; }
            else { o[0] = (unsigned char)r; o[1] = (unsigned char)g; o[2] = (unsigned char)bl; }
        }
    }
    free(coef); free(Y); free(Cb); free(Cr);
    return rc;
}
