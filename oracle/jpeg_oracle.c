/*
 * jpeg_oracle.c -- scalar CPU restatement of the baseline JFIF encode the
 * reference reaches through javax.imageio (ImageCompressionJpg.java:136-147)
 * plus the reference's own search / scale loops.  TEST INFRASTRUCTURE ONLY
 * (see oracle.h for the scope, the third-party pin and the parity status).
 *
 * Section markers a2.1 .. a2.8 follow SURVEY.md section 8(a).
 */
#include "oracle.h"
#include "jpeg_tables.h"

#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ a2.1 */
/* JDK JPEG.convertToLinearQuality + JPEGQTable.getScaledInstance(s, true)
 * [JDK-mem]: float32 throughout, (int) truncation, clamp to [1,255].
 * Alternative profile: IJG jcparam.c jpeg_quality_scaling/jpeg_add_quant_table
 * on quality = round(q*100). */
void orc_build_qtables(float q, const orc_profile *prof, uint8_t lum[64], uint8_t chr[64])
{
    int mode = prof ? prof->table_mode : 0;
    if (mode == 0) {
        volatile float s;
        float qq = q;
        if (qq <= 0.0f) qq = 0.01f;
        if (qq > 1.00f) qq = 1.00f;
        if (qq < 0.5f) s = 0.5f / qq; else s = 2.0f - (qq * 2.0f);
        for (int i = 0; i < 64; i++) {
            volatile float a = (float)k_base_lum[i] * s;   /* rounded to float32 */
            volatile float b = a + 0.5f;
            int sv = (int)b;
            if (sv < 1) sv = 1;
            if (sv > 255) sv = 255;
            lum[i] = (uint8_t)sv;
            a = (float)k_base_chr[i] * s;
            b = a + 0.5f;
            sv = (int)b;
            if (sv < 1) sv = 1;
            if (sv > 255) sv = 255;
            chr[i] = (uint8_t)sv;
        }
    } else {
        int quality = (int)(q * 100.0f + 0.5f);
        if (quality <= 0) quality = 1;
        if (quality > 100) quality = 100;
        int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
        for (int i = 0; i < 64; i++) {
            long t = ((long)k_base_lum[i] * scale + 50L) / 100L;
            if (t <= 0) t = 1;
            if (t > 255) t = 255;
            lum[i] = (uint8_t)t;
            t = ((long)k_base_chr[i] * scale + 50L) / 100L;
            if (t <= 0) t = 1;
            if (t > 255) t = 255;
            chr[i] = (uint8_t)t;
        }
    }
}

/* IJG jcmaster.c initial_setup / per_scan_setup for 3 components 2x2,1x1,1x1 */
void orc_geometry(int w, int h, orc_geom *g)
{
    g->w = w;
    g->h = h;
    g->mcu_cols = (w + 15) / 16;
    g->mcu_rows = (h + 15) / 16;
    g->yblk_w = (w + 7) / 8;
    g->yblk_h = (h + 7) / 8;
    g->cblk_w = ((w + 1) / 2 + 7) / 8;
    g->cblk_h = ((h + 1) / 2 + 7) / 8;
    g->nblocks = 6L * g->mcu_cols * g->mcu_rows;
}

/* ------------------------------------------------------------ a2.2, a2.3 */
/* IJG jccolor.c rgb_ycc_convert (SCALEBITS 16) into edge-replicated planes,
 * then jcsample.c h2v2_downsample with expand_right_edge on the full-size
 * chroma and jcprepct.c expand_bottom_edge: an odd last input row is paired
 * with itself; rows past the last DOWNSAMPLED row replicate that row. */
typedef struct {
    int pw, ph;          /* padded luma plane dims (multiples of 16) */
    uint8_t *y, *cb, *cr;
} planes_t;

static void fmt_offsets(int fmt, int *bpp, int *ro, int *go, int *bo)
{
    switch (fmt) {
    case ORC_FMT_BGR8:  *bpp = 3; *ro = 2; *go = 1; *bo = 0; break;
    case ORC_FMT_RGBA8: *bpp = 4; *ro = 0; *go = 1; *bo = 2; break;
    case ORC_FMT_ABGR8: *bpp = 4; *ro = 3; *go = 2; *bo = 1; break;
    default:            *bpp = 3; *ro = 0; *go = 1; *bo = 2; break;
    }
}

static int planes_build(planes_t *p, const uint8_t *px, int w, int h, int pitch, int fmt)
{
    orc_geom g;
    orc_geometry(w, h, &g);
    int pw = g.mcu_cols * 16, ph = g.mcu_rows * 16;
    int cw = pw / 2, chh = ph / 2;
    p->pw = pw;
    p->ph = ph;
    p->y = (uint8_t *)malloc((size_t)pw * ph);
    p->cb = (uint8_t *)malloc((size_t)cw * chh);
    p->cr = (uint8_t *)malloc((size_t)cw * chh);
    uint8_t *fcb = (uint8_t *)malloc((size_t)pw * 2);
    uint8_t *fcr = (uint8_t *)malloc((size_t)pw * 2);
    if (!p->y || !p->cb || !p->cr || !fcb || !fcr) return -1;
    int bpp, ro, go, bo;
    fmt_offsets(fmt, &bpp, &ro, &go, &bo);

    int real_crows = (h + 1) / 2;
    for (int cy = 0; cy < real_crows; cy++) {
        for (int k = 0; k < 2; k++) {
            int yy = 2 * cy + k;
            int sy = yy < h ? yy : h - 1;          /* odd tail row pairs with itself */
            const uint8_t *row = px + (size_t)sy * pitch;
            uint8_t *yo = p->y + (size_t)yy * pw;
            uint8_t *cbo = fcb + (size_t)k * pw, *cro = fcr + (size_t)k * pw;
            for (int x = 0; x < w; x++) {
                int r = row[x * bpp + ro], gg = row[x * bpp + go], b = row[x * bpp + bo];
                yo[x] = (uint8_t)((19595 * r + 38470 * gg + 7471 * b + 32768) >> 16);
                cbo[x] = (uint8_t)((-11059 * r - 21709 * gg + 32768 * b + (128 << 16) + 32767) >> 16);
                cro[x] = (uint8_t)((32768 * r - 27439 * gg - 5329 * b + (128 << 16) + 32767) >> 16);
            }
            for (int x = w; x < pw; x++) {           /* expand_right_edge */
                yo[x] = yo[w - 1];
                cbo[x] = cbo[w - 1];
                cro[x] = cro[w - 1];
            }
        }
        uint8_t *cbd = p->cb + (size_t)cy * cw, *crd = p->cr + (size_t)cy * cw;
        int bias = 1;                                  /* 1,2,1,2,... per output column */
        for (int x = 0; x < cw; x++) {
            cbd[x] = (uint8_t)((fcb[2 * x] + fcb[2 * x + 1] + fcb[pw + 2 * x] + fcb[pw + 2 * x + 1] + bias) >> 2);
            crd[x] = (uint8_t)((fcr[2 * x] + fcr[2 * x + 1] + fcr[pw + 2 * x] + fcr[pw + 2 * x + 1] + bias) >> 2);
            bias ^= 3;
        }
    }
    /* expand_bottom_edge on the downsampled output (luma rows too) */
    for (int yy = 2 * real_crows; yy < ph; yy++)
        memcpy(p->y + (size_t)yy * pw, p->y + (size_t)(2 * real_crows - 1) * pw, (size_t)pw);
    for (int cy = real_crows; cy < chh; cy++) {
        memcpy(p->cb + (size_t)cy * cw, p->cb + (size_t)(real_crows - 1) * cw, (size_t)cw);
        memcpy(p->cr + (size_t)cy * cw, p->cr + (size_t)(real_crows - 1) * cw, (size_t)cw);
    }
    free(fcb);
    free(fcr);
    return 0;
}

static void planes_free(planes_t *p)
{
    free(p->y);
    free(p->cb);
    free(p->cr);
}

/* ------------------------------------------------------------------ a2.4 */
/* IJG jfdctint.c jpeg_fdct_islow: CONST_BITS 13, PASS1_BITS 2. */
#define CONST_BITS 13
#define PASS1_BITS 2
#define FIX_0_298631336 2446
#define FIX_0_390180644 3196
#define FIX_0_541196100 4433
#define FIX_0_765366865 6270
#define FIX_0_899976223 7373
#define FIX_1_175875602 9633
#define FIX_1_501321110 12299
#define FIX_1_847759065 15137
#define FIX_1_961570560 16069
#define FIX_2_053119869 16819
#define FIX_2_562915447 20995
#define FIX_3_072711026 25172
#define DESCALE(x, n) (((x) + (1 << ((n)-1))) >> (n))

static void fdct_islow(int *data)
{
    int tmp0, tmp1, tmp2, tmp3, tmp4, tmp5, tmp6, tmp7, tmp10, tmp11, tmp12, tmp13;
    int z1, z2, z3, z4, z5;
    int *d = data;
    for (int ctr = 0; ctr < 8; ctr++, d += 8) {
        tmp0 = d[0] + d[7]; tmp7 = d[0] - d[7];
        tmp1 = d[1] + d[6]; tmp6 = d[1] - d[6];
        tmp2 = d[2] + d[5]; tmp5 = d[2] - d[5];
        tmp3 = d[3] + d[4]; tmp4 = d[3] - d[4];
        tmp10 = tmp0 + tmp3; tmp13 = tmp0 - tmp3;
        tmp11 = tmp1 + tmp2; tmp12 = tmp1 - tmp2;
        d[0] = (tmp10 + tmp11) << PASS1_BITS;
        d[4] = (tmp10 - tmp11) << PASS1_BITS;
        z1 = (tmp12 + tmp13) * FIX_0_541196100;
        d[2] = DESCALE(z1 + tmp13 * FIX_0_765366865, CONST_BITS - PASS1_BITS);
        d[6] = DESCALE(z1 + tmp12 * (-FIX_1_847759065), CONST_BITS - PASS1_BITS);
        z1 = tmp4 + tmp7; z2 = tmp5 + tmp6; z3 = tmp4 + tmp6; z4 = tmp5 + tmp7;
        z5 = (z3 + z4) * FIX_1_175875602;
        tmp4 *= FIX_0_298631336; tmp5 *= FIX_2_053119869;
        tmp6 *= FIX_3_072711026; tmp7 *= FIX_1_501321110;
        z1 *= -FIX_0_899976223; z2 *= -FIX_2_562915447;
        z3 *= -FIX_1_961570560; z4 *= -FIX_0_390180644;
        z3 += z5; z4 += z5;
        d[7] = DESCALE(tmp4 + z1 + z3, CONST_BITS - PASS1_BITS);
        d[5] = DESCALE(tmp5 + z2 + z4, CONST_BITS - PASS1_BITS);
        d[3] = DESCALE(tmp6 + z2 + z3, CONST_BITS - PASS1_BITS);
        d[1] = DESCALE(tmp7 + z1 + z4, CONST_BITS - PASS1_BITS);
    }
    d = data;
    for (int ctr = 0; ctr < 8; ctr++, d++) {
        tmp0 = d[0] + d[56]; tmp7 = d[0] - d[56];
        tmp1 = d[8] + d[48]; tmp6 = d[8] - d[48];
        tmp2 = d[16] + d[40]; tmp5 = d[16] - d[40];
        tmp3 = d[24] + d[32]; tmp4 = d[24] - d[32];
        tmp10 = tmp0 + tmp3; tmp13 = tmp0 - tmp3;
        tmp11 = tmp1 + tmp2; tmp12 = tmp1 - tmp2;
        d[0] = DESCALE(tmp10 + tmp11, PASS1_BITS);
        d[32] = DESCALE(tmp10 - tmp11, PASS1_BITS);
        z1 = (tmp12 + tmp13) * FIX_0_541196100;
        d[16] = DESCALE(z1 + tmp13 * FIX_0_765366865, CONST_BITS + PASS1_BITS);
        d[48] = DESCALE(z1 + tmp12 * (-FIX_1_847759065), CONST_BITS + PASS1_BITS);
        z1 = tmp4 + tmp7; z2 = tmp5 + tmp6; z3 = tmp4 + tmp6; z4 = tmp5 + tmp7;
        z5 = (z3 + z4) * FIX_1_175875602;
        tmp4 *= FIX_0_298631336; tmp5 *= FIX_2_053119869;
        tmp6 *= FIX_3_072711026; tmp7 *= FIX_1_501321110;
        z1 *= -FIX_0_899976223; z2 *= -FIX_2_562915447;
        z3 *= -FIX_1_961570560; z4 *= -FIX_0_390180644;
        z3 += z5; z4 += z5;
        d[56] = DESCALE(tmp4 + z1 + z3, CONST_BITS + PASS1_BITS);
        d[40] = DESCALE(tmp5 + z2 + z4, CONST_BITS + PASS1_BITS);
        d[24] = DESCALE(tmp6 + z2 + z3, CONST_BITS + PASS1_BITS);
        d[8] = DESCALE(tmp7 + z1 + z4, CONST_BITS + PASS1_BITS);
    }
}

/* jcdctmgr.c forward_DCT: load with level shift, transform. */
static void block_fdct(const uint8_t *plane, int stride, int bx, int by, int *coef)
{
    const uint8_t *s = plane + (size_t)by * 8 * stride + (size_t)bx * 8;
    for (int r = 0; r < 8; r++)
        for (int c = 0; c < 8; c++) coef[r * 8 + c] = (int)s[(size_t)r * stride + c] - 128;
    fdct_islow(coef);
}

/* ------------------------------------------------------------------ a2.6 */
/* IJG jccoefct.c compress_data: MCU-interleaved order Y00 Y01 Y10 Y11 Cb Cr.
 * Walks every MCU and hands each block's UNQUANTISED coefficients (natural
 * order) to a sink.  Dummy blocks: AC = 0; DC copied from the block IJG names
 * (right edge: the block before it in the MCU buffer; bottom row: the last
 * block of the row above, MCU_buffer[blkn-1]).  IJG copies the quantised DC;
 * copying the unquantised one and quantising it gives the same value. */
typedef void (*block_sink)(void *ctx, int comp, const int *coef_natural);

static void walk_mcus(const planes_t *p, const orc_geom *g, block_sink sink, void *ctx)
{
    int mcu[6][64];
    int last_col_width = g->yblk_w % 2 ? g->yblk_w % 2 : 2;
    int last_row_height = g->yblk_h % 2 ? g->yblk_h % 2 : 2;
    for (int my = 0; my < g->mcu_rows; my++) {
        for (int mx = 0; mx < g->mcu_cols; mx++) {
            int blockcnt = (mx < g->mcu_cols - 1) ? 2 : last_col_width;
            int blkn = 0;
            for (int yi = 0; yi < 2; yi++) {
                if (my < g->mcu_rows - 1 || yi < last_row_height) {
                    for (int xi = 0; xi < blockcnt; xi++)
                        block_fdct(p->y, p->pw, 2 * mx + xi, 2 * my + yi, mcu[blkn + xi]);
                    for (int bi = blockcnt; bi < 2; bi++) {
                        memset(mcu[blkn + bi], 0, sizeof(mcu[0]));
                        mcu[blkn + bi][0] = mcu[blkn + bi - 1][0];
                    }
                } else {
                    for (int bi = 0; bi < 2; bi++) {
                        memset(mcu[blkn + bi], 0, sizeof(mcu[0]));
                        mcu[blkn + bi][0] = mcu[blkn - 1][0];
                    }
                }
                blkn += 2;
            }
            block_fdct(p->cb, p->pw / 2, mx, my, mcu[4]);
            block_fdct(p->cr, p->pw / 2, mx, my, mcu[5]);
            for (int b = 0; b < 6; b++) sink(ctx, b < 4 ? 0 : b - 3, mcu[b]);
        }
    }
}

typedef struct { int16_t *out; } dump_ctx;
static void dump_sink(void *vctx, int comp, const int *c)
{
    dump_ctx *d = (dump_ctx *)vctx;
    (void)comp;
    for (int k = 0; k < 64; k++) d->out[k] = (int16_t)c[k_zigzag[k]];
    d->out += 64;
}

int orc_forward_coefficients(const uint8_t *px, int w, int h, int pitch, int fmt, int16_t *out)
{
    if (!px || !out || w < 1 || h < 1) return -1;
    planes_t p;
    orc_geom g;
    orc_geometry(w, h, &g);
    if (planes_build(&p, px, w, h, pitch, fmt)) return -2;
    dump_ctx d = {out};
    walk_mcus(&p, &g, dump_sink, &d);
    planes_free(&p);
    return 0;
}

/* ------------------------------------------------------------ a2.5, a2.7 */
typedef struct { uint16_t code[256]; uint8_t size[256]; } huff_t;

/* jchuff.c jpeg_make_c_derived_tbl: canonical codes by increasing length. */
static void huff_derive(huff_t *t, const uint8_t bits[16], const uint8_t *vals)
{
    memset(t, 0, sizeof(*t));
    unsigned code = 0;
    int k = 0;
    for (int l = 1; l <= 16; l++) {
        for (int i = 0; i < bits[l - 1]; i++, k++) {
            t->code[vals[k]] = (uint16_t)code++;
            t->size[vals[k]] = (uint8_t)l;
        }
        code <<= 1;
    }
}

typedef struct {
    uint8_t *out;
    size_t cap, pos;
    uint32_t put_buffer;
    int put_bits;
    int qdiv[2][64];        /* divisors = quantval << 3, natural order (islow) */
    huff_t dc[2], ac[2];
    int last_dc[3];
} enc_t;

static inline void put_byte(enc_t *e, int v)
{
    if (e->pos < e->cap) e->out[e->pos] = (uint8_t)v;
    e->pos++;
}

/* jchuff.c emit_bits: MSB first, every 0xFF byte followed by a stuffed 0x00. */
static inline void emit_bits(enc_t *e, unsigned code, int size)
{
    uint32_t pb = code & ((1u << size) - 1);
    int bits = e->put_bits + size;
    pb <<= 24 - bits;
    pb |= e->put_buffer;
    while (bits >= 8) {
        int c = (int)((pb >> 16) & 0xFF);
        put_byte(e, c);
        if (c == 0xFF) put_byte(e, 0);
        pb <<= 8;
        bits -= 8;
    }
    e->put_buffer = pb;
    e->put_bits = bits;
}

static void encode_sink(void *vctx, int comp, const int *c)
{
    enc_t *e = (enc_t *)vctx;
    int tb = comp ? 1 : 0;
    const int *qd = e->qdiv[tb];
    int q[64];
    /* a2.5 jcdctmgr.c forward_DCT quantise: round half away from zero */
    for (int i = 0; i < 64; i++) {
        int qval = qd[i], temp = c[i];
        if (temp < 0) {
            temp = -temp;
            temp += qval >> 1;
            temp = temp >= qval ? temp / qval : 0;
            temp = -temp;
        } else {
            temp += qval >> 1;
            temp = temp >= qval ? temp / qval : 0;
        }
        q[i] = temp;
    }
    /* a2.7 jchuff.c encode_one_block */
    const huff_t *dct = &e->dc[tb], *act = &e->ac[tb];
    int temp = q[0] - e->last_dc[comp], temp2 = temp;
    e->last_dc[comp] = q[0];
    if (temp < 0) { temp = -temp; temp2--; }
    int nbits = 0;
    while (temp) { nbits++; temp >>= 1; }
    emit_bits(e, dct->code[nbits], dct->size[nbits]);
    if (nbits) emit_bits(e, (unsigned)temp2, nbits);
    int r = 0;
    for (int k = 1; k < 64; k++) {
        if ((temp = q[k_zigzag[k]]) == 0) {
            r++;
        } else {
            while (r > 15) { emit_bits(e, act->code[0xF0], act->size[0xF0]); r -= 16; }
            temp2 = temp;
            if (temp < 0) { temp = -temp; temp2--; }
            nbits = 1;
            while ((temp >>= 1)) nbits++;
            int i = (r << 4) + nbits;
            emit_bits(e, act->code[i], act->size[i]);
            emit_bits(e, (unsigned)temp2, nbits);
            r = 0;
        }
    }
    if (r > 0) emit_bits(e, act->code[0], act->size[0]);
}

/* ------------------------------------------------------------------ a2.8 */
/* jcmarker.c write_file_header/write_frame_header/write_scan_header with the
 * JDK writer's choices: JFIF APP0 (version 1.<jfif_minor>, units 0, 1x1, no
 * thumbnail), one DQT segment per table, SOF0, one DHT segment per table in
 * the order DC0 AC0 DC1 AC1, SOS.  623 bytes. */
static void put_u16(enc_t *e, int v) { put_byte(e, (v >> 8) & 0xFF); put_byte(e, v & 0xFF); }

static void put_dht(enc_t *e, int idx, const uint8_t bits[16], const uint8_t *vals, int nvals)
{
    put_u16(e, 0xFFC4);
    put_u16(e, 2 + 1 + 16 + nvals);
    put_byte(e, idx);
    for (int i = 0; i < 16; i++) put_byte(e, bits[i]);
    for (int i = 0; i < nvals; i++) put_byte(e, vals[i]);
}

static void write_headers(enc_t *e, int w, int h, const uint8_t lum[64], const uint8_t chr[64], int jfif_minor)
{
    put_u16(e, 0xFFD8);
    put_u16(e, 0xFFE0);
    put_u16(e, 16);
    put_byte(e, 'J'); put_byte(e, 'F'); put_byte(e, 'I'); put_byte(e, 'F'); put_byte(e, 0);
    put_byte(e, 1); put_byte(e, jfif_minor);
    put_byte(e, 0);
    put_u16(e, 1); put_u16(e, 1);
    put_byte(e, 0); put_byte(e, 0);
    for (int t = 0; t < 2; t++) {
        const uint8_t *tab = t ? chr : lum;
        put_u16(e, 0xFFDB);
        put_u16(e, 67);
        put_byte(e, t);
        for (int k = 0; k < 64; k++) put_byte(e, tab[k_zigzag[k]]);
    }
    put_u16(e, 0xFFC0);
    put_u16(e, 17);
    put_byte(e, 8);
    put_u16(e, h);
    put_u16(e, w);
    put_byte(e, 3);
    put_byte(e, 1); put_byte(e, 0x22); put_byte(e, 0);
    put_byte(e, 2); put_byte(e, 0x11); put_byte(e, 1);
    put_byte(e, 3); put_byte(e, 0x11); put_byte(e, 1);
    put_dht(e, 0x00, k_dc_lum_bits, k_dc_vals, 12);
    put_dht(e, 0x10, k_ac_lum_bits, k_ac_lum_vals, 162);
    put_dht(e, 0x01, k_dc_chr_bits, k_dc_vals, 12);
    put_dht(e, 0x11, k_ac_chr_bits, k_ac_chr_vals, 162);
    put_u16(e, 0xFFDA);
    put_u16(e, 12);
    put_byte(e, 3);
    put_byte(e, 1); put_byte(e, 0x00);
    put_byte(e, 2); put_byte(e, 0x11);
    put_byte(e, 3); put_byte(e, 0x11);
    put_byte(e, 0); put_byte(e, 0x3F); put_byte(e, 0);
}

int orc_jpeg_encode_tables(const uint8_t *px, int w, int h, int pitch, int fmt,
                           const uint8_t lum[64], const uint8_t chr[64], int jfif_minor,
                           uint8_t *out, size_t cap, size_t *out_len)
{
    if (!px || w < 1 || h < 1 || w > 65535 || h > 65535 || !out_len) return -1;
    planes_t p;
    orc_geom g;
    orc_geometry(w, h, &g);
    if (planes_build(&p, px, w, h, pitch, fmt)) return -2;
    enc_t *e = (enc_t *)calloc(1, sizeof(enc_t));
    if (!e) { planes_free(&p); return -2; }
    e->out = out;
    e->cap = out ? cap : 0;
    for (int i = 0; i < 64; i++) {
        e->qdiv[0][i] = lum[i] << 3;
        e->qdiv[1][i] = chr[i] << 3;
    }
    huff_derive(&e->dc[0], k_dc_lum_bits, k_dc_vals);
    huff_derive(&e->dc[1], k_dc_chr_bits, k_dc_vals);
    huff_derive(&e->ac[0], k_ac_lum_bits, k_ac_lum_vals);
    huff_derive(&e->ac[1], k_ac_chr_bits, k_ac_chr_vals);
    write_headers(e, w, h, lum, chr, jfif_minor);
    walk_mcus(&p, &g, encode_sink, e);
    emit_bits(e, 0x7F, 7);                  /* jchuff.c flush_bits: pad with 1s */
    put_u16(e, 0xFFD9);
    *out_len = e->pos;
    free(e);
    planes_free(&p);
    return 0;
}

int orc_jpeg_encode(const uint8_t *px, int w, int h, int pitch, int fmt, float q,
                    const orc_profile *prof, uint8_t *out, size_t cap, size_t *out_len)
{
    uint8_t lum[64], chr[64];
    orc_build_qtables(q, prof, lum, chr);
    return orc_jpeg_encode_tables(px, w, h, pitch, fmt, lum, chr, prof ? prof->jfif_minor : 2,
                                  out, cap, out_len);
}

/* -------------------------------------------------------------------- a1 */
/* ImageCompressionJpg.java:158-200 findBestQualityByBinarySearch, float32
 * control exactly as written: lo=0, hi=initial, <=8 iterations, break when
 * mid < 0.01f (before encoding) or hi-lo < 0.01f (after the update). */
float orc_find_best_quality(const uint8_t *px, int w, int h, int pitch, int fmt,
                            long target, float initial_q, const orc_profile *prof,
                            float *trace_q, long *trace_size, int *n_trace)
{
    float low = 0.0f, high = initial_q, best = -1.0f;
    int n = 0;
    for (int i = 0; i < 8; i++) {
        float mid = (low + high) / 2.0f;
        if (mid < 0.01f) break;
        size_t len = 0;
        if (orc_jpeg_encode(px, w, h, pitch, fmt, mid, prof, NULL, 0, &len)) break;
        if (trace_q) trace_q[n] = mid;
        if (trace_size) trace_size[n] = (long)len;
        n++;
        if ((long)len <= target) {
            best = mid;
            low = mid;
        } else {
            high = mid;
        }
        if ((high - low) < 0.01f) break;
    }
    if (n_trace) *n_trace = n;
    return best;
}

/* ------------------------------------------------------------ a3, a4, a5 */
/* ImageCompressionJpg.java:77-122 compressJpgWithTargetSize with :216-238
 * tryCachedParams and :255-260 saveCompressedImage.  File write and cache
 * mutation stay with the caller: the bytes and (best_q, scale) come back. */
static uint8_t *resized_copy(const uint8_t *px, int w, int h, int pitch, int fmt, double scale,
                             int *nw, int *nh, int *npitch)
{
    orc_resize_dims(w, h, scale, nw, nh);
    int bpp = (fmt == ORC_FMT_RGBA8 || fmt == ORC_FMT_ABGR8) ? 4 : 3;
    *npitch = *nw * bpp;
    uint8_t *dst = (uint8_t *)malloc((size_t)*npitch * *nh);
    if (!dst) return NULL;
    if (orc_resize_bilinear(px, w, h, pitch, fmt, dst, *nw, *nh, *npitch)) { free(dst); return NULL; }
    return dst;
}

int orc_jpg_fit(const uint8_t *px, int w, int h, int pitch, int fmt,
                float initial_q, long target,
                int has_cached, float cached_q, double cached_scale,
                const orc_profile *prof,
                uint8_t *out, size_t cap, size_t *out_len,
                float *best_q, double *scale_out, int *used_cache, int *n_encodes)
{
    int encodes = 0;
    if (used_cache) *used_cache = 0;
    if (n_encodes) *n_encodes = 0;
    if (!px || w < 1 || h < 1 || !out_len) return -1;

    if (has_cached) {                                   /* :82-89 -> :216-238 */
        const uint8_t *src = px;
        uint8_t *tmp = NULL;
        int cw = w, chh = h, cp = pitch;
        if (cached_scale < 1.0) {
            tmp = resized_copy(px, w, h, pitch, fmt, cached_scale, &cw, &chh, &cp);
            if (!tmp) return -2;
            src = tmp;
        }
        size_t len = 0;
        int rc = orc_jpeg_encode(src, cw, chh, cp, fmt, cached_q, prof, out, cap, &len);
        encodes++;
        free(tmp);
        if (rc) return rc;
        if ((long)len <= target) {
            *out_len = len;
            if (best_q) *best_q = cached_q;
            if (scale_out) *scale_out = cached_scale;
            if (used_cache) *used_cache = 1;
            if (n_encodes) *n_encodes = encodes;
            return 0;
        }
    }

    const double SCALE_STEP = 0.85;
    for (double scale = 1.0; scale > 0.1; scale = (scale == 1.0) ? SCALE_STEP : scale * SCALE_STEP) {
        const uint8_t *cur = px;
        uint8_t *tmp = NULL;
        int cw = w, chh = h, cp = pitch;
        if (scale < 1.0) {                               /* :97-102, from the ORIGINAL */
            tmp = resized_copy(px, w, h, pitch, fmt, scale, &cw, &chh, &cp);
            if (!tmp) return -2;
            cur = tmp;
        }
        int nt = 0;
        float best = orc_find_best_quality(cur, cw, chh, cp, fmt, target, initial_q, prof, NULL, NULL, &nt);
        encodes += nt;
        if (best > 0) {                                  /* :109-113 */
            size_t len = 0;
            int rc = orc_jpeg_encode(cur, cw, chh, cp, fmt, best, prof, out, cap, &len);
            encodes++;
            free(tmp);
            if (rc) return rc;
            *out_len = len;
            if (best_q) *best_q = best;
            if (scale_out) *scale_out = scale;
            if (n_encodes) *n_encodes = encodes;
            return 0;
        }
        free(tmp);
    }
    *out_len = 0;
    if (best_q) *best_q = -1.0f;
    if (scale_out) *scale_out = 0.0;
    if (n_encodes) *n_encodes = encodes;
    return 1;                                            /* :120-121 return false */
}
