// jpeg_parse.h -- host side of the GPU JPEG reader (k_decode.cuh): marker parse of a JFIF file (jdmarker.c's
// part of what core/ImageCompression.java:119-155 has ImageIO do) and the decode tables the kernels use.
// Takes exactly what the kernels decode -- SOF0, 8 bit, three components, luma sampled 2x2 / 2x1 / 1x1 over 1x1 chroma, 8-bit
// quantisation tables, one interleaved scan over the full spectrum, with or without restart intervals -- and says
// "unsupported" for everything else, so that the caller keeps the stock host decode for such files.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cstring>

#include "k_decode.cuh"

namespace icb {

enum { JP_OK = 0, JP_UNSUPPORTED = 1, JP_CORRUPT = 2 };

struct JpegHeader {
    int w = 0, h = 0;
    int ri = 0;                             // restart interval in MCUs (DRI); 0 = none
    int hs = 2, vs = 2;                     // luma sampling factors: 2x2 (4:2:0), 2x1 (4:2:2) or 1x1 (4:4:4); chroma 1x1
    size_t scan_off = 0;                    // first entropy-coded byte
    int comp_tq[3] = {0, 0, 0}, comp_td[3] = {0, 0, 0}, comp_ta[3] = {0, 0, 0};
    unsigned char qt[4][64];                // zigzag order as stored in the file
    bool qt_set[4] = {false, false, false, false};
    unsigned char bits[4][16];              // index = class * 2 + id
    unsigned char vals[4][256];
    int nvals[4] = {0, 0, 0, 0};
    bool huf_set[4] = {false, false, false, false};
};

inline int jpeg_parse(const uint8_t *f, size_t n, JpegHeader &hd)
{
    if (!f || n < 4 || f[0] != 0xFF || f[1] != 0xD8) return JP_CORRUPT;
    size_t p = 2;
    bool have_sof = false;
    int comp_id[3] = {0, 0, 0};
    while (p + 4 <= n) {
        if (f[p] != 0xFF) return JP_CORRUPT;
        while (p < n && f[p] == 0xFF) p++;
        if (p >= n) return JP_CORRUPT;
        const int m = f[p++];
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;      // stand-alone markers
        if (m == 0xD9 || p + 2 > n) return JP_CORRUPT;
        const size_t len = ((size_t)f[p] << 8) | f[p + 1];
        if (len < 2 || p + len > n) return JP_CORRUPT;
        const uint8_t *s = f + p + 2;
        const size_t sl = len - 2;
        switch (m) {
        case 0xDB:
            for (size_t o = 0; o < sl; o += 65) {
                const int pq = s[o] >> 4, tq = s[o] & 15;
                if (tq > 3) return JP_CORRUPT;
                if (pq) return JP_UNSUPPORTED;                                  // 16-bit tables
                if (o + 65 > sl) return JP_CORRUPT;
                memcpy(hd.qt[tq], s + o + 1, 64);
                hd.qt_set[tq] = true;
            }
            break;
        case 0xC4:
            for (size_t o = 0; o < sl;) {
                if (o + 17 > sl) return JP_CORRUPT;
                const int tc = s[o] >> 4, th = s[o] & 15;
                if (tc > 1 || th > 1) return JP_UNSUPPORTED;
                int cnt = 0;
                for (int i = 0; i < 16; i++) cnt += s[o + 1 + i];
                if (cnt > 256 || o + 17 + (size_t)cnt > sl) return JP_CORRUPT;
                const int t = tc * 2 + th;
                memcpy(hd.bits[t], s + o + 1, 16);
                memcpy(hd.vals[t], s + o + 17, (size_t)cnt);
                hd.nvals[t] = cnt;
                hd.huf_set[t] = true;
                o += 17 + (size_t)cnt;
            }
            break;
        case 0xC0:
            if (sl < 15 || s[0] != 8 || s[5] != 3) return JP_UNSUPPORTED;
            hd.h = (s[1] << 8) | s[2];
            hd.w = (s[3] << 8) | s[4];
            if (hd.w < 1 || hd.h < 1) return JP_UNSUPPORTED;
            for (int c = 0; c < 3; c++) {
                comp_id[c] = s[6 + 3 * c];
                const int hv = s[7 + 3 * c];
                if (c == 0) {
                    if (hv != 0x22 && hv != 0x21 && hv != 0x11) return JP_UNSUPPORTED;     // 4:2:0, 4:2:2, 4:4:4
                    hd.hs = hv >> 4; hd.vs = hv & 15;
                } else if (hv != 0x11) return JP_UNSUPPORTED;
                hd.comp_tq[c] = s[8 + 3 * c] & 3;
            }
            have_sof = true;
            break;
        case 0xDD:
            if (sl < 2) return JP_CORRUPT;
            hd.ri = (s[0] << 8) | s[1];
            break;
        case 0xDA:
            if (!have_sof) return JP_CORRUPT;
            if (sl < 10 || s[0] != 3) return JP_UNSUPPORTED;
            for (int c = 0; c < 3; c++) {
                if (s[1 + 2 * c] != comp_id[c]) return JP_UNSUPPORTED;
                hd.comp_td[c] = s[2 + 2 * c] >> 4;
                hd.comp_ta[c] = s[2 + 2 * c] & 15;
                if (hd.comp_td[c] > 1 || hd.comp_ta[c] > 1) return JP_UNSUPPORTED;
                if (!hd.huf_set[hd.comp_td[c]] || !hd.huf_set[2 + hd.comp_ta[c]] || !hd.qt_set[hd.comp_tq[c]]) return JP_CORRUPT;
            }
            if (s[7] != 0 || s[8] != 63 || s[9] != 0) return JP_UNSUPPORTED;
            hd.scan_off = p + len;
            return hd.scan_off < n ? JP_OK : JP_CORRUPT;
        default:
            if (m >= 0xC1 && m <= 0xCF && m != 0xC8 && m != 0xCC) return JP_UNSUPPORTED;   // every other frame type
            break;                                                              // APPn, COM, ...: skipped
        }
        p += len;
    }
    return JP_CORRUPT;
}

// jdhuff.c jpeg_make_d_derived_tbl, plus the look-up by the next DEC_LUT_BITS bits
inline void jpeg_build_tables(const JpegHeader &hd, DecTables &t)
{
    static const unsigned char zz[64] = {
        0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
        35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
    memset(&t, 0, sizeof(t));
    for (int tb = 0; tb < 4; tb++) {
        if (!hd.huf_set[tb]) continue;
        int code = 0, k = 0, n2 = 0;
        for (int l = 1; l <= 16; l++) {
            t.valoff[tb][l] = k - code;
            for (int i = 0; i < hd.bits[tb][l - 1]; i++, k++, code++) {
                const unsigned int ent = dec_entry(tb >> 1, (unsigned int)l, hd.vals[tb][k]);
                if (l <= DEC_LUT_BITS) {
                    const int lo = code << (DEC_LUT_BITS - l), cnt = 1 << (DEC_LUT_BITS - l);
                    for (int j = 0; j < cnt && lo + j < (1 << DEC_LUT_BITS); j++) t.lut[tb][lo + j] = ent;
                } else {
                    const int pre = code >> (l - DEC_LUT_BITS);
                    if (pre >= (1 << DEC_LUT_BITS)) continue;                     // an over-subscribed table: the loop's business
                    unsigned int &slot = t.lut[tb][pre];
                    if (slot == 0 && n2 < DEC_LUT2) slot = 0x80000000u | (unsigned int)n2++;
                    if (!(slot & 0x80000000u)) continue;                          // no second-level table left: the loop decodes it
                    const int sub = (code << (16 - l)) & 63, cnt = 1 << (16 - l);
                    for (int j = 0; j < cnt && sub + j < 64; j++) t.lut2[tb][slot & 0xff][sub + j] = ent;
                }
            }
            t.maxcode[tb][l] = hd.bits[tb][l - 1] ? code - 1 : -1;
            code <<= 1;
        }
        t.maxcode[tb][17] = 0xFFFFF;
        memcpy(t.huffval[tb], hd.vals[tb], 256);
    }
    for (int c = 0; c < 3; c++)
        for (int i = 0; i < 64; i++) t.qt[c][zz[i]] = hd.qt[hd.comp_tq[c]][i];
    const int ny = hd.hs * hd.vs;
    for (int k = 0; k < ny + 2; k++) {
        const int c = k < ny ? 0 : k - ny + 1;
        t.tab_dc[k] = (unsigned char)hd.comp_td[c];
        t.tab_ac[k] = (unsigned char)(2 + hd.comp_ta[c]);
    }
}

}  // namespace icb
