// Host check of the transform kernel's arithmetic forms (csrc/xf_arith.cuh) against the CPU oracle.
// Test infrastructure: built and run by tests/test_xf_arith.py with g++ (dp2a / dp4a / prmt emulated
// on the host as the PTX ISA defines them).  Walks whole MCUs the way k_transform does -- 8x2-pixel
// strips into a byte sample buffer, then row pass on bytes, column pass, zigzag int16 pack -- on
// images whose sides are multiples of 16 (image edges take the kernel's scalar path, which the GPU
// parity tests cover).
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../image_compression_b200/csrc/xf_arith.cuh"
extern "C" int orc_forward_coefficients(const uint8_t *px, int w, int h, int pitch, int fmt, int16_t *out);

static const unsigned char kZZ[64] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

template <bool BGR>
static long run(const uint8_t *px, int w, int h, int pitch, const int16_t *ref)
{
    using namespace icb;
    long bad = 0;
    const int mcu_cols = w / 16, mcu_rows = h / 16;
    for (int my = 0; my < mcu_rows; my++)
        for (int mx = 0; mx < mcu_cols; mx++) {
            uint32_t samp[6][16];       // 6 blocks x 64 sample bytes, row r = words 2r, 2r+1
            for (int sr = 0; sr < 8; sr++)
                for (int xi = 0; xi < 2; xi++) {
                    int ncb[4] = {0, 0, 0, 0}, ncr[4] = {0, 0, 0, 0};
                    uint32_t y[4];
                    alignas(8) unsigned char row[2][24];
                    for (int r = 0; r < 2; r++)
                        memcpy(row[r], px + (size_t)(my * 16 + 2 * sr + r) * pitch + (mx * 16 + xi * 8) * 3, 24);
                    convert_row8<BGR>(row[0], y[0], y[1], ncb, ncr);
                    convert_row8<BGR>(row[1], y[2], y[3], ncb, ncr);
                    const int yblk = (sr >> 2) * 2 + xi, yrow = (2 * sr) & 7;
                    for (int k = 0; k < 4; k++) samp[yblk][2 * yrow + k] = y[k];
                    uint32_t cbv = 0, crv = 0;
                    for (int j = 0; j < 4; j++) {
                        cbv |= (uint32_t)((1025 + (j & 1) - ncb[j]) >> 2) << (8 * j);
                        crv |= (uint32_t)((1025 + (j & 1) - ncr[j]) >> 2) << (8 * j);
                    }
                    samp[4][2 * sr + xi] = cbv;
                    samp[5][2 * sr + xi] = crv;
                }
            for (int b = 0; b < 6; b++) {
                int d[64];
                for (int r = 0; r < 8; r++) fdct_row_bytes(samp[b][2 * r], samp[b][2 * r + 1], &d[8 * r]);
                for (int c = 0; c < 8; c++)
                    fdct_col_hi16(d[c], d[8 + c], d[16 + c], d[24 + c], d[32 + c], d[40 + c], d[48 + c], d[56 + c]);
                const int16_t *rb = ref + ((size_t)(my * mcu_cols + mx) * 6 + b) * 64;
                for (int k = 0; k < 64; k += 2) {
                    const uint32_t pk = prmt((uint32_t)d[kZZ[k]], (uint32_t)d[kZZ[k + 1]], 0x7632);
                    if ((int16_t)(pk & 0xffff) != rb[k] || (int16_t)(pk >> 16) != rb[k + 1]) bad++;
                }
                // the column-at-a-time form the kernel runs (packed int16 columns, zigzag picked by PRMT)
                uint32_t P[32], o[32];
                fdct_block_bytes(samp[b], P);
                fdct_zz_chunk<0>(P, o[0], o[1], o[2], o[3]);     fdct_zz_chunk<1>(P, o[4], o[5], o[6], o[7]);
                fdct_zz_chunk<2>(P, o[8], o[9], o[10], o[11]);   fdct_zz_chunk<3>(P, o[12], o[13], o[14], o[15]);
                fdct_zz_chunk<4>(P, o[16], o[17], o[18], o[19]); fdct_zz_chunk<5>(P, o[20], o[21], o[22], o[23]);
                fdct_zz_chunk<6>(P, o[24], o[25], o[26], o[27]); fdct_zz_chunk<7>(P, o[28], o[29], o[30], o[31]);
                for (int k = 0; k < 32; k++)
                    if ((int16_t)(o[k] & 0xffff) != rb[2 * k] || (int16_t)(o[k] >> 16) != rb[2 * k + 1]) bad++;
            }
        }
    return bad;
}

int main(int argc, char **argv)
{
    const int w = argc > 1 ? atoi(argv[1]) : 64, h = argc > 2 ? atoi(argv[2]) : 48;
    const unsigned seed = argc > 3 ? (unsigned)atoi(argv[3]) : 1u;
    const int mode = argc > 4 ? atoi(argv[4]) : 0;      // 0 noise, 1 extremes (0/255 only), 2 smooth
    if (w % 16 || h % 16) return 2;
    const int pitch = w * 3;
    std::vector<uint8_t> px((size_t)pitch * h);
    uint32_t s = seed * 2654435761u + 12345u;
    for (size_t i = 0; i < px.size(); i++) {
        s = s * 1664525u + 1013904223u;
        const uint32_t v = s >> 24;
        px[i] = mode == 0 ? (uint8_t)v : mode == 1 ? (uint8_t)((v & 1) ? 255 : 0) : (uint8_t)(((i / 3) % w * 255 / w + (v & 7)) & 255);
    }
    std::vector<int16_t> ref((size_t)(w / 16) * (h / 16) * 6 * 64);
    long bad = 0;
    for (int fmt = 0; fmt < 2; fmt++) {                 // 0 = RGB8, 1 = BGR8 (include/imgcomp_b200.h)
        if (orc_forward_coefficients(px.data(), w, h, pitch, fmt, ref.data())) return 3;
        bad += fmt ? run<true>(px.data(), w, h, pitch, ref.data()) : run<false>(px.data(), w, h, pitch, ref.data());
    }
    printf("mismatches %ld\n", bad);
    return bad ? 1 : 0;
}
