/* oracle/jpeg_oracle_enc.c -- TEST INFRASTRUCTURE, NOT PRODUCT.
 * Plain-C restatement of libjpeg-turbo 3.0.3's baseline compress path as the reference tree vendors it
 * ("ljt/" = /root/reference/external/libjpeg-turbo), for jpeg_set_quality(q, TRUE), JDCT_ISLOW, optimize_coding=FALSE,
 * JFIF header, one interleaved scan:
 *   quality -> tables      ljt/jcparam.c jpeg_quality_scaling:132-154, jpeg_add_quant_table:30-63, std tables :75-94
 *   RGB -> YCbCr           ljt/jccolor.c rgb_ycc_start:214-246, ljt/jccolext.c rgb_ycc_convert_internal:30-76
 *   edge expansion         ljt/jcsample.c expand_right_edge:99-116, ljt/jcprepct.c expand_bottom_edge:104-118, pre_process_data:126-195
 *   downsampling           ljt/jcsample.c h2v1_downsample:217-252, h2v2_downsample:264-295, fullsize_downsample:189-204
 *   forward DCT            ljt/jfdctint.c _jpeg_fdct_islow:143-288, level shift ljt/jcdctmgr.c convsamp:384-412
 *   quantisation           ljt/jcdctmgr.c compute_reciprocal:173-228 (32-bit DCTELEM, the non-SIMD build), quantize:419-452
 *   dummy blocks           ljt/jccoefct.c compress_data:142-225
 *   entropy coding         ljt/jchuff.c jpeg_make_c_derived_tbl:232-312, encode_one_block:558-655, flush_bits:480, emit_restart:663,
 *                          encode_mcu_huff:688; std Huffman tables ljt/jstdhuff.c:54-144 (ITU T.81 Annex K.3)
 *   markers                ljt/jcmarker.c write_file_header:475, write_frame_header:500, write_scan_header:563, emit_dqt:151,
 *                          emit_dht:191, emit_sof:290, emit_sos:320, emit_jfif_app0:355, emit_dri:278
 * Parity status: PINNED -- tests/test_oracle.py compares the produced files byte for byte with the compiled reference
 * (oracle/_ref, refh_ljt_encode) and with libjpeg-turbo's own golden MD5s (MD5_JPEG_420_ISLOW etc.). */
#include <stdlib.h>
#include <string.h>
#include "jpeg_oracle.h"

static const uint8_t zz_nat[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21,
    28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

static const uint8_t base_q[2][64] = {
    {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57, 69, 56, 14, 17, 22, 29, 51, 87, 80, 62,
        18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99},
    {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99, 99, 99, 47, 66, 99, 99, 99, 99, 99, 99,
        99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99}};

/* T.81 Annex K.3 typical Huffman tables: BITS (16 counts) then HUFFVAL, as hex text */
static const char* std_bits[4] = {"00010501010101010100000000000000", "00030101010101010101010000000000", "0002010303020403050504040000017d",
    "00020102040403040705040400010277"};
static const char* std_vals[4] = {"000102030405060708090a0b", "000102030405060708090a0b",
    "01020300041105122131410613516107227114328191a1082342b1c11552d1f02433627282090a161718191a25262728292a3435363738393a434445464748494a53545556"
    "5758595a636465666768696a737475767778797a838485868788898a92939495969798999aa2a3a4a5a6a7a8a9aab2b3b4b5b6b7b8b9bac2c3c4c5c6c7c8c9cad2d3d4d5"
    "d6d7d8d9dae1e2e3e4e5e6e7e8e9eaf1f2f3f4f5f6f7f8f9fa",
    "000102031104052131061241510761711322328108144291a1b1c109233352f0156272d10a162434e125f11718191a262728292a35363738393a434445464748494a5354"
    "55565758595a636465666768696a737475767778797a82838485868788898a92939495969798999aa2a3a4a5a6a7a8a9aab2b3b4b5b6b7b8b9bac2c3c4c5c6c7c8c9cad2"
    "d3d4d5d6d7d8d9dae2e3e4e5e6e7e8e9eaf2f3f4f5f6f7f8f9fa"};
/* table index: 0 DC luma, 1 DC chroma, 2 AC luma, 3 AC chroma */

typedef struct
{
    uint8_t bits[17];
    uint8_t vals[256];
    int nvals;
    uint32_t code[256];
    uint8_t size[256];
} enc_table;

static int hexv(char c) { return c <= '9' ? c - '0' : c - 'a' + 10; }

static void make_table(int which, enc_table* t)
{
    memset(t, 0, sizeof(*t));
    for (int i = 0; i < 16; ++i)
        t->bits[i + 1] = (uint8_t)(hexv(std_bits[which][2 * i]) * 16 + hexv(std_bits[which][2 * i + 1]));
    const char* v = std_vals[which];
    t->nvals = (int)strlen(v) / 2;
    for (int i = 0; i < t->nvals; ++i)
        t->vals[i] = (uint8_t)(hexv(v[2 * i]) * 16 + hexv(v[2 * i + 1]));
    /* canonical code assignment (T.81 C.2; jchuff.c:232-312) */
    uint32_t code = 0;
    int k = 0;
    for (int len = 1; len <= 16; ++len) {
        for (int i = 0; i < t->bits[len]; ++i, ++k) {
            t->code[t->vals[k]] = code++;
            t->size[t->vals[k]] = (uint8_t)len;
        }
        code <<= 1;
    }
}

typedef struct
{
    uint8_t* out;
    size_t cap, n;
    uint64_t acc;
    int nbits;
    int overflow;
} bitw;

static void put_byte(bitw* w, int b)
{
    if (w->n < w->cap)
        w->out[w->n] = (uint8_t)b;
    else
        w->overflow = 1;
    w->n++;
}
static void put_u16(bitw* w, int v)
{
    put_byte(w, v >> 8);
    put_byte(w, v & 255);
}
static void put_bits(bitw* w, uint32_t code, int size)
{
    w->acc = (w->acc << size) | (code & ((1u << size) - 1u));
    w->nbits += size;
    while (w->nbits >= 8) {
        const int b = (int)((w->acc >> (w->nbits - 8)) & 255);
        put_byte(w, b);
        if (b == 0xFF)
            put_byte(w, 0);
        w->nbits -= 8;
    }
}
static void flush_ones(bitw* w)
{
    if (w->nbits)
        put_bits(w, 0x7F, 8 - w->nbits);
    w->acc = 0;
    w->nbits = 0;
}

/* 1-D LL&M butterfly shared by both passes; pass 1 keeps 2 fractional bits, pass 2 removes them (+3 bits overall scale) */
static void fdct_1d(const int* in, int* o, int pass2)
{
    const int C = 13, P = 2;
    const int s07 = in[0] + in[7], d07 = in[0] - in[7], s16 = in[1] + in[6], d16 = in[1] - in[6];
    const int s25 = in[2] + in[5], d25 = in[2] - in[5], s34 = in[3] + in[4], d34 = in[3] - in[4];
    const int e0 = s07 + s34, e3 = s07 - s34, e1 = s16 + s25, e2 = s16 - s25;
    const int sh = pass2 ? C + P : C - P;
    const int rnd = 1 << (sh - 1);
    if (pass2) {
        o[0] = (e0 + e1 + (1 << (P - 1))) >> P;
        o[4] = (e0 - e1 + (1 << (P - 1))) >> P;
    } else {
        o[0] = (e0 + e1) * (1 << P);
        o[4] = (e0 - e1) * (1 << P);
    }
    const int r = (e2 + e3) * 4433;
    o[2] = (r + e3 * 6270 + rnd) >> sh;
    o[6] = (r - e2 * 15137 + rnd) >> sh;
    const int a = d34 + d07, b = d25 + d16, c = d34 + d16, d = d25 + d07;
    const int w = (c + d) * 9633;
    const int t4 = d34 * 2446, t5 = d25 * 16819, t6 = d16 * 25172, t7 = d07 * 12299;
    const int za = -a * 7373, zb = -b * 20995, zc = -c * 16069 + w, zd = -d * 3196 + w;
    o[7] = (t4 + za + zc + rnd) >> sh;
    o[5] = (t5 + zb + zd + rnd) >> sh;
    o[3] = (t6 + zb + zc + rnd) >> sh;
    o[1] = (t7 + za + zd + rnd) >> sh;
}

typedef struct
{
    uint32_t recip[64], corr[64];
    int shift[64];
} divisors;

static void make_divisors(const uint16_t* q, divisors* dv)
{
    for (int i = 0; i < 64; ++i) {
        const uint32_t d = (uint32_t)q[i] << 3;
        int b = 0;
        while ((d >> (b + 1)) != 0)
            ++b; /* significant bits - 1 */
        int r = 32 + b;
        uint64_t fq = ((uint64_t)1 << r) / d, fr = ((uint64_t)1 << r) % d;
        uint32_t c = d / 2;
        if (fr == 0) {
            fq >>= 1;
            r--;
        } else if (fr <= d / 2)
            c++;
        else
            fq++;
        dv->recip[i] = (uint32_t)fq;
        dv->corr[i] = c;
        dv->shift[i] = r - 32;
    }
}

static void fdct_quant(const uint8_t* samples, int pitch, const divisors* dv, int16_t* coef)
{
    int ws[64], row[8], o[8];
    for (int y = 0; y < 8; ++y) {
        for (int x = 0; x < 8; ++x)
            row[x] = (int)samples[y * pitch + x] - 128;
        fdct_1d(row, o, 0);
        for (int x = 0; x < 8; ++x)
            ws[y * 8 + x] = o[x];
    }
    for (int x = 0; x < 8; ++x) {
        for (int y = 0; y < 8; ++y)
            row[y] = ws[y * 8 + x];
        fdct_1d(row, o, 1);
        for (int y = 0; y < 8; ++y)
            ws[y * 8 + x] = o[y];
    }
    for (int i = 0; i < 64; ++i) {
        const int v = ws[i];
        const uint32_t mag = (uint32_t)(v < 0 ? -v : v);
        const uint32_t qv = (uint32_t)((((uint64_t)(mag + dv->corr[i]) * dv->recip[i]) >> 32) >> dv->shift[i]);
        coef[i] = (int16_t)(v < 0 ? -(int)qv : (int)qv);
    }
}

static int nbits_of(int v)
{
    int n = 0;
    while (v) {
        ++n;
        v >>= 1;
    }
    return n;
}

static void encode_block(bitw* w, const int16_t* c, int last_dc, const enc_table* dc, const enc_table* ac)
{
    int diff = c[0] - last_dc;
    int mag = diff < 0 ? -diff : diff, n = nbits_of(mag);
    put_bits(w, dc->code[n], dc->size[n]);
    if (n)
        put_bits(w, (uint32_t)(diff < 0 ? diff - 1 : diff), n);
    int run = 0;
    for (int k = 1; k < 64; ++k) {
        const int v = c[zz_nat[k]];
        if (v == 0) {
            ++run;
            continue;
        }
        while (run > 15) {
            put_bits(w, ac->code[0xF0], ac->size[0xF0]);
            run -= 16;
        }
        mag = v < 0 ? -v : v;
        n = nbits_of(mag);
        put_bits(w, ac->code[run * 16 + n], ac->size[run * 16 + n]);
        put_bits(w, (uint32_t)(v < 0 ? v - 1 : v), n);
        run = 0;
    }
    if (run)
        put_bits(w, ac->code[0], ac->size[0]);
}

long jo_encode(const uint8_t* pixels, int width, int height, int channels, int64_t stride, int quality, int hs, int vs,
    int restart_interval, uint8_t* out, size_t cap)
{
    if (!pixels || !out || width < 1 || height < 1 || width > 65535 || height > 65535 || (channels != 1 && channels != 3))
        return -1;
    const int ncomp = channels;
    if (ncomp == 1)
        hs = vs = 1;
    if (hs < 1 || hs > 2 || vs < 1 || vs > 2)
        return -1;
    /* quality -> quant tables (natural order here; DQT writes zigzag) */
    int q = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
    const int scale = q < 50 ? 5000 / q : 200 - q * 2;
    uint16_t qt[2][64];
    for (int t = 0; t < 2; ++t)
        for (int i = 0; i < 64; ++i) {
            long v = ((long)base_q[t][i] * scale + 50) / 100;
            if (v <= 0)
                v = 1;
            if (v > 255)
                v = 255;
            qt[t][i] = (uint16_t)v;
        }
    divisors dv[2];
    make_divisors(qt[0], &dv[0]);
    make_divisors(qt[1], &dv[1]);
    enc_table tabs[4];
    for (int i = 0; i < 4; ++i)
        make_table(i, &tabs[i]);

    /* geometry */
    const int ch[3] = {hs, 1, 1}, cv[3] = {vs, 1, 1};
    const int mcus_x = ncomp == 1 ? (width + 7) / 8 : (width + 8 * hs - 1) / (8 * hs);
    const int mcus_y = ncomp == 1 ? (height + 7) / 8 : (height + 8 * vs - 1) / (8 * vs);
    int wib[3], hib[3], pw[3], ph[3];
    uint8_t* plane[3] = {0, 0, 0};
    for (int c = 0; c < ncomp; ++c) {
        wib[c] = (width * ch[c] + hs * 8 - 1) / (hs * 8);  /* real blocks per row (width_in_blocks) */
        hib[c] = (height * cv[c] + vs * 8 - 1) / (vs * 8); /* real block rows (height_in_blocks) */
        pw[c] = wib[c] * 8;
        ph[c] = mcus_y * cv[c] * 8;
        plane[c] = (uint8_t*)malloc((size_t)pw[c] * ph[c]);
        if (!plane[c])
            return -1;
    }
    /* colour conversion + downsampling with the reference's edge rules */
    for (int c = 0; c < ncomp; ++c) {
        const int fx = hs / ch[c], fy = vs / cv[c]; /* downsampling factors of this component */
        const int rows_in = (height + vs - 1) / vs * vs; /* input rows after padding to a full row group */
        const int rows_ds = rows_in / fy;
        for (int y = 0; y < ph[c]; ++y) {
            const int yd = y < rows_ds ? y : rows_ds - 1;
            for (int x = 0; x < pw[c]; ++x) {
                int sum = 0;
                for (int dy = 0; dy < fy; ++dy)
                    for (int dx = 0; dx < fx; ++dx) {
                        int sy = yd * fy + dy, sx = x * fx + dx;
                        if (sy > height - 1)
                            sy = height - 1;
                        if (sx > width - 1)
                            sx = width - 1;
                        const uint8_t* p = pixels + (size_t)sy * stride + (size_t)sx * channels;
                        int v;
                        if (ncomp == 1)
                            v = p[0];
                        else if (c == 0)
                            v = (19595 * p[0] + 38470 * p[1] + 7471 * p[2] + 32768) >> 16;
                        else if (c == 1)
                            v = (-11059 * p[0] - 21709 * p[1] + 32768 * p[2] + (128 << 16) + 32767) >> 16;
                        else
                            v = (32768 * p[0] - 27439 * p[1] - 5329 * p[2] + (128 << 16) + 32767) >> 16;
                        sum += v;
                    }
                int v = sum;
                if (fx == 2 && fy == 2)
                    v = (sum + ((x & 1) ? 2 : 1)) >> 2;
                else if (fx == 2 || fy == 2) {
                    /* h2v1: bias 0,1,0,1 along the row; v-only subsampling (1x2 luma) is an integral box with rounding */
                    v = fx == 2 ? (sum + (x & 1)) >> 1 : (sum + 1) >> 1;
                }
                plane[c][(size_t)y * pw[c] + x] = (uint8_t)v;
            }
        }
    }

    bitw w = {out, cap, 0, 0, 0, 0};
    /* ---- headers ---- */
    put_u16(&w, 0xFFD8);
    put_u16(&w, 0xFFE0);
    put_u16(&w, 16);
    put_byte(&w, 'J'); put_byte(&w, 'F'); put_byte(&w, 'I'); put_byte(&w, 'F'); put_byte(&w, 0);
    put_byte(&w, 1); put_byte(&w, 1); put_byte(&w, 0);
    put_u16(&w, 1); put_u16(&w, 1); put_byte(&w, 0); put_byte(&w, 0);
    for (int t = 0; t < (ncomp == 3 ? 2 : 1); ++t) {
        put_u16(&w, 0xFFDB);
        put_u16(&w, 67);
        put_byte(&w, t);
        for (int i = 0; i < 64; ++i)
            put_byte(&w, qt[t][zz_nat[i]]);
    }
    put_u16(&w, 0xFFC0);
    put_u16(&w, 8 + 3 * ncomp);
    put_byte(&w, 8);
    put_u16(&w, height);
    put_u16(&w, width);
    put_byte(&w, ncomp);
    for (int c = 0; c < ncomp; ++c) {
        put_byte(&w, c + 1);
        put_byte(&w, ch[c] * 16 + cv[c]);
        put_byte(&w, c ? 1 : 0);
    }
    for (int t = 0; t < (ncomp == 3 ? 2 : 1); ++t)
        for (int k = 0; k < 2; ++k) { /* DC then AC of table set t */
            const enc_table* e = &tabs[t + 2 * k];
            put_u16(&w, 0xFFC4);
            put_u16(&w, 2 + 1 + 16 + e->nvals);
            put_byte(&w, t + 16 * k);
            for (int i = 1; i <= 16; ++i)
                put_byte(&w, e->bits[i]);
            for (int i = 0; i < e->nvals; ++i)
                put_byte(&w, e->vals[i]);
        }
    if (restart_interval) {
        put_u16(&w, 0xFFDD);
        put_u16(&w, 4);
        put_u16(&w, restart_interval);
    }
    put_u16(&w, 0xFFDA);
    put_u16(&w, 6 + 2 * ncomp);
    put_byte(&w, ncomp);
    for (int c = 0; c < ncomp; ++c) {
        put_byte(&w, c + 1);
        put_byte(&w, c ? 0x11 : 0x00);
    }
    put_byte(&w, 0); put_byte(&w, 63); put_byte(&w, 0);

    /* ---- scan ---- */
    int last_dc[3] = {0, 0, 0};
    int to_go = restart_interval, next_rst = 0;
    int16_t mcu[6][64];
    for (int my = 0; my < mcus_y; ++my)
        for (int mx = 0; mx < mcus_x; ++mx) {
            if (restart_interval && to_go == 0) {
                flush_ones(&w);
                put_byte(&w, 0xFF);
                put_byte(&w, 0xD0 + next_rst);
                next_rst = (next_rst + 1) & 7;
                last_dc[0] = last_dc[1] = last_dc[2] = 0;
                to_go = restart_interval;
            }
            int b = 0;
            for (int c = 0; c < ncomp; ++c) {
                const int mw = ncomp == 1 ? 1 : ch[c], mh = ncomp == 1 ? 1 : cv[c];
                const enc_table *dct = &tabs[c ? 1 : 0], *act = &tabs[c ? 3 : 2];
                for (int by = 0; by < mh; ++by)
                    for (int bx = 0; bx < mw; ++bx, ++b) {
                        const int gx = mx * mw + bx, gy = my * mh + by;
                        if (gy < hib[c] && gx < wib[c]) {
                            fdct_quant(plane[c] + (size_t)gy * 8 * pw[c] + gx * 8, pw[c], &dv[c ? 1 : 0], mcu[b]);
                        } else { /* dummy block: AC zero, DC of the block before it in the MCU */
                            memset(mcu[b], 0, sizeof(mcu[b]));
                            mcu[b][0] = mcu[b - 1][0];
                            if (gy >= hib[c] && bx > 0)
                                mcu[b][0] = mcu[b - bx - 1][0]; /* whole dummy row copies the block before the row */
                        }
                        encode_block(&w, mcu[b], last_dc[c], dct, act);
                        last_dc[c] = mcu[b][0];
                    }
            }
            if (restart_interval)
                --to_go;
        }
    flush_ones(&w);
    put_u16(&w, 0xFFD9);
    for (int c = 0; c < ncomp; ++c)
        free(plane[c]);
    return w.overflow ? -1 : (long)w.n;
}
