/* oracle/jpeg_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT.  (see jpeg_oracle.h for the parity status)
 *
 * Plain-C restatement of the reference's baseline JPEG path.  "ljt/" = /root/reference/external/libjpeg-turbo
 * (3.0.3), the arithmetic behind extensions/libjpeg_turbo/jpeg_mem.cpp:102-460.
 * Written for clarity, not speed: whole-image coefficient arrays, then whole planes, then pixels. */
#include "jpeg_oracle.h"
#include <stdlib.h>
#include <string.h>

/* ljt/jutils.c:59 jpeg_natural_order (zigzag index -> natural index), +16 guard entries */
static const uint8_t ZZ[64 + 16] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13,
    6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47,
    55, 62, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};

typedef struct
{
    int present;
    uint8_t bits[17];
    uint8_t vals[256];
    /* derived, ljt/jdhuff.c:143-285 jpeg_make_d_derived_tbl */
    int32_t maxcode[18];
    int32_t valoffset[18];
} jo_htab;

typedef struct
{
    int id, hs, vs, tq, td, ta;
    int wib, hib;       /* MCU-padded block grid */
    int ds_w, ds_h;     /* downsampled_width / height: ljt/jdinput.c:130-135 */
    int16_t* coef;      /* hib*wib*64 natural order */
    uint8_t* plane;     /* (hib*8) x (wib*8) */
} jo_comp;

typedef struct
{
    int width, height, ncomp, sof, hmax, vmax;
    int mcus_x, mcus_y, restart_interval;
    int adobe, adobe_transform, jfif;
    uint16_t quant[4][64]; /* natural order */
    int quant_present[4];
    jo_htab dc[4], ac[4];
    jo_comp comp[3];
    const uint8_t* scan;
    size_t scan_len;
} jo_image;

static void jo_free(jo_image* im)
{
    for (int c = 0; c < 3; ++c) {
        free(im->comp[c].coef);
        free(im->comp[c].plane);
    }
}

static void derive(jo_htab* t)
{
    int code = 0, p = 0;
    for (int l = 1; l <= 16; ++l) {
        if (t->bits[l]) {
            t->valoffset[l] = p - code;
            p += t->bits[l];
            code += t->bits[l];
            t->maxcode[l] = code - 1;
        } else
            t->maxcode[l] = -1;
        code <<= 1;
    }
    t->maxcode[17] = 0xFFFFF; /* ljt/jdhuff.c:236 sentinel */
}

/* Marker parse: ljt/jdmarker.c get_sof:242 get_sos:308 get_dht:437 get_dqt:511 get_dri:569;
 * geometry: ljt/jdinput.c initial_setup:47, per_scan_setup:156 */
static int jo_parse(const uint8_t* d, size_t n, jo_image* im)
{
    memset(im, 0, sizeof(*im));
    if (n < 4 || d[0] != 0xFF || d[1] != 0xD8)
        return -1;
    size_t p = 2;
    int got_sof = 0;
    while (p + 4 <= n) {
        if (d[p] != 0xFF)
            return -2;
        while (p < n && d[p] == 0xFF)
            ++p;
        int m = d[p++];
        if (m == 0xD8 || (m >= 0xD0 && m <= 0xD7) || m == 0x01)
            continue;
        if (m == 0xD9)
            return -3;
        size_t L = (d[p] << 8) | d[p + 1];
        if (L < 2 || p + L > n)
            return -4;
        const uint8_t* s = d + p + 2;
        size_t sl = L - 2;
        if (m == 0xDB) {
            while (sl > 0) {
                int pq = s[0] >> 4, tq = s[0] & 15;
                if (tq > 3)
                    return -5;
                for (int i = 0; i < 64; ++i)
                    im->quant[tq][ZZ[i]] = pq ? (uint16_t)((s[1 + 2 * i] << 8) | s[2 + 2 * i]) : s[1 + i];
                im->quant_present[tq] = 1;
                size_t used = 1 + (pq ? 128 : 64);
                s += used;
                sl -= used;
            }
        } else if (m == 0xC4) {
            while (sl > 0) {
                int tc = s[0] >> 4, th = s[0] & 15;
                if (tc > 1 || th > 3)
                    return -6;
                jo_htab* t = tc ? &im->ac[th] : &im->dc[th];
                int cnt = 0;
                t->bits[0] = 0;
                for (int i = 1; i <= 16; ++i)
                    cnt += (t->bits[i] = s[i]);
                if (cnt > 256 || (size_t)(17 + cnt) > sl)
                    return -7;
                memset(t->vals, 0, 256);
                memcpy(t->vals, s + 17, cnt);
                t->present = 1;
                derive(t);
                s += 17 + cnt;
                sl -= 17 + cnt;
            }
        } else if (m == 0xC0 || m == 0xC1) {
            im->sof = m;
            if (s[0] != 8)
                return -8;
            im->height = (s[1] << 8) | s[2];
            im->width = (s[3] << 8) | s[4];
            im->ncomp = s[5];
            if (im->ncomp != 1 && im->ncomp != 3)
                return -9;
            for (int c = 0; c < im->ncomp; ++c) {
                im->comp[c].id = s[6 + 3 * c];
                im->comp[c].hs = s[7 + 3 * c] >> 4;
                im->comp[c].vs = s[7 + 3 * c] & 15;
                im->comp[c].tq = s[8 + 3 * c];
            }
            got_sof = 1;
        } else if (m >= 0xC2 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
            return -10; /* progressive / lossless / arithmetic: not this path */
        } else if (m == 0xDD) {
            im->restart_interval = (s[0] << 8) | s[1];
        } else if (m == 0xE0) {
            if (sl >= 5 && !memcmp(s, "JFIF", 5))
                im->jfif = 1;
        } else if (m == 0xEE) {
            if (sl >= 12 && !memcmp(s, "Adobe", 5)) {
                im->adobe = 1;
                im->adobe_transform = s[11];
            }
        } else if (m == 0xDA) {
            if (!got_sof)
                return -11;
            int ns = s[0];
            if (ns != im->ncomp)
                return -12; /* non-interleaved multi-scan baseline: not this path */
            for (int i = 0; i < ns; ++i) {
                int cid = s[1 + 2 * i], found = -1;
                for (int c = 0; c < im->ncomp; ++c)
                    if (im->comp[c].id == cid)
                        found = c;
                if (found != i)
                    return -13;
                im->comp[i].td = s[2 + 2 * i] >> 4;
                im->comp[i].ta = s[2 + 2 * i] & 15;
            }
            im->scan = d + p + L;
            im->scan_len = n - (p + L);
            break;
        }
        p += L;
    }
    if (!im->scan)
        return -14;
    im->hmax = im->vmax = 1;
    if (im->ncomp == 1)
        im->comp[0].hs = im->comp[0].vs = 1; /* single-component scan is non-interleaved: ljt/jdinput.c:166-186 */
    for (int c = 0; c < im->ncomp; ++c) {
        if (im->comp[c].hs > im->hmax)
            im->hmax = im->comp[c].hs;
        if (im->comp[c].vs > im->vmax)
            im->vmax = im->comp[c].vs;
    }
    im->mcus_x = (im->width + 8 * im->hmax - 1) / (8 * im->hmax);
    im->mcus_y = (im->height + 8 * im->vmax - 1) / (8 * im->vmax);
    for (int c = 0; c < im->ncomp; ++c) {
        jo_comp* k = &im->comp[c];
        k->wib = im->mcus_x * k->hs;
        k->hib = im->mcus_y * k->vs;
        k->ds_w = (im->width * k->hs + im->hmax - 1) / im->hmax;
        k->ds_h = (im->height * k->vs + im->vmax - 1) / im->vmax;
        k->coef = (int16_t*)calloc((size_t)k->wib * k->hib * 64, sizeof(int16_t));
        k->plane = (uint8_t*)calloc((size_t)k->wib * k->hib * 64, 1);
        if (!im->quant_present[k->tq & 3] || !im->dc[k->td & 3].present || !im->ac[k->ta & 3].present)
            return -15;
    }
    return 0;
}

/* ---------------- entropy decode: ljt/jdhuff.c decode_mcu_slow:552, jpeg_fill_bit_buffer:287, process_restart:512 --------- */
typedef struct
{
    const uint8_t* d;
    size_t n, pos;
    uint64_t buf;
    int nbits;
    int hit_marker; /* 0, or marker code seen (data exhausted => zeros are fed: ljt/jdhuff.c:365-390) */
} jo_bits;

static void fill(jo_bits* b)
{
    while (b->nbits <= 56) {
        int c = 0;
        if (!b->hit_marker) {
            if (b->pos >= b->n) {
                b->hit_marker = 0xD9;
            } else {
                c = b->d[b->pos++];
                if (c == 0xFF) {
                    int c2;
                    do {
                        c2 = b->pos < b->n ? b->d[b->pos++] : 0xD9;
                    } while (c2 == 0xFF); /* fill bytes */
                    if (c2 == 0)
                        c = 0xFF; /* stuffed */
                    else {
                        b->hit_marker = c2;
                        c = 0;
                    }
                }
            }
        }
        b->buf |= (uint64_t)c << (56 - b->nbits);
        b->nbits += 8;
    }
}
static inline int peek(jo_bits* b, int n) { return (int)(b->buf >> (64 - n)); }
static inline void skip(jo_bits* b, int n)
{
    b->buf <<= n;
    b->nbits -= n;
}
static int getbits(jo_bits* b, int n)
{
    if (!n)
        return 0;
    fill(b);
    int v = peek(b, n);
    skip(b, n);
    return v;
}
/* canonical decode, equivalent to HUFF_DECODE / jpeg_huff_decode (ljt/jdhuff.c:398-450) */
static int decode_sym(jo_bits* b, const jo_htab* t)
{
    fill(b);
    int code = 0;
    for (int l = 1; l <= 16; ++l) {
        code = peek(b, l);
        if (code <= t->maxcode[l] && t->maxcode[l] >= 0) {
            skip(b, l);
            return t->vals[(code + t->valoffset[l]) & 255];
        }
    }
    skip(b, 16); /* ljt/jdhuff.c:443-447: bad code => warning, return 0 */
    return 0;
}
/* ljt/jdhuff.c:483-488 HUFF_EXTEND */
static inline int extend(int x, int s) { return x < (1 << (s - 1)) ? x - (1 << s) + 1 : x; }

static int jo_entropy_decode(jo_image* im)
{
    jo_bits b = {im->scan, im->scan_len, 0, 0, 0, 0};
    int pred[3] = {0, 0, 0};
    int total = im->mcus_x * im->mcus_y;
    int next_rst = 0;
    for (int mcu = 0; mcu < total; ++mcu) {
        if (im->restart_interval && mcu && mcu % im->restart_interval == 0) {
            /* process_restart: discard partial byte, expect RSTn, reset predictors */
            b.buf = 0;
            b.nbits = 0;
            if (!b.hit_marker) {
                /* advance to next marker */
                while (b.pos + 1 < b.n && !(b.d[b.pos] == 0xFF && b.d[b.pos + 1] != 0 && b.d[b.pos + 1] != 0xFF))
                    b.pos++;
                if (b.pos + 1 < b.n) {
                    b.hit_marker = b.d[b.pos + 1];
                    b.pos += 2;
                }
            }
            if (b.hit_marker != 0xD0 + next_rst)
                return -20; /* the reference would try to resync (jdmarker.c:1188); we call it corrupt */
            b.hit_marker = 0;
            next_rst = (next_rst + 1) & 7;
            pred[0] = pred[1] = pred[2] = 0;
        }
        int mx = mcu % im->mcus_x, my = mcu / im->mcus_x;
        for (int c = 0; c < im->ncomp; ++c) {
            jo_comp* k = &im->comp[c];
            for (int v = 0; v < k->vs; ++v)
                for (int h = 0; h < k->hs; ++h) {
                    int16_t* blk = k->coef + ((size_t)(my * k->vs + v) * k->wib + (mx * k->hs + h)) * 64;
                    int s = decode_sym(&b, &im->dc[k->td]);
                    if (s) {
                        int r = getbits(&b, s);
                        s = extend(r, s);
                    }
                    pred[c] += s;
                    blk[0] = (int16_t)pred[c];
                    for (int kk = 1; kk < 64; ++kk) {
                        int rs = decode_sym(&b, &im->ac[k->ta]);
                        int r = rs >> 4;
                        s = rs & 15;
                        if (s) {
                            kk += r;
                            r = getbits(&b, s);
                            blk[ZZ[kk]] = (int16_t)extend(r, s);
                        } else {
                            if (r != 15)
                                break;
                            kk += 15;
                        }
                    }
                }
        }
    }
    return 0;
}

/* ---------------- dequant + IDCT: ljt/jidctint.c:173-410 _jpeg_idct_islow ---------------- */
#define CONST_BITS 13
#define PASS1_BITS 2
#define F_0_298631336 2446
#define F_0_390180644 3196
#define F_0_541196100 4433
#define F_0_765366865 6270
#define F_0_899976223 7373
#define F_1_175875602 9633
#define F_1_501321110 12299
#define F_1_847759065 15137
#define F_1_961570560 16069
#define F_2_053119869 16819
#define F_2_562915447 20995
#define F_3_072711026 25172
#define DESCALE(x, n) (((x) + ((int64_t)1 << ((n)-1))) >> (n)) /* ljt/jdct.h:191, arithmetic shift */

/* post-IDCT range limit incl. +128 level shift and the &1023 wrap: ljt/jdmaster.c:386-407,474-495 */
static inline uint8_t range_limit(int64_t v)
{
    int i = (int)(v & 1023);
    return (uint8_t)(i < 128 ? i + 128 : i < 512 ? 255 : i < 896 ? 0 : i - 896);
}

static void idct_1d(const int64_t in[8], int64_t out[8])
{
    /* returns the 8 un-descaled sums (tmp10+tmp3 ... tmp13-tmp0) in output-index order */
    int64_t z2 = in[2], z3 = in[6];
    int64_t z1 = (z2 + z3) * F_0_541196100;
    int64_t tmp2 = z1 + z3 * -F_1_847759065;
    int64_t tmp3 = z1 + z2 * F_0_765366865;
    int64_t tmp0 = (in[0] + in[4]) * ((int64_t)1 << CONST_BITS);
    int64_t tmp1 = (in[0] - in[4]) * ((int64_t)1 << CONST_BITS);
    int64_t tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    tmp0 = in[7];
    tmp1 = in[5];
    tmp2 = in[3];
    tmp3 = in[1];
    z1 = tmp0 + tmp3;
    z2 = tmp1 + tmp2;
    z3 = tmp0 + tmp2;
    int64_t z4 = tmp1 + tmp3;
    int64_t z5 = (z3 + z4) * F_1_175875602;
    tmp0 *= F_0_298631336;
    tmp1 *= F_2_053119869;
    tmp2 *= F_3_072711026;
    tmp3 *= F_1_501321110;
    z1 *= -F_0_899976223;
    z2 *= -F_2_562915447;
    z3 *= -F_1_961570560;
    z4 *= -F_0_390180644;
    z3 += z5;
    z4 += z5;
    tmp0 += z1 + z3;
    tmp1 += z2 + z4;
    tmp2 += z2 + z3;
    tmp3 += z1 + z4;
    out[0] = tmp10 + tmp3;
    out[7] = tmp10 - tmp3;
    out[1] = tmp11 + tmp2;
    out[6] = tmp11 - tmp2;
    out[2] = tmp12 + tmp1;
    out[5] = tmp12 - tmp1;
    out[3] = tmp13 + tmp0;
    out[4] = tmp13 - tmp0;
}

void jo_idct_islow(const int16_t* coef, const uint16_t* q, uint8_t* out, int stride)
{
    int ws[64]; /* `int workspace[DCTSIZE2]`: values are truncated to int between the passes like the reference */
    for (int c = 0; c < 8; ++c) {
        int64_t in[8], o[8];
        for (int r = 0; r < 8; ++r)
            in[r] = (int64_t)((int)coef[r * 8 + c] * (int)q[r * 8 + c]); /* DEQUANTIZE, MULTIPLIER=int w/o SIMD */
        idct_1d(in, o);
        for (int r = 0; r < 8; ++r)
            ws[r * 8 + c] = (int)DESCALE(o[r], CONST_BITS - PASS1_BITS);
        /* the all-AC-zero column shortcut (jidctint.c:205-227) gives dc<<PASS1_BITS == the general result */
    }
    for (int r = 0; r < 8; ++r) {
        int64_t in[8], o[8];
        for (int c = 0; c < 8; ++c)
            in[c] = ws[r * 8 + c];
        idct_1d(in, o);
        for (int c = 0; c < 8; ++c)
            out[r * stride + c] = range_limit(DESCALE(o[c], CONST_BITS + PASS1_BITS + 3));
        /* zero-AC row shortcut (jidctint.c:302-321): DESCALE(ws0, PASS1_BITS+3) == general path since
         * (ws0<<13 + 2^17)>>18 == (ws0 + 2^4)>>5 */
    }
}

/* ljt/jdcoefct.c:86-166 decompress_onepass: every block of the padded grid gets an IDCT here; the reference skips the
 * dummy right/bottom blocks, whose samples are never read for real output pixels */
static void jo_idct_all(jo_image* im)
{
    for (int c = 0; c < im->ncomp; ++c) {
        jo_comp* k = &im->comp[c];
        int pitch = k->wib * 8;
        for (int by = 0; by < k->hib; ++by)
            for (int bx = 0; bx < k->wib; ++bx)
                jo_idct_islow(k->coef + ((size_t)by * k->wib + bx) * 64, im->quant[k->tq], k->plane + (size_t)by * 8 * pitch + bx * 8,
                    pitch);
    }
}

/* ---------------- upsampling: ljt/jdsample.c (fancy h2v1:276 h1v2:316 h2v2:362, box h2v1:205 h2v2:233 int:161),
 * edge rows: ljt/jdmainct.c:217-250 (first/last *real* row is duplicated), method choice: jdsample.c:414-535 ---------- */
static int upsampled(const jo_image* im, int c, int x, int y, int fancy)
{
    const jo_comp* k = &im->comp[c];
    int pitch = k->wib * 8;
    const uint8_t* P = k->plane;
    int hx = im->hmax / k->hs, vy = im->vmax / k->vs; /* expansion factors 1 or 2 (4:1:1: 4 x 1) */
    if (hx == 1 && vy == 1)
        return P[(size_t)y * pitch + x];
    if (hx == 4 && vy == 1) /* neither a 2:1 case nor merged: int_upsample (jdsample.c:161-203, chosen at :501-520) replicates */
        return P[(size_t)y * pitch + (x >> 2)];
    if (hx == 2 && vy == 1) {
        const uint8_t* in = P + (size_t)y * pitch;
        int cx = x >> 1;
        if (!(fancy && k->ds_w > 2))
            return in[cx];
        if (!(x & 1))
            return cx == 0 ? in[0] : (in[cx] * 3 + in[cx - 1] + 1) >> 2;
        return cx == k->ds_w - 1 ? in[cx] : (in[cx] * 3 + in[cx + 1] + 2) >> 2;
    }
    int r = y >> 1;
    int far = (y & 1) ? (r + 1 > k->ds_h - 1 ? k->ds_h - 1 : r + 1) : (r - 1 < 0 ? 0 : r - 1);
    const uint8_t* n0 = P + (size_t)r * pitch;
    const uint8_t* n1 = P + (size_t)far * pitch;
    if (hx == 1 && vy == 2) {
        if (!fancy)
            return n0[x];
        return (n0[x] * 3 + n1[x] + ((y & 1) ? 2 : 1)) >> 2;
    }
    /* h2v2 */
    int cx = x >> 1;
    if (!(fancy && k->ds_w > 2))
        return n0[cx];
    int cs = n0[cx] * 3 + n1[cx];
    if (!(x & 1)) {
        if (cx == 0)
            return (cs * 4 + 8) >> 4;
        return (cs * 3 + (n0[cx - 1] * 3 + n1[cx - 1]) + 8) >> 4;
    }
    if (cx == k->ds_w - 1)
        return (cs * 4 + 7) >> 4;
    return (cs * 3 + (n0[cx + 1] * 3 + n1[cx + 1]) + 7) >> 4;
}

/* ---------------- colour: ljt/jdcolor.c:215-251 tables, ljt/jdcolext.c:30-76 inner loop ---------------- */
static inline uint8_t clamp255(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }
static void ycc_to_rgb(int y, int cb, int cr, uint8_t* r, uint8_t* g, uint8_t* b)
{
    int x_cb = cb - 128, x_cr = cr - 128;
    int cr_r = (int)((91881LL * x_cr + 32768) >> 16);
    int cb_b = (int)((116130LL * x_cb + 32768) >> 16);
    int g_off = (int)((-22554LL * x_cb + 32768 + -46802LL * x_cr) >> 16);
    *r = clamp255(y + cr_r);
    *g = clamp255(y + g_off);
    *b = clamp255(y + cb_b);
}

static int jo_run(const uint8_t* jpg, size_t len, jo_image* im, int stage)
{
    int rc = jo_parse(jpg, len, im);
    if (rc == 0 && stage >= 1)
        rc = jo_entropy_decode(im);
    if (rc == 0 && stage >= 2)
        jo_idct_all(im);
    return rc;
}

int jo_info(const uint8_t* jpg, size_t len, int* out)
{
    jo_image im;
    int rc = jo_run(jpg, len, &im, 0);
    if (rc == 0) {
        out[0] = im.width;
        out[1] = im.height;
        out[2] = im.ncomp;
        out[3] = im.comp[0].hs;
        out[4] = im.comp[0].vs;
        out[5] = im.restart_interval;
        out[6] = im.sof;
        out[7] = (int)im.scan_len;
    }
    jo_free(&im);
    return rc;
}

int jo_read_coefficients(const uint8_t* jpg, size_t len, int comp, int16_t* out, size_t cap_blocks, int* dims)
{
    jo_image im;
    int rc = jo_run(jpg, len, &im, 1);
    if (rc == 0) {
        if (comp >= im.ncomp)
            rc = -30;
        else {
            dims[0] = im.comp[comp].wib;
            dims[1] = im.comp[comp].hib;
            if ((size_t)dims[0] * dims[1] > cap_blocks)
                rc = -31;
            else
                memcpy(out, im.comp[comp].coef, (size_t)dims[0] * dims[1] * 128);
        }
    }
    jo_free(&im);
    return rc;
}

int jo_read_plane(const uint8_t* jpg, size_t len, int comp, uint8_t* out, size_t cap, int* dims)
{
    jo_image im;
    int rc = jo_run(jpg, len, &im, 2);
    if (rc == 0) {
        if (comp >= im.ncomp)
            rc = -30;
        else {
            dims[0] = im.comp[comp].wib * 8;
            dims[1] = im.comp[comp].hib * 8;
            if ((size_t)dims[0] * dims[1] > cap)
                rc = -31;
            else
                memcpy(out, im.comp[comp].plane, (size_t)dims[0] * dims[1]);
        }
    }
    jo_free(&im);
    return rc;
}

/* Output mapping: extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:335-353 (format -> JCS_*), :409-436 (planar = decode
 * interleaved then de-interleave), jpeg_mem.cpp:154-172 */
int jo_decode(const uint8_t* jpg, size_t len, int fmt, int fancy, uint8_t* out, int64_t stride)
{
    jo_image im;
    int rc = jo_run(jpg, len, &im, 2);
    if (rc) {
        jo_free(&im);
        return rc;
    }
    int W = im.width, H = im.height;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            uint8_t r, g, b;
            int Y = upsampled(&im, 0, x, y, fancy);
            if (fmt == JO_FMT_Y) {
                out[(size_t)y * stride + x] = (uint8_t)Y; /* grayscale_convert: ljt/jdcolor.c:438 */
                continue;
            }
            if (im.ncomp == 1)
                r = g = b = (uint8_t)Y; /* gray_rgb_convert: ljt/jdcolor.c:451 */
            else
                ycc_to_rgb(Y, upsampled(&im, 1, x, y, fancy), upsampled(&im, 2, x, y, fancy), &r, &g, &b);
            uint8_t c0 = (fmt == JO_FMT_I_BGR || fmt == JO_FMT_P_BGR) ? b : r;
            uint8_t c2 = (fmt == JO_FMT_I_BGR || fmt == JO_FMT_P_BGR) ? r : b;
            if (fmt == JO_FMT_I_RGB || fmt == JO_FMT_I_BGR) {
                uint8_t* p = out + (size_t)y * stride + 3 * x;
                p[0] = c0;
                p[1] = g;
                p[2] = c2;
            } else {
                size_t plane = (size_t)stride * H;
                out[(size_t)y * stride + x] = c0;
                out[plane + (size_t)y * stride + x] = g;
                out[2 * plane + (size_t)y * stride + x] = c2;
            }
        }
    jo_free(&im);
    return 0;
}
