// enc_math.cuh -- per-thread arithmetic of the b200jpeg encode path as host+device inline functions (the CUDA kernels in
// encode_kernels.cu wrap them with the memory staging; compiled with g++ the same bodies run in the CPU emulation of the
// "not gpu" tests, tests/emu).
//
// Reference arithmetic restated here ("ljt/" = external/libjpeg-turbo in the reference tree):
//   RGB -> YCbCr        ljt/jccolor.c rgb_ycc_start:214-246 (16.16 fixed point), ljt/jccolext.c rgb_ycc_convert_internal:30-76
//   chroma downsampling ljt/jcsample.c h2v1_downsample:217-252 (bias 0,1), h2v2_downsample:264-295 (bias 1,2),
//                       int_downsample:141-178 (1x2: plain rounding); edges expand_right_edge:99-116,
//                       ljt/jcprepct.c expand_bottom_edge:104-118 / pre_process_data:126-195
//   forward DCT         ljt/jfdctint.c _jpeg_fdct_islow:143-288 (LL&M, CONST_BITS 13, PASS1_BITS 2), level shift
//                       ljt/jcdctmgr.c convsamp:384-412
//   quantisation        ljt/jcdctmgr.c quantize:419-452 with the reciprocals of compute_reciprocal:173-228
//   entropy coding      ljt/jchuff.c encode_one_block:558-655
#pragma once
#include "enc_types.h"
#if !defined(__CUDACC__)
#include "host_shim.h"
#endif

namespace b200jpeg {

// the same three numbers per coefficient in one 8-byte word (one shared-memory load in the kernel): x = recip, y = corr | shift << 16
struct EncQuantPacked
{
    const uint2* q;
    B2J_HD uint32_t get_recip(int n) const { return q[n].x; }
    B2J_HD uint32_t get_corr(int n) const { return q[n].y & 0xFFFFu; }
    B2J_HD uint32_t get_shift(int n) const { return q[n].y >> 16; }
};


B2J_HD constexpr int enc_zigzag(int k) // natural-order index of zigzag position k (ljt/jutils.c jpeg_natural_order)
{
    constexpr uint8_t t[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21,
        28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
    return t[k];
}

// ---- colour ----
B2J_HD uint32_t enc_luma(uint32_t r, uint32_t g, uint32_t b) { return (19595u * r + 38470u * g + 7471u * b + 32768u) >> 16; }
B2J_HD uint32_t enc_cb(int r, int g, int b) { return uint32_t((-11059 * r - 21709 * g + 32768 * b + (128 << 16) + 32767) >> 16); }
B2J_HD uint32_t enc_cr(int r, int g, int b) { return uint32_t((32768 * r - 27439 * g - 5329 * b + (128 << 16) + 32767) >> 16); }

// ---- forward DCT ----
// One 1-D pass over v[0..7] in place.  PASS2 = false: row pass, outputs keep 2 extra fractional bits;
// PASS2 = true: column pass, those bits are removed (results stay scaled by 8, which the quantiser divides out).
template <bool PASS2>
B2J_HD void enc_fdct_1d(int& v0, int& v1, int& v2, int& v3, int& v4, int& v5, int& v6, int& v7)
{
    constexpr int SH = PASS2 ? 15 : 11, RND = 1 << (SH - 1);
    const int s07 = v0 + v7, d07 = v0 - v7, s16 = v1 + v6, d16 = v1 - v6;
    const int s25 = v2 + v5, d25 = v2 - v5, s34 = v3 + v4, d34 = v3 - v4;
    const int e0 = s07 + s34, e3 = s07 - s34, e1 = s16 + s25, e2 = s16 - s25;
    if (PASS2) {
        v0 = (e0 + e1 + 2) >> 2;
        v4 = (e0 - e1 + 2) >> 2;
    } else {
        v0 = (e0 + e1) * 4;
        v4 = (e0 - e1) * 4;
    }
    const int rot = (e2 + e3) * 4433;
    v2 = (rot + e3 * 6270 + RND) >> SH;
    v6 = (rot - e2 * 15137 + RND) >> SH;
    const int a = d34 + d07, b = d25 + d16, c = d34 + d16, d = d25 + d07;
    const int w = (c + d) * 9633;
    const int za = a * -7373, zb = b * -20995, zc = c * -16069 + w, zd = d * -3196 + w;
    v7 = (d34 * 2446 + za + zc + RND) >> SH;
    v5 = (d25 * 16819 + zb + zd + RND) >> SH;
    v3 = (d16 * 25172 + zb + zc + RND) >> SH;
    v1 = (d07 * 12299 + za + zd + RND) >> SH;
}

B2J_HD int enc_quantise(int v, uint32_t recip, uint32_t corr, uint32_t shift)
{
    const uint32_t mag = uint32_t(v < 0 ? -v : v) + corr;
#if defined(__CUDA_ARCH__)
    const uint32_t q = __umulhi(mag, recip) >> shift;
#else
    const uint32_t q = uint32_t((uint64_t(mag) * recip) >> 32) >> shift;
#endif
    return v < 0 ? -int(q) : int(q);
}

// Level shift, 2-D DCT and quantisation of one block.  rows[2 * r], rows[2 * r + 1]: the 8 samples of row r packed
// little-endian; zz[k]: quantised coefficient at zigzag position k.
template <class Quant>
B2J_HD void enc_fdct_quant(const uint32_t* rows, const Quant& qt, int16_t* zz)
{
    int ws[64];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const uint32_t lo = rows[2 * r], hi = rows[2 * r + 1];
        int v0 = int(lo & 255u) - 128, v1 = int((lo >> 8) & 255u) - 128, v2 = int((lo >> 16) & 255u) - 128, v3 = int(lo >> 24) - 128;
        int v4 = int(hi & 255u) - 128, v5 = int((hi >> 8) & 255u) - 128, v6 = int((hi >> 16) & 255u) - 128, v7 = int(hi >> 24) - 128;
        enc_fdct_1d<false>(v0, v1, v2, v3, v4, v5, v6, v7);
        ws[r * 8 + 0] = v0; ws[r * 8 + 1] = v1; ws[r * 8 + 2] = v2; ws[r * 8 + 3] = v3;
        ws[r * 8 + 4] = v4; ws[r * 8 + 5] = v5; ws[r * 8 + 6] = v6; ws[r * 8 + 7] = v7;
    }
#pragma unroll
    for (int c = 0; c < 8; ++c)
        enc_fdct_1d<true>(ws[c], ws[8 + c], ws[16 + c], ws[24 + c], ws[32 + c], ws[40 + c], ws[48 + c], ws[56 + c]);
#pragma unroll
    for (int k = 0; k < 64; ++k) {
        const int n = enc_zigzag(k);
        zz[k] = int16_t(enc_quantise(ws[n], qt.get_recip(n), qt.get_corr(n), qt.get_shift(n)));
    }
}

// ---- entropy coding ----
B2J_HD int enc_category(int v) // bits needed for |v| (JPEG_NBITS)
{
    const uint32_t m = uint32_t(v < 0 ? -v : v);
#if defined(__CUDA_ARCH__)
    return 32 - __clz(int(m));
#else
    return m ? 32 - __builtin_clz(m) : 0;
#endif
}

// DC part of a block: Huffman code of the category followed by the value bits of the difference.
template <class Sink>
B2J_HD void enc_put_dc(int diff, const uint32_t* dc_tab, Sink& sink)
{
    const int n = enc_category(diff);
    const uint32_t e = dc_tab[n];
    const uint32_t vbits = uint32_t(diff < 0 ? diff - 1 : diff) & ((1u << n) - 1u);
    sink.put(((e >> 8) << n) | vbits, int(e & 255u) + n);
}

// AC part: run/size symbols with value bits, ZRL for runs > 15, EOB when the block ends in zeros.
// get(k) returns the coefficient at zigzag position k.
template <class Sink, class Get>
B2J_HD void enc_put_ac(const Get& get, const uint32_t* ac_tab, Sink& sink)
{
    int run = 0;
#pragma unroll
    for (int k = 1; k < 64; ++k) {
        const int v = get(k);
        if (v == 0) {
            ++run;
            continue;
        }
        while (run > 15) {
            const uint32_t z = ac_tab[0xF0];
            sink.put(z >> 8, int(z & 255u));
            run -= 16;
        }
        const int n = enc_category(v);
        const uint32_t e = ac_tab[run * 16 + n];
        const uint32_t vbits = uint32_t(v < 0 ? v - 1 : v) & ((1u << n) - 1u);
        sink.put(((e >> 8) << n) | vbits, int(e & 255u) + n);
        run = 0;
    }
    if (run) {
        const uint32_t e = ac_tab[0];
        sink.put(e >> 8, int(e & 255u));
    }
}

// The same AC part driven by the block's non-zero mask (bit k = coefficient at zigzag position k is non-zero): one loop
// iteration per non-zero coefficient instead of a test of all 63 positions.
B2J_HD int enc_lowest_bit(uint64_t m) // index of the lowest set bit, m != 0
{
#if defined(__CUDA_ARCH__)
    return __ffsll((long long)m) - 1;
#else
    return __builtin_ctzll(m);
#endif
}

template <class Sink, class Get>
B2J_HD void enc_put_ac_sparse(uint64_t mask, const Get& get, const uint32_t* ac_tab, Sink& sink)
{
    int prev = 0;
    mask &= ~uint64_t(1);
    while (mask) {
        const int k = enc_lowest_bit(mask);
        mask &= mask - 1;
        int run = k - prev - 1;
        prev = k;
        for (; run > 15; run -= 16) {
            const uint32_t z = ac_tab[0xF0];
            sink.put(z >> 8, int(z & 255u));
        }
        const int v = get(k);
        const int n = enc_category(v);
        const uint32_t e = ac_tab[run * 16 + n];
        const uint32_t vbits = uint32_t(v < 0 ? v - 1 : v) & ((1u << n) - 1u);
        sink.put(((e >> 8) << n) | vbits, int(e & 255u) + n);
    }
    if (prev != 63) {
        const uint32_t e = ac_tab[0];
        sink.put(e >> 8, int(e & 255u));
    }
}

struct EncBitCounter
{
    uint32_t bits = 0;
    B2J_HD void put(uint32_t, int len) { bits += uint32_t(len); }
};

// Writes big-endian bit order into 32-bit words that other threads may share at both ends: every word goes out with an
// OR (atomic on the device).  The words hold the stream as big-endian integers; the copy-out byte-swaps them.
struct EncWordSink
{
    uint32_t* words;
    uint32_t widx;
    uint64_t acc; // bits left-aligned
    int n;        // valid bits in acc
    bool first;   // the next word written is the first one: its leading bits may belong to the previous block
    B2J_HD void start(uint32_t* w, uint32_t bitpos)
    {
        words = w;
        widx = bitpos >> 5;
        acc = 0;
        n = int(bitpos & 31u);
        first = true;
    }
    B2J_HD void set_word(uint32_t v) // a word in the middle of the block's bits: nobody else writes it
    {
        words[widx] = v;
        ++widx;
    }
    B2J_HD void or_word(uint32_t v)
    {
        if (v) {
#if defined(__CUDA_ARCH__)
            atomicOr(words + widx, v);
#else
            words[widx] |= v;
#endif
        }
        ++widx;
    }
    B2J_HD void put(uint32_t code, int len) // len <= 27
    {
        acc |= uint64_t(code) << (64 - n - len);
        n += len;
        if (n >= 32) {
            if (first)
                or_word(uint32_t(acc >> 32));
            else
                set_word(uint32_t(acc >> 32));
            first = false;
            acc <<= 32;
            n -= 32;
        }
    }
    B2J_HD void finish()
    {
        if (n)
            or_word(uint32_t(acc >> 32));
    }
};

} // namespace b200jpeg
