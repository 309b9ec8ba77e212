// jpeg_math.cuh -- the arithmetic of the b200jpeg decode path as host+device inline functions.
// Each function is the per-thread body of one kernel in decode_kernels.cu; compiled with g++ (B2J_HD = inline) the same
// bodies run in the CPU emulation used by the "not gpu" tests (tests/emu), so the logic is checked without a GPU.
//
// Reference arithmetic restated here ("ljt/" = external/libjpeg-turbo in the reference tree):
//   Huffman decode      ljt/jdhuff.c decode_mcu_slow:552-660, HUFF_EXTEND:483-488, jpeg_huff_decode:398-450
//   dequant + IDCT      ljt/jidctint.c:173-410 (_jpeg_idct_islow), range limit ljt/jdmaster.c:386-407
//   upsampling          ljt/jdsample.c h2v1_fancy:276 h1v2_fancy:316 h2v2_fancy:362 h2v1:205 h2v2:233,
//                       edge rows ljt/jdmainct.c:217-250
//   colour              ljt/jdcolor.c:215-251, ljt/jdcolext.c:30-76
#pragma once
#include "device_types.h"
#if !defined(__CUDACC__)
#include "host_shim.h"
#endif

namespace b200jpeg {

#if defined(__CUDA_ARCH__)
#define B2J_ZZ(i) c_zigzag[i]
static __constant__ uint8_t c_zigzag[80] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,
    7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55,
    62, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};
#else
#if defined(__CUDACC__)
static __constant__ uint8_t c_zigzag[80] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,
    7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55,
    62, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};
#endif
static const uint8_t h_zigzag[80] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,
    7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55,
    62, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};
#define B2J_ZZ(i) h_zigzag[i]
#endif

B2J_HD uint32_t bswap32(uint32_t v)
{
#if defined(__CUDA_ARCH__)
    return __byte_perm(v, 0, 0x0123);
#else
    return __builtin_bswap32(v);
#endif
}

// ------------------------------------------------------------------------------------------------------------------
// Bit window over the unstuffed stream: 32 valid bits starting at absolute bit position p (MSB first).
// ------------------------------------------------------------------------------------------------------------------
struct BitWindow
{
    const uint32_t* words;
    uint32_t idx;
    uint32_t w0, w1;
};

B2J_HD void bw_init(BitWindow& bw, const uint32_t* words)
{
    bw.words = words;
    bw.idx = 0xFFFFFFF0u;
    bw.w0 = bw.w1 = 0;
}

B2J_HD uint32_t bw_peek(BitWindow& bw, uint32_t p)
{
    const uint32_t i = p >> 5;
    if (i != bw.idx) {
        if (i == bw.idx + 1) {
            bw.w0 = bw.w1;
            bw.w1 = bswap32(bw.words[i + 1]);
        } else {
            bw.w0 = bswap32(bw.words[i]);
            bw.w1 = bswap32(bw.words[i + 1]);
        }
        bw.idx = i;
    }
    const uint32_t sh = p & 31;
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(bw.w1, bw.w0, sh);
#else
    return sh ? (bw.w0 << sh) | (bw.w1 >> (32 - sh)) : bw.w0;
#endif
}

// One Huffman symbol from the top bits of win. Returns symbol, sets len (1..16). ok=false on an invalid code.
B2J_HD int huff_symbol(const DevHuffTable* t, uint32_t win, int& len, bool& ok)
{
    const uint32_t e = t->lut[win >> (32 - kLutBits)];
    if (e) {
        len = int(e >> 8);
        return int(e & 255);
    }
    for (int l = kLutBits + 1; l <= 16; ++l) {
        const int code = int(win >> (32 - l));
        if (code <= t->maxcode[l]) {
            len = l;
            return t->vals[(code + t->valoffset[l]) & 255];
        }
    }
    ok = false;
    len = 16;
    return 0;
}

B2J_HD int huff_extend(int x, int s) { return x < (1 << (s - 1)) ? x - (1 << s) + 1 : x; }

struct SubseqState
{
    uint32_t p;    // bit position of the next symbol (relative to the image's unstuffed stream)
    uint32_t slot; // coefficient slot inside the current MCU: block_in_mcu * 64 + zigzag index
    uint32_t n;    // slots consumed: absolute (from image start) if abs, else relative to the entry of this decode
    bool abs;
};

// ------------------------------------------------------------------------------------------------------------------
// Decode every symbol that starts inside subsequence k of image `im`, beginning in state `s`.
// WRITE=false: synchronisation pass (nothing is stored).  WRITE=true: final pass, s.abs must be true; coefficients are
// written de-zigzagged into the image's coefficient area (DC still as the difference).
// Restart segments: segment g covers unstuffed bytes [S[g], S[g+1]); crossing into a new segment makes the state exact.
// ------------------------------------------------------------------------------------------------------------------
template <bool WRITE>
B2J_HD void decode_subsequence(const BatchDev& b, const ImageDesc& im, ImageState* st, uint32_t k, uint32_t clean_len, SubseqState& s)
{
    const uint32_t clean_bits = clean_len * 8u;
    uint32_t end = (k + 1) * uint32_t(b.subseq_bits);
    if (end > clean_bits)
        end = clean_bits;
    const uint32_t* S = b.seg_start + im.seg_first;
    const int nseg = im.num_segments;

    // locate the segment containing p
    int g = 0;
    if (nseg > 1) {
        int lo = 0, hi = nseg - 1;
        const uint32_t pb = s.p >> 3;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (S[mid] <= pb)
                lo = mid;
            else
                hi = mid - 1;
        }
        g = lo;
    }
    if (!s.abs && s.p == S[g] * 8u) { // exactly at a segment start: the state is known
        s.abs = true;
        s.n = uint32_t(g) * im.seg_slots;
        s.slot = 0;
    }
    bool last = (g == nseg - 1);
    uint32_t B = last ? clean_bits : S[g + 1] * 8u;
    uint32_t seg_end = last ? im.total_slots : uint32_t(g + 1) * im.seg_slots;

    BitWindow bw;
    bw_init(bw, reinterpret_cast<const uint32_t*>(b.clean + im.clean_off));
    int16_t* coef = WRITE ? b.coef + im.coef_off : nullptr;
    uint32_t err = 0;

    while (true) {
        bool advance = false;
        if (s.abs && s.n >= seg_end) { // segment complete: the rest of its bytes is padding
            if (last) {
                if (WRITE)
                    st->done_slots = s.n;
                break;
            }
            advance = true;
        } else if (s.p >= B) {
            if (last)
                break;
            if (WRITE)
                err |= kErrOverrun; // segment data ended before its blocks were complete
            advance = true;
        }
        if (!advance) {
            if (s.p >= end)
                break;
            const uint32_t win = bw_peek(bw, s.p);
            const uint32_t blk = s.slot >> 6;
            uint32_t z = s.slot & 63;
            const int comp = im.blk_comp[blk];
            const DevHuffTable* t = b.huff + (z == 0 ? im.dc_tab[comp] : im.ac_tab[comp]);
            int len;
            bool ok = true;
            const int sym = huff_symbol(t, win, len, ok);
            const int sz = sym & 15;
            const uint32_t total = uint32_t(len + sz);
            if (s.p + total > B) { // symbol would cross the segment end: it was padding, not data
                if (last) {
                    if (WRITE)
                        err |= kErrOverrun;
                    s.p = B;
                    break;
                }
                advance = true;
            } else {
                if (WRITE && !ok)
                    err |= kErrBadCode;
                int val = 0;
                if (sz)
                    val = huff_extend(int((win << len) >> (32 - sz)), sz);
                uint32_t adv;
                if (z == 0) {
                    if (WRITE)
                        coef[(s.n & ~63u)] = int16_t(val);
                    adv = 1;
                } else {
                    const int r = sym >> 4;
                    if (sz) {
                        z += r;
                        if (z > 63) {
                            z = 63;
                            if (WRITE)
                                err |= kErrBadCode;
                        }
                        if (WRITE)
                            coef[(s.n & ~63u) + B2J_ZZ(z)] = int16_t(val);
                        adv = z + 1 - (s.slot & 63);
                    } else if (r == 15) {
                        adv = 16;
                        if ((s.slot & 63) + 16 > 64)
                            adv = 64 - (s.slot & 63);
                    } else {
                        adv = 64 - (s.slot & 63); // EOB
                    }
                }
                s.slot += adv;
                s.n += adv;
                if (s.slot >= uint32_t(im.mcu_slots))
                    s.slot = 0;
                s.p += total;
            }
        }
        if (advance) {
            ++g;
            s.p = S[g] * 8u;
            s.slot = 0;
            s.abs = true;
            s.n = uint32_t(g) * im.seg_slots;
            last = (g == nseg - 1);
            B = last ? clean_bits : S[g + 1] * 8u;
            seg_end = last ? im.total_slots : uint32_t(g + 1) * im.seg_slots;
        }
    }
    if (WRITE && err) {
#if defined(__CUDA_ARCH__)
        atomicOr(&st->error, err);
#else
        st->error |= err;
#endif
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Dequantise + 8x8 inverse DCT ("islow") of one block: in[64] natural-order coefficients, q[64] natural-order quant
// values, out = 8 rows of 8 samples packed little-endian in two uint32 each.  int32 arithmetic: exact for every stream
// an 8-bit encoder can produce (the reference's JLONG is wider only for corrupt input).
// ------------------------------------------------------------------------------------------------------------------
B2J_HD int idct_descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

B2J_HD uint32_t idct_range_limit(int v)
{
    const int i = v & 1023;
    return uint32_t(i < 128 ? i + 128 : (i < 512 ? 255 : (i < 896 ? 0 : i - 896)));
}

B2J_HD void idct_1d(int i0, int i1, int i2, int i3, int i4, int i5, int i6, int i7, int* o)
{
    int z2 = i2, z3 = i6;
    int z1 = (z2 + z3) * 4433;
    int tmp2 = z1 + z3 * -15137;
    int tmp3 = z1 + z2 * 6270;
    int tmp0 = (i0 + i4) * 8192;
    int tmp1 = (i0 - i4) * 8192;
    const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    tmp0 = i7;
    tmp1 = i5;
    tmp2 = i3;
    tmp3 = i1;
    z1 = tmp0 + tmp3;
    z2 = tmp1 + tmp2;
    z3 = tmp0 + tmp2;
    int z4 = tmp1 + tmp3;
    const int z5 = (z3 + z4) * 9633;
    tmp0 *= 2446;
    tmp1 *= 16819;
    tmp2 *= 25172;
    tmp3 *= 12299;
    z1 *= -7373;
    z2 *= -20995;
    z3 = z3 * -16069 + z5;
    z4 = z4 * -3196 + z5;
    tmp0 += z1 + z3;
    tmp1 += z2 + z4;
    tmp2 += z2 + z3;
    tmp3 += z1 + z4;
    o[0] = tmp10 + tmp3;
    o[7] = tmp10 - tmp3;
    o[1] = tmp11 + tmp2;
    o[6] = tmp11 - tmp2;
    o[2] = tmp12 + tmp1;
    o[5] = tmp12 - tmp1;
    o[3] = tmp13 + tmp0;
    o[4] = tmp13 - tmp0;
}

B2J_HD void idct_block(const int16_t* in, const uint16_t* q, uint32_t* out /*[16]*/)
{
    int ws[64];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        int o[8];
        idct_1d(in[c] * q[c], in[8 + c] * q[8 + c], in[16 + c] * q[16 + c], in[24 + c] * q[24 + c], in[32 + c] * q[32 + c],
            in[40 + c] * q[40 + c], in[48 + c] * q[48 + c], in[56 + c] * q[56 + c], o);
#pragma unroll
        for (int r = 0; r < 8; ++r)
            ws[r * 8 + c] = idct_descale(o[r], 11);
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        int o[8];
        idct_1d(ws[r * 8], ws[r * 8 + 1], ws[r * 8 + 2], ws[r * 8 + 3], ws[r * 8 + 4], ws[r * 8 + 5], ws[r * 8 + 6], ws[r * 8 + 7], o);
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            lo |= idct_range_limit(idct_descale(o[c], 18)) << (8 * c);
            hi |= idct_range_limit(idct_descale(o[c + 4], 18)) << (8 * c);
        }
        out[2 * r] = lo;
        out[2 * r + 1] = hi;
    }
}

// Body of the IDCT kernel: thread u of image im handles one block, enumerated in plane order (component, block row,
// block column) so that neighbouring threads write neighbouring 8-byte row pieces.
B2J_HD void idct_thread(const BatchDev& b, const ImageDesc& im, uint32_t u)
{
    int c = 0;
    uint32_t rem = u;
    for (; c < im.ncomp_out; ++c) {
        const uint32_t nb = uint32_t(im.wib[c]) * uint32_t(im.hib[c]);
        if (rem < nb)
            break;
        rem -= nb;
    }
    if (c >= im.ncomp_out)
        return;
    const int bx = int(rem % uint32_t(im.wib[c])), by = int(rem / uint32_t(im.wib[c]));
    const int mcu = (by / im.vs[c]) * im.mcus_x + (bx / im.hs[c]);
    const int j = im.blk_first[c] + (by % im.vs[c]) * im.hs[c] + (bx % im.hs[c]);
    const int16_t* in = b.coef + im.coef_off + (uint64_t(mcu) * im.bpm + j) * 64;
    const uint16_t* q = b.quant + im.qt[c] * 64;
    alignas(16) int16_t blk[64];
    const uint4* in4 = reinterpret_cast<const uint4*>(in);
    uint4* blk4 = reinterpret_cast<uint4*>(blk);
#pragma unroll
    for (int i = 0; i < 8; ++i)
        blk4[i] = in4[i];
    uint32_t out[16];
    idct_block(blk, q, out);
    uint8_t* dst = b.planes + im.plane_off[c] + uint64_t(by) * 8 * im.plane_pitch[c] + bx * 8;
#pragma unroll
    for (int r = 0; r < 8; ++r)
        *reinterpret_cast<uint2*>(dst + r * im.plane_pitch[c]) = make_uint2(out[2 * r], out[2 * r + 1]);
}

// ------------------------------------------------------------------------------------------------------------------
// Upsampling + colour conversion.
// ------------------------------------------------------------------------------------------------------------------
B2J_HD int upsampled_sample(const BatchDev& b, const ImageDesc& im, int c, int x, int y)
{
    const uint8_t* P = b.planes + im.plane_off[c];
    const int pitch = im.plane_pitch[c];
    const int hx = im.hmax / im.hs[c], vy = im.vmax / im.vs[c];
    if (hx == 1 && vy == 1)
        return P[uint64_t(y) * pitch + x];
    const bool fancy = im.fancy != 0;
    if (vy == 1) { // h2v1
        const uint8_t* in = P + uint64_t(y) * pitch;
        const int cx = x >> 1;
        if (!(fancy && im.ds_w[c] > 2))
            return in[cx];
        if (!(x & 1))
            return cx == 0 ? in[0] : (in[cx] * 3 + in[cx - 1] + 1) >> 2;
        return cx == im.ds_w[c] - 1 ? in[cx] : (in[cx] * 3 + in[cx + 1] + 2) >> 2;
    }
    const int r = y >> 1;
    int far_row = (y & 1) ? r + 1 : r - 1;
    if (far_row < 0)
        far_row = 0;
    if (far_row > im.ds_h[c] - 1)
        far_row = im.ds_h[c] - 1;
    const uint8_t* n0 = P + uint64_t(r) * pitch;
    const uint8_t* n1 = P + uint64_t(far_row) * pitch;
    if (hx == 1) { // h1v2
        if (!fancy)
            return n0[x];
        return (n0[x] * 3 + n1[x] + ((y & 1) ? 2 : 1)) >> 2;
    }
    const int cx = x >> 1; // h2v2
    if (!(fancy && im.ds_w[c] > 2))
        return n0[cx];
    const int cs = n0[cx] * 3 + n1[cx];
    if (!(x & 1)) {
        if (cx == 0)
            return (cs * 4 + 8) >> 4;
        return (cs * 3 + (n0[cx - 1] * 3 + n1[cx - 1]) + 8) >> 4;
    }
    if (cx == im.ds_w[c] - 1)
        return (cs * 4 + 7) >> 4;
    return (cs * 3 + (n0[cx + 1] * 3 + n1[cx + 1]) + 7) >> 4;
}

B2J_HD int clamp_u8(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }

B2J_HD void ycc_to_rgb(int y, int cb, int cr, int& r, int& g, int& bl)
{
    const int xcb = cb - 128, xcr = cr - 128;
    r = clamp_u8(y + ((91881 * xcr + 32768) >> 16));
    bl = clamp_u8(y + ((116130 * xcb + 32768) >> 16));
    g = clamp_u8(y + ((-22554 * xcb - 46802 * xcr + 32768) >> 16));
}

constexpr int kColorPixelsPerThread = 16;

// Body of the colour kernel: thread u of image im produces 16 consecutive pixels of one output row.
B2J_HD void color_thread(const BatchDev& b, const ImageDesc& im, uint32_t u)
{
    const uint32_t units_per_row = (uint32_t(im.width) + kColorPixelsPerThread - 1) / kColorPixelsPerThread;
    const int y = int(u / units_per_row);
    if (y >= im.height)
        return;
    const int x0 = int(u % units_per_row) * kColorPixelsPerThread;
    const int npx = im.width - x0 < kColorPixelsPerThread ? im.width - x0 : kColorPixelsPerThread;
    const int fmt = im.out_fmt;
    uint8_t c0[kColorPixelsPerThread], c1[kColorPixelsPerThread], c2[kColorPixelsPerThread];
    const bool want_rgb = (fmt != 7);
    const bool bgr = (fmt == 5 || fmt == 6);
#pragma unroll 4
    for (int i = 0; i < kColorPixelsPerThread; ++i) {
        int r = 0, g = 0, bl = 0;
        if (i < npx) {
            const int yy = upsampled_sample(b, im, 0, x0 + i, y);
            if (!want_rgb || im.ncomp == 1)
                r = g = bl = yy;
            else
                ycc_to_rgb(yy, upsampled_sample(b, im, 1, x0 + i, y), upsampled_sample(b, im, 2, x0 + i, y), r, g, bl);
        }
        c0[i] = uint8_t(bgr ? bl : r);
        c1[i] = uint8_t(g);
        c2[i] = uint8_t(bgr ? r : bl);
    }
    if (fmt == 4 || fmt == 6) { // interleaved
        uint8_t* dst = im.out + int64_t(y) * im.out_stride + int64_t(x0) * 3;
        if (npx == kColorPixelsPerThread && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
            uint32_t w[12];
#pragma unroll
            for (int k = 0; k < 12; ++k) {
                uint32_t v = 0;
#pragma unroll
                for (int bb = 0; bb < 4; ++bb) {
                    const int idx = k * 4 + bb, px = idx / 3, ch = idx % 3;
                    const uint32_t s = ch == 0 ? c0[px] : (ch == 1 ? c1[px] : c2[px]);
                    v |= s << (8 * bb);
                }
                w[k] = v;
            }
            uint4* d4 = reinterpret_cast<uint4*>(dst);
            d4[0] = make_uint4(w[0], w[1], w[2], w[3]);
            d4[1] = make_uint4(w[4], w[5], w[6], w[7]);
            d4[2] = make_uint4(w[8], w[9], w[10], w[11]);
        } else {
            for (int i = 0; i < npx; ++i) {
                dst[3 * i] = c0[i];
                dst[3 * i + 1] = c1[i];
                dst[3 * i + 2] = c2[i];
            }
        }
    } else { // planar (P_RGB / P_BGR: 3 planes; P_Y: 1 plane)
        const int nplanes = want_rgb ? 3 : 1;
        for (int pl = 0; pl < nplanes; ++pl) {
            uint8_t* dst = im.out + (int64_t(pl) * im.height + y) * im.out_stride + x0;
            const uint8_t* src = pl == 0 ? c0 : (pl == 1 ? c1 : c2);
            if (npx == kColorPixelsPerThread && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
                uint32_t w[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    w[k] = uint32_t(src[4 * k]) | (uint32_t(src[4 * k + 1]) << 8) | (uint32_t(src[4 * k + 2]) << 16) |
                           (uint32_t(src[4 * k + 3]) << 24);
                *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
            } else {
                for (int i = 0; i < npx; ++i)
                    dst[i] = src[i];
            }
        }
    }
}

} // namespace b200jpeg
