// jpeg_math.cuh -- sample-domain arithmetic of the b200jpeg decode path as host+device inline functions: dequantisation
// + integer "islow" IDCT of one block, and chroma upsampling + YCbCr->RGB of one 16-pixel unit.  The CUDA kernels in
// decode_kernels.cu wrap these bodies with the memory staging; compiled with g++ (B2J_HD = inline) the same bodies run
// in the CPU emulation used by the "not gpu" tests (tests/emu).
//
// Reference arithmetic restated here ("ljt/" = external/libjpeg-turbo in the reference tree):
//   dequant + IDCT      ljt/jidctint.c:173-410 (_jpeg_idct_islow), range limit ljt/jdmaster.c:386-407
//   upsampling          ljt/jdsample.c h2v1_fancy:276 h1v2_fancy:316 h2v2_fancy:362 h2v1:205 h2v2:233, selector :414-535,
//                       edge rows ljt/jdmainct.c:217-250; merged (non-fancy) path ljt/jdmrgext.c:24-184
//   colour              ljt/jdcolor.c:215-251, ljt/jdcolext.c:30-76
#pragma once
#include "device_types.h"
#if !defined(__CUDACC__)
#include "host_shim.h"
#endif

namespace b200jpeg {

// ------------------------------------------------------------------------------------------------------------------
// Dequantise + 8x8 inverse DCT ("islow") of one block: in[64] natural-order coefficients (in[0] ignored, the DC value is
// passed separately), q[64] natural-order quant values, out = 8 rows of 8 samples packed little-endian in two uint32
// each.  int32 arithmetic: exact for every stream an 8-bit encoder can produce (the reference's JLONG is wider only for
// corrupt input).
// ------------------------------------------------------------------------------------------------------------------
B2J_HD int idct_descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// range_limit[(x) & RANGE_MASK] of the reference (ljt/jdmaster.c:386-407): a clamp of x + 128 to [0, 255] for
// x in [-512, 511], and the table's wrap-around outside of that
B2J_HD uint32_t idct_range_limit(int v)
{
    int t = int(uint32_t(v) << 22) >> 22; // sign-extend the low 10 bits: wraps exactly like the & 1023 table index
    t += 128;
    t = t < 0 ? 0 : (t > 255 ? 255 : t);
    return uint32_t(t);
}

// One 1-D pass (ljt/jidctint.c:197-291 / :303-399).  rnd is added to both even-part sums: every output contains exactly one
// of them exactly once, so the rounding constant of the descale that follows (1 << (n - 1)) is added here twice instead of
// eight times after the butterflies -- the same integers mod 2^32.
B2J_HD void idct_1d(int i0, int i1, int i2, int i3, int i4, int i5, int i6, int i7, int rnd, int* o)
{
    int z2 = i2, z3 = i6;
    int z1 = (z2 + z3) * 4433;
    int tmp2 = z1 + z3 * -15137;
    int tmp3 = z1 + z2 * 6270;
    int tmp0 = (i0 + i4) * 8192 + rnd;
    int tmp1 = (i0 - i4) * 8192 + rnd;
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

// Sample from a second-pass output that already carries the rounding constant: descale by 18 bits, range_limit.  Bits
// 18..27 of v, sign-extended, are exactly the table index the reference forms ((x >> 18) & RANGE_MASK read as a signed
// 10-bit number); + 128 and a clamp to [0, 255] are what the table holds there.
B2J_HD uint32_t idct_sample(int v)
{
    const int t = (int(uint32_t(v) << 4) >> 22) + 128;
#if defined(__CUDA_ARCH__)
    return uint32_t(__vimin_s32_relu(t, 255));
#else
    return uint32_t(t < 0 ? 0 : (t > 255 ? 255 : t));
#endif
}

B2J_HD void idct_block(const int16_t* in, int dcval, const uint16_t* q, uint32_t* out /*[16]*/)
{
    int ws[64];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        int o[8];
        const int i0 = (c == 0 ? dcval : int(in[c])) * q[c];
        idct_1d(i0, in[8 + c] * q[8 + c], in[16 + c] * q[16 + c], in[24 + c] * q[24 + c], in[32 + c] * q[32 + c],
            in[40 + c] * q[40 + c], in[48 + c] * q[48 + c], in[56 + c] * q[56 + c], 1 << 10, o);
#pragma unroll
        for (int r = 0; r < 8; ++r)
            ws[r * 8 + c] = o[r] >> 11;
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        int o[8];
        idct_1d(ws[r * 8], ws[r * 8 + 1], ws[r * 8 + 2], ws[r * 8 + 3], ws[r * 8 + 4], ws[r * 8 + 5], ws[r * 8 + 6], ws[r * 8 + 7], 1 << 17, o);
        const uint32_t a = idct_sample(o[0]) + (idct_sample(o[1]) << 8), b = idct_sample(o[2]) + (idct_sample(o[3]) << 8);
        const uint32_t c = idct_sample(o[4]) + (idct_sample(o[5]) << 8), d = idct_sample(o[6]) + (idct_sample(o[7]) << 8);
        out[2 * r] = a + (b << 16);
        out[2 * r + 1] = c + (d << 16);
    }
}

// Where block u (MCU scan order, the order of the stream: ljt/jdcoefct.c decompress_onepass:86-169) of image im goes to.
struct IdctBlockRef
{
    int comp;          // -1: no such block, or a component the output does not need
    uint32_t src_blk;  // = u
    uint64_t dst;      // byte offset of the block's top-left sample in the planes buffer
    int pitch;
    int bx, by;        // block coordinates inside the component plane
};

B2J_HD IdctBlockRef idct_locate(const ImageDesc& im, uint32_t u)
{
    IdctBlockRef r;
    r.comp = -1;
    r.src_blk = u;
    r.dst = 0;
    r.pitch = 0;
    r.bx = r.by = 0;
    if (u >= im.total_blocks)
        return r;
    const uint32_t mcu = u / uint32_t(im.bpm), j = u - mcu * uint32_t(im.bpm);
    const int c = im.blk_comp[j];
    if (c >= im.ncomp_out)
        return r;
    const int my = int(mcu / uint32_t(im.mcus_x)), mx = int(mcu - uint32_t(my) * uint32_t(im.mcus_x));
    const int jj = int(j) - im.blk_first[c];
    const int hs = im.hs[c], lh = hs >> 1; // sampling factors are 1, 2 or 4: log2 = f >> 1
    r.comp = c;
    r.bx = mx * hs + (jj & (hs - 1));
    r.by = my * im.vs[c] + (jj >> lh);
    r.pitch = im.plane_pitch[c];
    r.dst = im.plane_off[c] + uint64_t(r.by) * 8 * uint64_t(r.pitch) + uint64_t(r.bx) * 8;
    return r;
}

// ------------------------------------------------------------------------------------------------------------------
// Upsampling + colour conversion of one unit: 16 horizontally adjacent pixels of one output row (VY == 1) or of the
// output row pair (2q+1, 2q+2) (VY == 2; this pairing needs only chroma rows q and q+1).
// ------------------------------------------------------------------------------------------------------------------
B2J_HD int clamp_u8(int v)
{
#if defined(__CUDA_ARCH__)
    return __vimin_s32_relu(v, 255);
#else
    return v < 0 ? 0 : (v > 255 ? 255 : v);
#endif
}

B2J_HD void ycc_to_rgb(int y, int cb, int cr, int& r, int& g, int& bl)
{
    const int xcb = cb - 128, xcr = cr - 128;
    r = clamp_u8(y + ((91881 * xcr + 32768) >> 16));
    bl = clamp_u8(y + ((116130 * xcb + 32768) >> 16));
    g = clamp_u8(y + ((-22554 * xcb - 46802 * xcr + 32768) >> 16));
}

B2J_HD void load16(const uint8_t* p, int* v) // p is 8-byte aligned
{
    const uint2 a = *reinterpret_cast<const uint2*>(p), c = *reinterpret_cast<const uint2*>(p + 8);
    const uint32_t w[4] = {a.x, a.y, c.x, c.y};
#pragma unroll
    for (int i = 0; i < 16; ++i)
        v[i] = int((w[i >> 2] >> (8 * (i & 3))) & 255u);
}

// chroma samples [c0 - 1, c0 + N] of a plane row with the column index clamped to [0, dsw - 1]; v[0] is column c0 - 1
template <int N>
B2J_HD void load_chroma_row(const uint8_t* row, int c0, int dsw, int* v /*[N + 2]*/)
{
    if (c0 >= 1 && c0 + N <= dsw - 1) {
        v[0] = row[c0 - 1];
        v[N + 1] = row[c0 + N];
        if constexpr (N == 8) {
            const uint2 a = *reinterpret_cast<const uint2*>(row + c0);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                v[1 + i] = int((a.x >> (8 * i)) & 255u);
                v[5 + i] = int((a.y >> (8 * i)) & 255u);
            }
        } else {
            load16(row + c0, v + 1);
        }
    } else {
#pragma unroll
        for (int i = 0; i < N + 2; ++i) {
            int c = c0 - 1 + i;
            c = c < 0 ? 0 : (c > dsw - 1 ? dsw - 1 : c);
            v[i] = row[c];
        }
    }
}

// Full-resolution chroma of the unit for one component: up[r][i], r < VY, i < 16.
template <int HX, int VY>
B2J_HD void upsample_unit(const uint8_t* P, int pitch, int dsw, int dsh, bool fancy, int q, int x0, int up[][16])
{
    constexpr int NC = 16 / HX;
    const int c0 = x0 / HX;
    if constexpr (VY == 1) {
        const uint8_t* row = P + uint64_t(q) * pitch;
        if constexpr (HX == 1) {
            load16(row + c0, up[0]);
        } else if constexpr (HX == 4) { // 4:1:1: each chroma sample covers four pixels (jdsample.c int_upsample:161-203)
            const uint32_t w = *reinterpret_cast<const uint32_t*>(row + c0);
#pragma unroll
            for (int i = 0; i < 16; ++i)
                up[0][i] = int((w >> (8 * (i >> 2))) & 255u);
        } else {
            int a[NC + 2];
            load_chroma_row<NC>(row, c0, dsw, a);
            const bool f = fancy && dsw > 2;
#pragma unroll
            for (int i = 0; i < NC; ++i) {
                up[0][2 * i] = f ? (3 * a[i + 1] + a[i] + 1) >> 2 : a[i + 1];
                up[0][2 * i + 1] = f ? (3 * a[i + 1] + a[i + 2] + 2) >> 2 : a[i + 1];
            }
        }
    } else {
        int ra = q < 0 ? 0 : (q > dsh - 1 ? dsh - 1 : q), rb = q + 1 > dsh - 1 ? dsh - 1 : q + 1;
        const uint8_t* rowa = P + uint64_t(ra) * pitch;
        const uint8_t* rowb = P + uint64_t(rb) * pitch;
        if constexpr (HX == 4) { // 4:1:0: plain replication in both directions (jdsample.c int_upsample:161-203)
            const uint32_t wa = *reinterpret_cast<const uint32_t*>(rowa + c0), wb = *reinterpret_cast<const uint32_t*>(rowb + c0);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                up[0][i] = int((wa >> (8 * (i >> 2))) & 255u);
                up[1][i] = int((wb >> (8 * (i >> 2))) & 255u);
            }
        } else if constexpr (HX == 1) {
            int a[16], bb[16];
            load16(rowa + c0, a);
            load16(rowb + c0, bb);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                up[0][i] = fancy ? (3 * a[i] + bb[i] + 2) >> 2 : a[i];
                up[1][i] = fancy ? (3 * bb[i] + a[i] + 1) >> 2 : bb[i];
            }
        } else {
            int a[NC + 2], bb[NC + 2];
            load_chroma_row<NC>(rowa, c0, dsw, a);
            load_chroma_row<NC>(rowb, c0, dsw, bb);
            const bool f = fancy && dsw > 2;
            if (f) {
                int ca[NC + 2], cb[NC + 2];
#pragma unroll
                for (int i = 0; i < NC + 2; ++i) {
                    ca[i] = 3 * a[i] + bb[i];
                    cb[i] = 3 * bb[i] + a[i];
                }
#pragma unroll
                for (int i = 0; i < NC; ++i) {
                    up[0][2 * i] = (3 * ca[i + 1] + ca[i] + 8) >> 4;
                    up[0][2 * i + 1] = (3 * ca[i + 1] + ca[i + 2] + 7) >> 4;
                    up[1][2 * i] = (3 * cb[i + 1] + cb[i] + 8) >> 4;
                    up[1][2 * i + 1] = (3 * cb[i + 1] + cb[i + 2] + 7) >> 4;
                }
            } else {
#pragma unroll
                for (int i = 0; i < NC; ++i) {
                    up[0][2 * i] = up[0][2 * i + 1] = a[i + 1];
                    up[1][2 * i] = up[1][2 * i + 1] = bb[i + 1];
                }
            }
        }
    }
}

// px[r * 16 + i] = c0 | c1 << 8 | c2 << 16 of pixel x0 + i of output row yrow[r]; channel order already swapped for BGR
// outputs, Y replicated for gray streams; Y-only output (P_Y) has just c0.  yrow[r] < 0 marks a row outside the image.
template <int HX, int VY>
B2J_HD void color_unit(const BatchDev& b, const ImageDesc& im, int q, int x0, uint32_t* px, int* yrow)
{
    if constexpr (VY == 2) {
        yrow[0] = 2 * q + 1;
        yrow[1] = 2 * q + 2;
    } else {
        yrow[0] = q;
    }
#pragma unroll
    for (int r = 0; r < VY; ++r)
        if (yrow[r] < 0 || yrow[r] >= im.height)
            yrow[r] = -1;
    int yv[VY][16];
#pragma unroll
    for (int r = 0; r < VY; ++r) {
        if (yrow[r] >= 0)
            load16(b.planes + im.plane_off[0] + uint64_t(yrow[r]) * im.plane_pitch[0] + x0, yv[r]);
        else {
#pragma unroll
            for (int i = 0; i < 16; ++i)
                yv[r][i] = 0;
        }
    }
    const int fmt = im.out_fmt;
    if (im.ncomp == 1 || fmt == 7) {
#pragma unroll
        for (int r = 0; r < VY; ++r)
#pragma unroll
            for (int i = 0; i < 16; ++i)
                px[r * 16 + i] = uint32_t(yv[r][i]) * 0x010101u;
        return;
    }
    int cbv[VY][16], crv[VY][16];
    upsample_unit<HX, VY>(b.planes + im.plane_off[1], im.plane_pitch[1], im.ds_w[1], im.ds_h[1], im.fancy != 0, q, x0, cbv);
    upsample_unit<HX, VY>(b.planes + im.plane_off[2], im.plane_pitch[2], im.ds_w[2], im.ds_h[2], im.fancy != 0, q, x0, crv);
    const bool bgr = (fmt == 5 || fmt == 6);
#pragma unroll
    for (int r = 0; r < VY; ++r)
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            int rr, gg, bl;
            ycc_to_rgb(yv[r][i], cbv[r][i], crv[r][i], rr, gg, bl);
            px[r * 16 + i] = bgr ? (uint32_t(bl) | (uint32_t(gg) << 8) | (uint32_t(rr) << 16))
                                 : (uint32_t(rr) | (uint32_t(gg) << 8) | (uint32_t(bl) << 16));
        }
}

// Where pixel (x, y) of the w x h image (region) lands in the oriented output, and that output's size (see ImageDesc::orient).
B2J_HD void orient_pixel(int orient, int w, int h, int x, int y, int& ox, int& oy)
{
    const int rot = orient & 3;
    int ow = w, oh = h;
    if (rot == 1) { // 90 degrees counter-clockwise
        ox = y;
        oy = w - 1 - x;
        ow = h;
        oh = w;
    } else if (rot == 2) {
        ox = w - 1 - x;
        oy = h - 1 - y;
    } else if (rot == 3) { // 270 degrees counter-clockwise = 90 clockwise
        ox = h - 1 - y;
        oy = x;
        ow = h;
        oh = w;
    } else {
        ox = x;
        oy = y;
    }
    if (orient & 4)
        ox = ow - 1 - ox;
    if (orient & 8)
        oy = oh - 1 - oy;
}

// Plain store of a unit (used by the CPU emulation; the CUDA kernel stages through shared memory instead).
template <int VY>
B2J_HD void store_unit_simple(const ImageDesc& im, int x0, const uint32_t* px, const int* yrow)
{
    const int fmt = im.out_fmt;
    const int npx = im.width - x0 < 16 ? im.width - x0 : 16;
    for (int r = 0; r < VY; ++r) {
        const int oy = yrow[r] - im.roi_y; // row inside the region of interest
        if (yrow[r] < 0 || oy < 0 || oy >= im.roi_h)
            continue;
        for (int i = 0; i < npx; ++i) {
            int ox = x0 + i - im.roi_x;
            if (ox < 0 || ox >= im.roi_w)
                continue;
            int oyy = oy, out_h = im.roi_h;
            if (im.orient) {
                orient_pixel(im.orient, im.roi_w, im.roi_h, ox, oy, ox, oyy);
                out_h = (im.orient & 1) ? im.roi_w : im.roi_h;
            }
            if (fmt == 4 || fmt == 6) {
                uint8_t* dst = im.out + int64_t(oyy) * im.out_stride + int64_t(ox) * 3;
                dst[0] = uint8_t(px[r * 16 + i]);
                dst[1] = uint8_t(px[r * 16 + i] >> 8);
                dst[2] = uint8_t(px[r * 16 + i] >> 16);
            } else {
                const int nplanes = fmt == 7 ? 1 : 3;
                for (int pl = 0; pl < nplanes; ++pl)
                    im.out[(int64_t(pl) * out_h + oyy) * im.out_stride + ox] = uint8_t(px[r * 16 + i] >> (8 * pl));
            }
        }
    }
}

// sampling class of an image for the colour kernel: chroma horizontal / vertical expansion
B2J_HD int color_hx(const ImageDesc& im) { return im.ncomp == 1 ? 1 : im.hmax / im.hs[1]; }
B2J_HD int color_vy(const ImageDesc& im) { return im.ncomp == 1 ? 1 : im.vmax / im.vs[1]; }

} // namespace b200jpeg
