// color_kernel.cu -- C1 k_color: chroma upsampling (fancy / box) + YCbCr->RGB of b200jpeg.
//
// Work unit: a warp tile = 512 pixels x VY output rows (VY = 2 when the chroma planes are vertically subsampled: the row
// pair (2q+1, 2q+2) needs only chroma rows q and q+1).  Each lane converts 16 pixels per row: all its loads (luma rows,
// chroma rows, the two edge neighbours per chroma row) are issued first, then the rows are converted one after the other,
// assembled in shared memory and written with one 4-byte word per lane per store instruction (128 contiguous bytes per
// instruction) whatever the alignment of the destination row.
// Arithmetic: jpeg_math.cuh states it sample by sample (that statement runs in the CPU emulation of the tests); this
// file is the register-level version of the same formulas.  Reference: ljt/jdsample.c h2v1_fancy:276 h1v2_fancy:316
// h2v2_fancy:362 h2v1:205 h2v2:233, ljt/jdcolor.c:215-251, ljt/jdcolext.c:30-76.
// Edge rule: a chroma column index is clamped to [0, ds_w - 1].  The left/right neighbours of a lane's 8 columns are
// loaded with clamped addresses; column ds_w inside a lane's 8 columns was overwritten with a copy of column ds_w - 1 by
// the IDCT kernel (decode_kernels.cu), so no per-sample clamping is needed here.
#include <cuda_runtime.h>
#include "decode_kernels.h"
#include "jpeg_math.cuh"
#include "cta_lookup.cuh"

namespace b200jpeg {

__device__ __forceinline__ int byte_of(uint32_t w, int i) { return int(__byte_perm(w, 0, 0x4440 + i)); }

constexpr int kRowWords = kColorTilePx * 3 / 4; // interleaved RGB row of a tile in 32-bit words
constexpr int kRowStride = kRowWords + 4;       // spare word for the unaligned copy, keeps 16-byte alignment
constexpr int kPlaneStride = kColorTilePx / 4 + 4;

// copy nbytes from shared memory (s_row is word aligned, one spare word readable past the end of the row; the copy starts
// sofs bytes into it) to dst (any alignment)
__device__ __forceinline__ void warp_store_row(uint8_t* dst, const uint32_t* s_row, int nbytes, uint32_t lane, int sofs = 0)
{
    const uint32_t m = uint32_t(reinterpret_cast<uintptr_t>(dst) & 3u);
    if (sofs) { // region of interest that starts inside the tile: source and destination are both unaligned
        const uint8_t* sb = reinterpret_cast<const uint8_t*>(s_row) + sofs;
        int head = int((4u - m) & 3u);
        if (head > nbytes)
            head = nbytes;
        if (int(lane) < head)
            dst[lane] = sb[lane];
        const int nw = (nbytes - head) >> 2;
        uint32_t* d32 = reinterpret_cast<uint32_t*>(dst + head);
        const int s0 = sofs + head;
        const uint32_t* sw = s_row + (s0 >> 2);
        const uint32_t sh = uint32_t(s0 & 3) * 8u;
        for (int w = int(lane); w < nw; w += 32)
            d32[w] = __funnelshift_r(sw[w], sw[w + 1], sh);
        const int done = head + (nw << 2);
        if (int(lane) < nbytes - done)
            dst[done + lane] = sb[done + lane];
        return;
    }
    const uint8_t* sb = reinterpret_cast<const uint8_t*>(s_row);
    if (m == 0) {
        const int nw = nbytes >> 2;
        uint32_t* d32 = reinterpret_cast<uint32_t*>(dst);
        for (int w = int(lane); w < nw; w += 32)
            d32[w] = s_row[w];
        const int done = nw << 2;
        if (int(lane) < nbytes - done)
            dst[done + lane] = sb[done + lane];
    } else {
        int head = int(4u - m);
        if (head > nbytes)
            head = nbytes;
        if (int(lane) < head)
            dst[lane] = sb[lane];
        const int nw = (nbytes - head) >> 2;
        uint32_t* d32 = reinterpret_cast<uint32_t*>(dst + head);
        const uint32_t sh = uint32_t(head) * 8u;
        for (int w = int(lane); w < nw; w += 32)
            d32[w] = __funnelshift_r(s_row[w], s_row[w + 1], sh);
        const int done = head + (nw << 2);
        if (int(lane) < nbytes - done)
            dst[done + lane] = sb[done + lane];
    }
}

// one chroma row segment of a lane: HX == 2: 8 samples + left/right neighbour; HX == 1: 16 samples; HX == 4: 4 samples
template <int HX>
struct ChromaSeg
{
    uint32_t w[HX == 2 ? 2 : (HX == 4 ? 1 : 4)];
    int left, right; // HX == 2 only
};

template <int HX>
__device__ __forceinline__ void load_seg(ChromaSeg<HX>& s, const uint8_t* row, int c0, int dsw)
{
    if constexpr (HX == 2) {
        const uint2 v = *reinterpret_cast<const uint2*>(row + c0);
        s.w[0] = v.x;
        s.w[1] = v.y;
        s.left = row[c0 > 0 ? c0 - 1 : 0];
        s.right = row[c0 + 8 < dsw ? c0 + 8 : dsw - 1];
    } else if constexpr (HX == 4) {
        s.w[0] = *reinterpret_cast<const uint32_t*>(row + c0);
        s.left = s.right = 0;
    } else {
        const uint2 v = *reinterpret_cast<const uint2*>(row + c0), u = *reinterpret_cast<const uint2*>(row + c0 + 8);
        s.w[0] = v.x;
        s.w[1] = v.y;
        s.w[2] = u.x;
        s.w[3] = u.y;
        s.left = s.right = 0;
    }
}

// full-resolution chroma of output row r (0: row 2q+1 / the only row, 1: row 2q+2) from the lane's segments
template <int HX, int VY>
__device__ __forceinline__ void upsample_row(const ChromaSeg<HX>& a, const ChromaSeg<HX>& bseg, int r, bool fancy_h, bool fancy_v, int* up)
{
    if constexpr (HX == 2) {
        int n[10], f[10]; // near / far row, columns c0-1 .. c0+8
        const ChromaSeg<HX>& nr = (VY == 2 && r == 1) ? bseg : a;
        const ChromaSeg<HX>& fr = (VY == 2 && r == 1) ? a : bseg;
        n[0] = nr.left;
        n[9] = nr.right;
#pragma unroll
        for (int i = 0; i < 8; ++i)
            n[1 + i] = byte_of(nr.w[i >> 2], i & 3);
        if constexpr (VY == 2) {
            if (fancy_h) { // h2v2 fancy: vertical 3:1 blend, then horizontal 3:1 blend, one rounding at the end
                f[0] = fr.left;
                f[9] = fr.right;
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    f[1 + i] = byte_of(fr.w[i >> 2], i & 3);
#pragma unroll
                for (int i = 0; i < 10; ++i)
                    n[i] = 3 * n[i] + f[i];
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    up[2 * c] = (3 * n[c + 1] + n[c] + 8) >> 4;
                    up[2 * c + 1] = (3 * n[c + 1] + n[c + 2] + 7) >> 4;
                }
                return;
            }
        } else {
            if (fancy_h) { // h2v1 fancy
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    up[2 * c] = (3 * n[c + 1] + n[c] + 1) >> 2;
                    up[2 * c + 1] = (3 * n[c + 1] + n[c + 2] + 2) >> 2;
                }
                return;
            }
        }
#pragma unroll
        for (int c = 0; c < 8; ++c)
            up[2 * c] = up[2 * c + 1] = n[c + 1]; // box
    } else if constexpr (HX == 4) { // 4:1:1 / 4:1:0: plain replication (jdsample.c int_upsample), whatever fancy_upsampling says
        const ChromaSeg<HX>& nr = (VY == 2 && r == 1) ? bseg : a;
#pragma unroll
        for (int i = 0; i < 16; ++i)
            up[i] = byte_of(nr.w[0], i >> 2);
    } else {
        const ChromaSeg<HX>& nr = (VY == 2 && r == 1) ? bseg : a;
        const ChromaSeg<HX>& fr = (VY == 2 && r == 1) ? a : bseg;
#pragma unroll
        for (int i = 0; i < 16; ++i)
            up[i] = byte_of(nr.w[i >> 2], i & 3);
        if constexpr (VY == 2) {
            if (fancy_v) { // h1v2 fancy: bias 2 for the odd output row (r == 0), 1 for the even one
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    up[i] = (3 * up[i] + byte_of(fr.w[i >> 2], i & 3) + (r == 0 ? 2 : 1)) >> 2;
            }
        }
    }
}

// 16 pixels -> px[i] = c0 | c1 << 8 | c2 << 16 (channel order already swapped for BGR outputs)
__device__ __forceinline__ void convert16(const uint32_t* yw, const int* cb, const int* cr, bool bgr, uint32_t* px)
{
    constexpr int kR = 32768 - 91881 * 128, kB = 32768 - 116130 * 128, kG = 32768 + (22554 + 46802) * 128;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int y = byte_of(yw[i >> 2], i & 3);
        const int r = clamp_u8(y + ((91881 * cr[i] + kR) >> 16));
        const int bl = clamp_u8(y + ((116130 * cb[i] + kB) >> 16));
        const int g = clamp_u8(y + ((-22554 * cb[i] - 46802 * cr[i] + kG) >> 16));
        const uint32_t lo = bgr ? uint32_t(bl) : uint32_t(r), hi = bgr ? uint32_t(r) : uint32_t(bl);
        px[i] = __byte_perm(__byte_perm(lo, uint32_t(g), 0x4440), hi, 0x4410); // lo | g << 8 | hi << 16
    }
}

// stage one converted row of the lane in shared memory in the output layout
__device__ __forceinline__ void stage_row(const uint32_t* px, int fmt, uint32_t* s_buf, uint32_t lane)
{
    if (fmt == 4 || fmt == 6) { // interleaved: 4 pixels -> 3 words
        uint32_t w[12];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint32_t a = px[4 * g], c = px[4 * g + 1], d = px[4 * g + 2], e = px[4 * g + 3];
            w[3 * g] = __byte_perm(a, c, 0x4210);     // a0 a1 a2 c0
            w[3 * g + 1] = __byte_perm(c, d, 0x5421); // c1 c2 d0 d1
            w[3 * g + 2] = __byte_perm(d, e, 0x6542); // d2 e0 e1 e2
        }
        uint4* s4 = reinterpret_cast<uint4*>(s_buf + lane * 12);
        s4[0] = make_uint4(w[0], w[1], w[2], w[3]);
        s4[1] = make_uint4(w[4], w[5], w[6], w[7]);
        s4[2] = make_uint4(w[8], w[9], w[10], w[11]);
    } else { // planar: plane pl of the tile row at s_buf + pl * kPlaneStride
        const int nplanes = fmt == 7 ? 1 : 3;
        for (int pl = 0; pl < nplanes; ++pl) {
            uint32_t w[4];
            const uint32_t sel = 0x40u + uint32_t(pl) * 0x11u; // result byte 0 = byte pl of x, byte 1 = byte pl of y
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const uint32_t lo = __byte_perm(px[4 * g], px[4 * g + 1], sel) & 0xFFFFu; // p0.pl | p1.pl << 8
                const uint32_t hi = __byte_perm(px[4 * g + 2], px[4 * g + 3], sel) & 0xFFFFu;
                w[g] = lo | (hi << 16);
            }
            *reinterpret_cast<uint4*>(s_buf + pl * kPlaneStride + lane * 4) = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
}

// Oriented output (ImageDesc::orient): the lane's 16 pixels of one image row go to wherever the rotation / mirroring puts
// them, one pixel at a time.  A rotated row is a column of the output, so there is nothing to coalesce; oriented files are
// rare and this path only has to be right.
__device__ __forceinline__ void store_row_oriented(const ImageDesc& im, const uint32_t* px, int x0, int y, int fmt)
{
    const int oy0 = y - im.roi_y, out_h = (im.orient & 1) ? im.roi_w : im.roi_h;
    for (int i = 0; i < 16; ++i) {
        const int x = x0 + i - im.roi_x;
        if (x < 0 || x >= im.roi_w || x0 + i >= im.width)
            continue;
        int ox, oy;
        orient_pixel(im.orient, im.roi_w, im.roi_h, x, oy0, ox, oy);
        if (fmt == 4 || fmt == 6) {
            uint8_t* dst = im.out + int64_t(oy) * im.out_stride + int64_t(ox) * 3;
            dst[0] = uint8_t(px[i]);
            dst[1] = uint8_t(px[i] >> 8);
            dst[2] = uint8_t(px[i] >> 16);
        } else {
            const int nplanes = fmt == 7 ? 1 : 3;
            for (int pl = 0; pl < nplanes; ++pl)
                im.out[(int64_t(pl) * out_h + oy) * im.out_stride + ox] = uint8_t(px[i] >> (8 * pl));
        }
    }
}

template <int HX, int VY>
__device__ __forceinline__ void color_tile(const BatchDev& b, const ImageDesc& im, uint32_t tile, uint32_t* s_warp, uint32_t lane)
{
    const int q = int(tile / uint32_t(im.color_xtiles)) - (VY == 2 ? 1 : 0);
    const int xt = int(tile % uint32_t(im.color_xtiles));
    const int xbase = xt * kColorTilePx;
    const int x0 = xbase + int(lane) * 16;
    const bool active = x0 < im.width;
    const int fmt = im.out_fmt;
    const bool gray = im.ncomp == 1 || fmt == 7;
    int yrow[VY];
    if constexpr (VY == 2) {
        yrow[0] = 2 * q + 1;
        yrow[1] = 2 * q + 2;
    } else {
        yrow[0] = q;
    }
    bool valid[VY]; // rows of the image that are inside the region of interest
#pragma unroll
    for (int r = 0; r < VY; ++r)
        valid[r] = yrow[r] >= im.roi_y && yrow[r] < im.roi_y + im.roi_h;
    // the part of the tile inside the region of interest
    const int tile_end = im.width - xbase < kColorTilePx ? im.width : xbase + kColorTilePx;
    const int xa = xbase > im.roi_x ? xbase : im.roi_x, xb = tile_end < im.roi_x + im.roi_w ? tile_end : im.roi_x + im.roi_w;
    if (xb <= xa || !(valid[0] || valid[VY - 1]))
        return;

    // ---- loads: everything the lane needs is requested before anything is used ----
    uint32_t yw[VY][4];
    ChromaSeg<HX> cba, cbb, cra, crb;
    if (active) {
#pragma unroll
        for (int r = 0; r < VY; ++r) {
            const int yy = yrow[r] < 0 ? 0 : (yrow[r] > im.height - 1 ? im.height - 1 : yrow[r]);
            const uint8_t* p = b.planes + im.plane_off[0] + uint64_t(yy) * im.plane_pitch[0] + x0;
            const uint2 v = *reinterpret_cast<const uint2*>(p), u = *reinterpret_cast<const uint2*>(p + 8);
            yw[r][0] = v.x;
            yw[r][1] = v.y;
            yw[r][2] = u.x;
            yw[r][3] = u.y;
        }
        if (!gray) {
            const int c0 = x0 / HX;
            const int dsh = im.ds_h[1], dsw = im.ds_w[1];
            int ra, rb;
            if constexpr (VY == 2) {
                ra = q < 0 ? 0 : (q > dsh - 1 ? dsh - 1 : q);
                rb = q + 1 > dsh - 1 ? dsh - 1 : q + 1;
            } else {
                ra = rb = q;
            }
            const uint8_t* pcb = b.planes + im.plane_off[1];
            const uint8_t* pcr = b.planes + im.plane_off[2];
            load_seg<HX>(cba, pcb + uint64_t(ra) * im.plane_pitch[1], c0, dsw);
            load_seg<HX>(cra, pcr + uint64_t(ra) * im.plane_pitch[2], c0, dsw);
            if constexpr (VY == 2) {
                load_seg<HX>(cbb, pcb + uint64_t(rb) * im.plane_pitch[1], c0, dsw);
                load_seg<HX>(crb, pcr + uint64_t(rb) * im.plane_pitch[2], c0, dsw);
            }
        }
    }
    const bool fancy_h = im.fancy != 0 && im.ds_w[1] > 2; // h2v1 / h2v2 (jdsample.c:472,501)
    const bool fancy_v = im.fancy != 0;                   // h1v2
    const bool bgr = (fmt == 5 || fmt == 6);
    const bool interleaved = (fmt == 4 || fmt == 6);
    const int row_stride_words = interleaved ? kRowStride : 3 * kPlaneStride;

    // ---- convert row by row into the staging buffer ----
#pragma unroll
    for (int r = 0; r < VY; ++r) {
        if (active && valid[r]) {
            uint32_t* s_buf = s_warp + r * row_stride_words;
            if (fmt == 7 && !im.orient) { // Y only: the luma bytes are the output
                *reinterpret_cast<uint4*>(s_buf + lane * 4) = make_uint4(yw[r][0], yw[r][1], yw[r][2], yw[r][3]);
            } else {
                uint32_t px[16];
                if (gray) {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        px[i] = uint32_t(byte_of(yw[r][i >> 2], i & 3)) * 0x010101u;
                } else {
                    int cb[16], cr[16];
                    upsample_row<HX, VY>(cba, cbb, r, fancy_h, fancy_v, cb);
                    upsample_row<HX, VY>(cra, crb, r, fancy_h, fancy_v, cr);
                    convert16(yw[r], cb, cr, bgr, px);
                }
                if (im.orient)
                    store_row_oriented(im, px, x0, yrow[r], fmt);
                else
                    stage_row(px, fmt, s_buf, lane);
            }
        }
    }
    __syncwarp();
    if (im.orient)
        return;
    // ---- coalesced copy-out ----
#pragma unroll
    for (int r = 0; r < VY; ++r) {
        if (!valid[r])
            continue;
        const uint32_t* s_buf = s_warp + r * row_stride_words;
        const int oy = yrow[r] - im.roi_y, ox = xa - im.roi_x, npx = xb - xa, skip = xa - xbase;
        if (interleaved) {
            warp_store_row(im.out + int64_t(oy) * im.out_stride + int64_t(ox) * 3, s_buf, npx * 3, lane, skip * 3);
        } else {
            const int nplanes = fmt == 7 ? 1 : 3;
            for (int pl = 0; pl < nplanes; ++pl)
                warp_store_row(im.out + (int64_t(pl) * im.roi_h + oy) * im.out_stride + ox, s_buf + pl * kPlaneStride, npx, lane, skip);
        }
    }
    __syncwarp();
}

constexpr int kColorWarpWords = 2 * 3 * kPlaneStride > 2 * kRowStride ? 2 * 3 * kPlaneStride : 2 * kRowStride;

#ifndef B2J_COLOR_MIN_CTAS
#define B2J_COLOR_MIN_CTAS 8 // resident CTAs per SM the register allocation aims at
#endif
__global__ void __launch_bounds__(kColorCta, B2J_COLOR_MIN_CTAS) k_color(BatchDev b)
{
    __shared__ __align__(16) uint32_t s_stage[(kColorCta / 32) * kColorWarpWords];
    const uint32_t cta_abs = blockIdx.x + b.base_color;
    const int img = find_image_warp(b.cta_map_color, b.num_images, cta_abs);
    const ImageDesc& im = b.images[img];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tile = (cta_abs - b.cta_map_color[img]) * (kColorCta / 32) + warp;
    if (tile >= uint32_t(im.color_rows) * uint32_t(im.color_xtiles))
        return;
    uint32_t* s_warp = s_stage + warp * kColorWarpWords;
    const int hx = color_hx(im), vy = color_vy(im);
    if (vy == 2) {
        if (hx == 4)
            color_tile<4, 2>(b, im, tile, s_warp, lane);
        else if (hx == 2)
            color_tile<2, 2>(b, im, tile, s_warp, lane);
        else
            color_tile<1, 2>(b, im, tile, s_warp, lane);
    } else {
        if (hx == 4)
            color_tile<4, 1>(b, im, tile, s_warp, lane);
        else if (hx == 2)
            color_tile<2, 1>(b, im, tile, s_warp, lane);
        else
            color_tile<1, 1>(b, im, tile, s_warp, lane);
    }
}

int launch_color(const BatchDev& b, const LaunchDims& d, cudaStream_t st)
{
    k_color<<<d.color_ctas, kColorCta, 0, st>>>(b);
    return 1;
}

} // namespace b200jpeg
