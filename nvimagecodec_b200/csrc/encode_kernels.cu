// encode_kernels.cu -- sm_100a kernels of the b200jpeg encode path.
//
//   E1a k_enc_color       4 pixels x vs rows per thread: RGB -> YCbCr (or gray) with the reference's edge replication and
//                         chroma box downsampling -> component planes padded to whole MCUs
//   E1b k_enc_fdct        one 8x8 block per thread (tiles of kEncCta blocks in MCU scan order): level shift + islow FDCT +
//                         reciprocal quantisation -> zigzag int16 coefficients, the block's DC and the entropy-coded length
//                         of its AC part; dummy edge blocks take the DC of their source block
//   E2 k_enc_bits         bits of every block (DC category code + value bits + AC part), exclusive scan inside each tile of
//                         kEncCta blocks; k_enc_tilescan: one warp per image scans the tile totals
//   E3 k_enc_segscan      one CTA per image: bytes of every restart segment, byte offsets of the segments, stream length
//   E4 k_enc_emit         one block per thread: Huffman-codes its coefficients into the CTA's shared-memory window of the
//                         stream (OR-ed words), pads segment ends with 1-bits; window -> global, shared edge words by atomicOr
//   E5 k_enc_stuff_count / k_enc_stuff_scan / k_enc_stuff_write
//                         0xFF -> 0xFF 0x00 and RSTn insertion: per-chunk output sizes, per-image scan, compaction
// The arithmetic lives in enc_math.cuh (shared with the CPU emulation of the tests).
#include <cuda_runtime.h>
#include "enc_math.cuh"
#include "encode_kernels.h"
#include "cta_lookup.cuh"

namespace b200jpeg {

// ------------------------------------------------------------------------------------------------------------------
// E1a: colour conversion + downsampling -> component planes padded to whole MCUs
// ------------------------------------------------------------------------------------------------------------------
// Four pixels of one source row (x clamped to the last column: expand_right_edge; the caller clamps the row).
// The source image as the colour kernel needs it, copied out of the descriptor once (the kernel's stores could alias the
// descriptor as far as the compiler knows, and it would re-read it after every store).
struct EncSrc
{
    const uint8_t* __restrict__ pixels;
    int64_t stride;
    int width, height, in_fmt;
    bool aligned4;
};

__device__ __forceinline__ void enc_load_rgb4(const EncSrc& im, int y, int sx, bool aligned4, uint32_t* r, uint32_t* g, uint32_t* bl)
{
    const int W = im.width;
    const uint8_t* row = im.pixels + int64_t(y) * im.stride;
    if (im.in_fmt == 4) { // B200JPEG_FMT_I_RGB
        if (sx + 3 < W && aligned4) {
            const uint32_t* p = reinterpret_cast<const uint32_t*>(row + sx * 3);
            const uint32_t w0 = __ldg(p), w1 = __ldg(p + 1), w2 = __ldg(p + 2);
            r[0] = __byte_perm(w0, 0, 0x4440); g[0] = __byte_perm(w0, 0, 0x4441); bl[0] = __byte_perm(w0, 0, 0x4442);
            r[1] = __byte_perm(w0, 0, 0x4443); g[1] = __byte_perm(w1, 0, 0x4440); bl[1] = __byte_perm(w1, 0, 0x4441);
            r[2] = __byte_perm(w1, 0, 0x4442); g[2] = __byte_perm(w1, 0, 0x4443); bl[2] = __byte_perm(w2, 0, 0x4440);
            r[3] = __byte_perm(w2, 0, 0x4441); g[3] = __byte_perm(w2, 0, 0x4442); bl[3] = __byte_perm(w2, 0, 0x4443);
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint8_t* p = row + min(sx + i, W - 1) * 3;
                r[i] = __ldg(p); g[i] = __ldg(p + 1); bl[i] = __ldg(p + 2);
            }
        }
    } else { // planar RGB
        const int64_t plane_stride = im.stride * int64_t(im.height);
        if (sx + 3 < W && aligned4) {
            const uint32_t wr = __ldg(reinterpret_cast<const uint32_t*>(row + sx));
            const uint32_t wg = __ldg(reinterpret_cast<const uint32_t*>(row + plane_stride + sx));
            const uint32_t wb = __ldg(reinterpret_cast<const uint32_t*>(row + 2 * plane_stride + sx));
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                r[i] = (wr >> (8 * i)) & 255u; g[i] = (wg >> (8 * i)) & 255u; bl[i] = (wb >> (8 * i)) & 255u;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int x = min(sx + i, W - 1);
                r[i] = __ldg(row + x); g[i] = __ldg(row + plane_stride + x); bl[i] = __ldg(row + 2 * plane_stride + x);
            }
        }
    }
}

// One thread: 4 pixels x vs source rows -> 4 x vs luma samples and 4 / hs chroma samples per chroma component.
// Rows past the image repeat the last row (expand_bottom_edge before downsampling); chroma rows past the last row group
// repeat that group's result (expand_bottom_edge after downsampling), so their source rows differ from the luma rows.
template <int HS, int VS>
__device__ __forceinline__ void enc_color_unit(const EncBatchDev& b, const EncImageDesc& desc, int ux, int uy)
{
    EncSrc im;
    im.pixels = desc.pixels;
    im.stride = desc.stride;
    im.width = desc.width;
    im.height = desc.height;
    im.in_fmt = desc.in_fmt;
    const int ncomp = desc.ncomp, pitch_y = desc.plane_pitch[0], pitch_c = desc.plane_pitch[1];
    uint8_t* const yp = b.planes + desc.plane_off[0];
    uint8_t* const cbp = b.planes + desc.plane_off[1] + int64_t(uy) * pitch_c;
    uint8_t* const crp = b.planes + desc.plane_off[2] + int64_t(uy) * pitch_c;
    const int W = im.width, H = im.height, sx = ux * 4;
    const bool aligned4 = ((reinterpret_cast<uintptr_t>(im.pixels) | uint64_t(im.stride)) & 3u) == 0;
    if (ncomp == 1) {
        const uint8_t* row = im.pixels + int64_t(min(uy, H - 1)) * im.stride;
        uint32_t yv = 0;
        if (sx + 3 < W && aligned4)
            yv = *reinterpret_cast<const uint32_t*>(row + sx);
        else
#pragma unroll
            for (int i = 0; i < 4; ++i)
                yv |= uint32_t(row[min(sx + i, W - 1)]) << (8 * i);
        *reinterpret_cast<uint32_t*>(yp + int64_t(uy) * pitch_y + sx) = yv;
        return;
    }
    const int rows_ds = (H + VS - 1) / VS;
    const int cgrp = min(uy, rows_ds - 1);
    uint32_t cb[VS][4], cr[VS][4];
#pragma unroll
    for (int dy = 0; dy < VS; ++dy) {
        const int ly = min(uy * VS + dy, H - 1), cy = min(cgrp * VS + dy, H - 1);
        uint32_t r[4], g[4], bl[4];
        enc_load_rgb4(im, ly, sx, aligned4, r, g, bl);
        uint32_t yv = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            yv |= enc_luma(r[i], g[i], bl[i]) << (8 * i);
        *reinterpret_cast<uint32_t*>(yp + int64_t(uy * VS + dy) * pitch_y + sx) = yv;
        if (cy != ly)
            enc_load_rgb4(im, cy, sx, aligned4, r, g, bl);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            cb[dy][i] = enc_cb(int(r[i]), int(g[i]), int(bl[i]));
            cr[dy][i] = enc_cr(int(r[i]), int(g[i]), int(bl[i]));
        }
    }
    if (HS == 2) { // two outputs at even / odd chroma columns: biases 1, 2 (h2v2) or 0, 1 (h2v1)
        uint32_t ob, orr;
        if (VS == 2) {
            ob = ((cb[0][0] + cb[0][1] + cb[VS - 1][0] + cb[VS - 1][1] + 1u) >> 2) | (((cb[0][2] + cb[0][3] + cb[VS - 1][2] + cb[VS - 1][3] + 2u) >> 2) << 8);
            orr = ((cr[0][0] + cr[0][1] + cr[VS - 1][0] + cr[VS - 1][1] + 1u) >> 2) | (((cr[0][2] + cr[0][3] + cr[VS - 1][2] + cr[VS - 1][3] + 2u) >> 2) << 8);
        } else {
            ob = ((cb[0][0] + cb[0][1]) >> 1) | (((cb[0][2] + cb[0][3] + 1u) >> 1) << 8);
            orr = ((cr[0][0] + cr[0][1]) >> 1) | (((cr[0][2] + cr[0][3] + 1u) >> 1) << 8);
        }
        *reinterpret_cast<uint16_t*>(cbp + ux * 2) = uint16_t(ob);
        *reinterpret_cast<uint16_t*>(crp + ux * 2) = uint16_t(orr);
    } else {
        uint32_t ob = 0, orr = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t vb = VS == 2 ? (cb[0][i] + cb[VS - 1][i] + 1u) >> 1 : cb[0][i];
            const uint32_t vr = VS == 2 ? (cr[0][i] + cr[VS - 1][i] + 1u) >> 1 : cr[0][i];
            ob |= vb << (8 * i);
            orr |= vr << (8 * i);
        }
        *reinterpret_cast<uint32_t*>(cbp + sx) = ob;
        *reinterpret_cast<uint32_t*>(crp + sx) = orr;
    }
}

__global__ void __launch_bounds__(kEncColorCta) k_enc_color(EncBatchDev b)
{
    const uint32_t cta_abs = blockIdx.x + b.cta_base_color;
    const int img = find_image_warp(b.cta_map_color, b.num_images, cta_abs);
    const EncImageDesc& im = b.images[img];
    const uint32_t u = (cta_abs - b.cta_map_color[img]) * kEncColorCta + threadIdx.x;
    const uint32_t units_x = uint32_t(im.plane_pitch[0]) >> 2;
    const uint32_t uy = u / units_x, ux = u - uy * units_x;
    const int hs = im.ncomp == 1 ? 1 : im.hs, vs = im.ncomp == 1 ? 1 : im.vs;
    if (uy >= uint32_t(im.mcus_y) * 8u) // one unit row per chroma row (gray: hs = vs = 1)
        return;
    if (hs == 2 && vs == 2)
        enc_color_unit<2, 2>(b, im, int(ux), int(uy));
    else if (hs == 2)
        enc_color_unit<2, 1>(b, im, int(ux), int(uy));
    else if (vs == 2)
        enc_color_unit<1, 2>(b, im, int(ux), int(uy));
    else
        enc_color_unit<1, 1>(b, im, int(ux), int(uy));
}

// ------------------------------------------------------------------------------------------------------------------
// E1b: level shift + islow FDCT + quantisation, one block per thread (tiles of kEncCta blocks in MCU scan order)
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kEncCta, 4) k_enc_fdct(EncBatchDev b)
{
    __shared__ uint2 s_q[2][64];
    __shared__ uint8_t s_aclen[2][256];
    const uint32_t cta_abs = blockIdx.x + b.cta_base_emit;
    const int img = find_image_warp(b.cta_map_emit, b.num_images, cta_abs);
    const EncImageDesc& im = b.images[img];
    {
        const EncQuantTable* q = b.quant + im.qt;
        for (int i = threadIdx.x; i < 128; i += kEncCta) {
            const EncQuantTable& t = q[i >> 6];
            s_q[i >> 6][i & 63] = make_uint2(t.recip[i & 63], uint32_t(t.corr[i & 63]) | (uint32_t(t.shift[i & 63]) << 16));
        }
        for (int i = threadIdx.x; i < 512; i += kEncCta)
            s_aclen[i >> 8][i & 255] = uint8_t(b.huff->ac[i >> 8][i & 255] & 255u);
    }
    __syncthreads();
    const uint32_t blk = (cta_abs - b.cta_map_emit[img]) * kEncCta + threadIdx.x;
    if (blk >= im.total_blocks)
        return;
    const int ncomp = im.ncomp, hs = im.hs, vs = im.vs;
    const uint32_t mcu = blk / uint32_t(im.bpm);
    const int j = int(blk - mcu * uint32_t(im.bpm));
    const int my = int(mcu / uint32_t(im.mcus_x)), mx = int(mcu - uint32_t(my) * uint32_t(im.mcus_x));
    int comp = 0, first_j = 0, count = 1;
    enc_block_comp(im, j, comp, first_j, count);
    const int mw = comp == 0 && ncomp == 3 ? hs : 1, mh = comp == 0 && ncomp == 3 ? vs : 1;
    int jj = j - first_j;
    // dummy block (ljt/jccoefct.c:176-196): no AC, the DC of the block before it in the MCU; a whole dummy row copies the
    // block before the row.  The source may itself be a dummy of the first kind; blocks at (0, 0) of a component are real.
    bool real = true;
    for (;;) {
        const int sby = jj / mw, sbx = jj - sby * mw;
        const bool row_dummy = my * mh + sby >= im.hib[comp];
        const bool col_dummy = mx * mw + sbx >= im.wib[comp];
        if (!row_dummy && !col_dummy)
            break;
        real = false;
        jj = row_dummy ? sby * mw - 1 : jj - 1;
    }
    const int by = jj / mw, bx = jj - by * mw;
    const int pitch = im.plane_pitch[comp];
    const uint8_t* src = b.planes + im.plane_off[comp] + int64_t((my * mh + by) * 8) * pitch + (mx * mw + bx) * 8;
    uint32_t rows[16];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const uint2 v = *reinterpret_cast<const uint2*>(src + r * pitch);
        rows[2 * r] = v.x;
        rows[2 * r + 1] = v.y;
    }
    const EncQuantPacked qt{s_q[comp ? 1 : 0]};
    uint4* dst = reinterpret_cast<uint4*>(b.coef + (im.blk_first + blk) * 64);
    if (!real) {
        // the DC term of the islow FDCT is the plain sum of the level-shifted samples: (4 * sum + 2) >> 2
        uint32_t sum = 0;
#pragma unroll
        for (int i = 0; i < 16; ++i)
            sum = __dp4a(rows[i], 0x01010101u, sum);
        const int dcq = enc_quantise(int(sum) - 128 * 64, qt.get_recip(0), qt.get_corr(0), qt.get_shift(0));
        dst[0] = make_uint4(uint32_t(uint16_t(dcq)), 0, 0, 0);
#pragma unroll
        for (int i = 1; i < 8; ++i)
            dst[i] = make_uint4(0, 0, 0, 0);
        b.dc[im.blk_first + blk] = int16_t(dcq);
        b.acbits[im.blk_first + blk] = uint16_t(s_aclen[comp ? 1 : 0][0]); // EOB only
        b.nzmask[im.blk_first + blk] = 0;
        return;
    }
    int16_t zz[64];
    enc_fdct_quant(rows, qt, zz);
    // entropy-coded length of the AC part (same walk as the emit kernel, lengths only)
    const uint8_t* aclen = s_aclen[comp ? 1 : 0];
    uint32_t acbits = 0, mlo = 0, mhi = 0;
    int run = 0;
#pragma unroll
    for (int k = 1; k < 64; ++k) {
        const int v = zz[k];
        if (v == 0) {
            ++run;
        } else {
            acbits += uint32_t(run >> 4) * aclen[0xF0];
            const int n = enc_category(v);
            acbits += uint32_t(aclen[(run & 15) * 16 + n]) + uint32_t(n);
            run = 0;
            if (k < 32)
                mlo |= 1u << k;
            else
                mhi |= 1u << (k - 32);
        }
    }
    if (run)
        acbits += aclen[0];
    b.nzmask[im.blk_first + blk] = uint64_t(mlo) | (uint64_t(mhi) << 32);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        uint32_t w[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
            w[k] = uint32_t(uint16_t(zz[8 * i + 2 * k])) | (uint32_t(uint16_t(zz[8 * i + 2 * k + 1])) << 16);
        dst[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
    b.dc[im.blk_first + blk] = zz[0];
    b.acbits[im.blk_first + blk] = uint16_t(acbits);
}

// ------------------------------------------------------------------------------------------------------------------
// E2/E3: bit offsets
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int enc_dc_diff(const EncBatchDev& b, const EncImageDesc& im, uint32_t blk)
{
    const int64_t p = enc_dc_pred_block(im, blk);
    const int cur = b.dc[im.blk_first + blk];
    return p < 0 ? cur : cur - int(b.dc[im.blk_first + uint64_t(p)]);
}

// Bit position of block k in the image's entropy-coded data if restart segments were not byte aligned: a plain prefix
// sum over all blocks of the image, kept as a per-tile base (tile = the kEncCta blocks of one emit CTA) plus the offset
// inside the tile.  k == total_blocks gives the image's total.  Arithmetic is modulo 2^32 (differences are what is used).
__device__ __forceinline__ uint32_t enc_prefix_bits(const EncBatchDev& b, const EncImageDesc& im, int img, uint32_t k)
{
    if (k >= im.total_blocks)
        return b.img_bits[img];
    return b.tile_off[im.emit_cta_first + k / kEncCta] + b.blk_off[im.blk_first + k];
}

// absolute bit position of block k inside the image's unstuffed stream (segments start on byte boundaries)
__device__ __forceinline__ uint64_t enc_block_bitpos(const EncBatchDev& b, const EncImageDesc& im, int img, uint32_t k)
{
    const uint32_t seg = k / im.seg_blocks;
    const uint32_t rel = enc_prefix_bits(b, im, img, k) - enc_prefix_bits(b, im, img, seg * im.seg_blocks);
    return uint64_t(b.seg_byte_off[im.seg_first + seg]) * 8u + rel;
}

// E2a: bits of every block (DC category code + value bits + AC part) and their exclusive scan inside the tile
__global__ void __launch_bounds__(kEncCta) k_enc_bits(EncBatchDev b)
{
    __shared__ uint32_t s_warp[kEncCta / 32];
    const uint32_t cta_abs = blockIdx.x + b.cta_base_emit;
    const int img = find_image_warp(b.cta_map_emit, b.num_images, cta_abs);
    const EncImageDesc& im = b.images[img];
    const uint32_t blk = (cta_abs - b.cta_map_emit[img]) * kEncCta + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t bits = 0;
    if (blk < im.total_blocks) {
        const int j = int(blk % uint32_t(im.bpm));
        int comp, first_j, count;
        enc_block_comp(im, j, comp, first_j, count);
        const int n = enc_category(enc_dc_diff(b, im, blk));
        bits = uint32_t(b.acbits[im.blk_first + blk]) + (b.huff->dc[comp ? 1 : 0][n] & 255u) + uint32_t(n);
    }
    uint32_t incl = bits;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= uint32_t(o))
            incl += v;
    }
    if (lane == 31)
        s_warp[warp] = incl;
    __syncthreads();
    uint32_t pre = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kEncCta / 32; ++w) {
        const uint32_t v = s_warp[w];
        pre += uint32_t(w) < warp ? v : 0u;
        total += v;
    }
    if (blk < im.total_blocks)
        b.blk_off[im.blk_first + blk] = pre + incl - bits;
    if (threadIdx.x == 0)
        b.tile_off[cta_abs] = total;
}

// E2b: one warp per image: exclusive scan of the tile totals
__global__ void __launch_bounds__(32) k_enc_tilescan(EncBatchDev b)
{
    const int img = blockIdx.x;
    const uint32_t first = b.cta_map_emit[img], ntiles = b.cta_map_emit[img + 1] - first;
    const uint32_t lane = threadIdx.x;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < ntiles; base += 32) {
        const uint32_t t = base + lane;
        const uint32_t v = t < ntiles ? b.tile_off[first + t] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= uint32_t(o))
                incl += u;
        }
        if (t < ntiles)
            b.tile_off[first + t] = carry + incl - v;
        carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
    if (lane == 0)
        b.img_bits[img] = carry;
}

// E3: one CTA per image: byte length of every restart segment -> byte offsets of the segments, unstuffed stream length
constexpr int kEncSegScanThreads = 256;
__global__ void __launch_bounds__(kEncSegScanThreads) k_enc_segscan(EncBatchDev b)
{
    __shared__ uint32_t s_warp[kEncSegScanThreads / 32];
    __shared__ uint32_t s_carry;
    const int img = blockIdx.x;
    const EncImageDesc& im = b.images[img];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0)
        s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < im.num_segments; base += kEncSegScanThreads) {
        const uint32_t s = base + threadIdx.x;
        uint32_t bytes = 0;
        if (s < im.num_segments) {
            const uint32_t lo = s * im.seg_blocks;
            const uint32_t bits = enc_prefix_bits(b, im, img, lo + im.seg_blocks) - enc_prefix_bits(b, im, img, lo);
            bytes = (bits + 7u) >> 3;
        }
        uint32_t incl = bytes;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= uint32_t(o))
                incl += v;
        }
        if (lane == 31)
            s_warp[warp] = incl;
        __syncthreads();
        uint32_t pre = s_carry, total = 0;
#pragma unroll
        for (int w = 0; w < kEncSegScanThreads / 32; ++w) {
            const uint32_t v = s_warp[w];
            pre += uint32_t(w) < warp ? v : 0u;
            total += v;
        }
        if (s < im.num_segments)
            b.seg_byte_off[im.seg_first + s] = pre + incl - bytes;
        __syncthreads();
        if (threadIdx.x == 0)
            s_carry += total;
        __syncthreads();
    }
    if (threadIdx.x == 0)
        b.img_stream_bytes[img] = s_carry;
}

// ------------------------------------------------------------------------------------------------------------------
// E4: Huffman emit
// ------------------------------------------------------------------------------------------------------------------
// The CTA's window of the stream in shared memory: 64 bytes per block on average, i.e. any tile whose blocks average no more
// than 512 bits (anything but near-lossless qualities).  A tile that needs more ORs its bits straight into global memory.
constexpr int kEncWindowWords = kEncCta * 16 + 8;

// Same interface as EncWordSink; the words go to the (zeroed) global stream in stream byte order, all by atomicOr.
struct EncGlobalSink
{
    uint32_t* words;
    uint32_t widx;
    uint64_t acc;
    int n;
    __device__ __forceinline__ void start(uint32_t* w, uint32_t bitpos)
    {
        words = w;
        widx = bitpos >> 5;
        acc = 0;
        n = int(bitpos & 31u);
    }
    __device__ __forceinline__ void or_word(uint32_t v)
    {
        if (v)
            atomicOr(words + widx, __byte_perm(v, 0, 0x0123));
        ++widx;
    }
    __device__ __forceinline__ void put(uint32_t code, int len)
    {
        acc |= uint64_t(code) << (64 - n - len);
        n += len;
        if (n >= 32) {
            or_word(uint32_t(acc >> 32));
            acc <<= 32;
            n -= 32;
        }
    }
    __device__ __forceinline__ void finish()
    {
        if (n)
            or_word(uint32_t(acc >> 32));
    }
};

template <class Sink>
__device__ __forceinline__ void enc_emit_block(const EncBatchDev& b, const EncImageDesc& im, uint32_t blk, const uint32_t (*s_dc)[16],
    const uint32_t (*s_ac)[256], Sink& sink)
{
    const int16_t* cf = b.coef + (im.blk_first + blk) * 64; // one 128-byte line: read through L1, non-zero positions only
    const uint64_t mask = b.nzmask[im.blk_first + blk];
    const int j = int(blk % uint32_t(im.bpm));
    int comp, first_j, count;
    enc_block_comp(im, j, comp, first_j, count);
    const int tsel = comp ? 1 : 0;
    enc_put_dc(enc_dc_diff(b, im, blk), s_dc[tsel], sink);
    enc_put_ac_sparse(mask, [cf](int k) { return int(__ldg(cf + k)); }, s_ac[tsel], sink);
    // last block of a restart segment (or of the image): fill the byte with 1-bits (ljt/jchuff.c flush_bits:480)
    const bool seg_last = (blk + 1 == im.total_blocks) || ((blk + 1) % im.seg_blocks == 0);
    if (seg_last) {
        const int used = sink.n & 7; // the window starts on a 32-bit boundary of a byte-aligned stream
        if (used)
            sink.put((1u << (8 - used)) - 1u, 8 - used);
    }
    sink.finish();
}

__global__ void __launch_bounds__(kEncCta) k_enc_emit(EncBatchDev b)
{
    __shared__ uint32_t s_win[kEncWindowWords];
    __shared__ uint32_t s_ac[2][256];
    __shared__ uint32_t s_dc[2][16];
    const uint32_t cta_abs = blockIdx.x + b.cta_base_emit;
    const int img = find_image_warp(b.cta_map_emit, b.num_images, cta_abs);
    const EncImageDesc& im = b.images[img];
    const EncImageDyn& dy = b.dyn[img];
    const uint32_t cta = cta_abs - b.cta_map_emit[img];
    const uint32_t blk0 = cta * kEncCta, blk = blk0 + threadIdx.x;
    const uint32_t blk_end = min(blk0 + kEncCta, im.total_blocks);
    if (blk < blk_end) // start fetching the block's coefficient line; it is first needed after the set-up below
        asm volatile("prefetch.global.L1 [%0];" ::"l"(b.coef + (im.blk_first + blk) * 64));
    for (int i = threadIdx.x; i < 512; i += kEncCta)
        s_ac[i >> 8][i & 255] = b.huff->ac[i >> 8][i & 255];
    if (threadIdx.x < 32)
        s_dc[threadIdx.x >> 4][threadIdx.x & 15] = b.huff->dc[threadIdx.x >> 4][threadIdx.x & 15];
    // absolute bit position of a block inside the image's unstuffed stream
    auto bitpos = [&](uint32_t k) -> uint64_t { return enc_block_bitpos(b, im, img, k); };
    const uint64_t win_lo_bit = bitpos(blk0) & ~uint64_t(31);
    const uint64_t win_hi_bit = blk_end == im.total_blocks ? uint64_t(dy.stream_bytes) * 8u : bitpos(blk_end);
    const uint32_t win_words = uint32_t((win_hi_bit - win_lo_bit + 31u) >> 5);
    const bool in_smem = win_words <= uint32_t(kEncWindowWords);
    const bool last_shared = (win_hi_bit & 31u) != 0; // the last word continues in the next tile (or is the image's last)
    uint32_t* gw = reinterpret_cast<uint32_t*>(b.stream + dy.stream_off) + (win_lo_bit >> 5);
    if (in_smem) {
        for (uint32_t i = threadIdx.x; i < win_words; i += kEncCta)
            s_win[i] = 0;
    } else { // the words only this tile writes; k_enc_zero_edges cleared the shared ones
        for (uint32_t i = 1 + threadIdx.x; i < win_words; i += kEncCta)
            if (!(i + 1 == win_words && last_shared))
                gw[i] = 0;
    }
    __syncthreads();
    if (blk < blk_end) {
        const uint32_t rel = uint32_t(bitpos(blk) - win_lo_bit);
        if (in_smem) {
            EncWordSink sink;
            sink.start(s_win, rel);
            enc_emit_block(b, im, blk, s_dc, s_ac, sink);
        } else {
            EncGlobalSink sink;
            sink.start(gw, rel);
            enc_emit_block(b, im, blk, s_dc, s_ac, sink);
        }
    }
    if (!in_smem)
        return;
    __syncthreads();
    // window -> global stream (bytes in stream order): interior words plain, the two edge words may be shared with the
    // neighbouring CTAs' windows -> atomicOr on memory the zeroing kernel cleared
    for (uint32_t i = threadIdx.x; i < win_words; i += kEncCta) {
        const uint32_t v = __byte_perm(s_win[i], 0, 0x0123);
        if (i == 0 || (i + 1 == win_words && last_shared))
            atomicOr(gw + i, v);
        else
            gw[i] = v;
    }
}

// clears the stream words two emit CTAs may both write (the first and last word of every CTA's window)
__global__ void k_enc_zero_edges(EncBatchDev b, uint32_t total_ctas)
{
    if (blockIdx.x * blockDim.x + threadIdx.x >= total_ctas)
        return;
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x + b.cta_base_emit;
    const int img = find_image_scalar(b.cta_map_emit, b.num_images, c); // one CTA per thread here
    const EncImageDesc& im = b.images[img];
    const EncImageDyn& dy = b.dyn[img];
    const uint32_t blk0 = (c - b.cta_map_emit[img]) * kEncCta;
    const uint64_t pos = enc_block_bitpos(b, im, img, blk0);
    uint32_t* gw = reinterpret_cast<uint32_t*>(b.stream + dy.stream_off);
    gw[pos >> 5] = 0;
    if (blk0 == 0) { // and the image's last word(s)
        const uint32_t words = (dy.stream_bytes + 3u) >> 2;
        if (words)
            gw[words - 1] = 0;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// E5: byte stuffing + restart markers
// ------------------------------------------------------------------------------------------------------------------
// number of restart segments (other than the first) that start inside [lo, hi) of the image's unstuffed stream
__device__ __forceinline__ uint32_t enc_seg_lower_bound(const uint32_t* seg_off, uint32_t nseg, uint32_t v)
{
    uint32_t lo = 0, hi = nseg; // first index with seg_off[i] >= v
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (seg_off[mid] < v)
            lo = mid + 1;
        else
            hi = mid;
    }
    return lo;
}

__device__ __forceinline__ uint32_t enc_count_ff(const uint4& v)
{
    auto ff = [](uint32_t w) { // bytes equal to 0xFF
        const uint32_t t = w & (w >> 4) & 0x0F0F0F0Fu; // low nibble & high nibble
        const uint32_t u = t & (t >> 2);
        const uint32_t s = u & (u >> 1) & 0x01010101u;
        return __popc(s);
    };
    return ff(v.x) + ff(v.y) + ff(v.z) + ff(v.w);
}

__global__ void __launch_bounds__(kEncStuffThreads) k_enc_stuff_count(EncBatchDev b)
{
    const int img = find_image_warp(b.cta_map_stuff, b.num_images, blockIdx.x);
    const EncImageDesc& im = b.images[img];
    const EncImageDyn& dy = b.dyn[img];
    const uint32_t chunk = blockIdx.x - dy.chunk_first;
    const uint32_t lo = chunk * kEncStuffChunk, hi = min(lo + kEncStuffChunk, dy.stream_bytes);
    const uint32_t off = lo + threadIdx.x * 16;
    uint32_t n = 0;
    if (off < hi) {
        uint4 v = *reinterpret_cast<const uint4*>(b.stream + dy.stream_off + off);
        const uint32_t valid = hi - off;
        if (valid < 16) { // clear the bytes past the end of the stream
            uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int keep = int(valid) - 4 * i;
                w[i] = keep >= 4 ? w[i] : (keep <= 0 ? 0u : (w[i] & ((1u << (8 * keep)) - 1u)));
            }
            v = make_uint4(w[0], w[1], w[2], w[3]);
        }
        n = (hi - off < 16 ? hi - off : 16u) + enc_count_ff(v);
    }
    __shared__ uint32_t s_sum[kEncStuffThreads / 32];
#pragma unroll
    for (int o = 16; o; o >>= 1)
        n += __shfl_xor_sync(0xFFFFFFFFu, n, o);
    if ((threadIdx.x & 31) == 0)
        s_sum[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t total = 0;
        for (int i = 0; i < kEncStuffThreads / 32; ++i)
            total += s_sum[i];
        // RSTn markers in front of the segments that start in this chunk
        const uint32_t* so = b.seg_byte_off + im.seg_first;
        const uint32_t a = enc_seg_lower_bound(so, im.num_segments, max(lo, 1u)), z = enc_seg_lower_bound(so, im.num_segments, hi);
        total += 2u * (z - a);
        b.chunk_out[blockIdx.x] = total;
    }
}

__global__ void __launch_bounds__(32) k_enc_stuff_scan(EncBatchDev b)
{
    const int img = blockIdx.x;
    const EncImageDyn& dy = b.dyn[img];
    const uint32_t nchunks = (dy.stream_bytes + kEncStuffChunk - 1) / kEncStuffChunk;
    const uint32_t lane = threadIdx.x;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < nchunks; base += 32) {
        const uint32_t c = base + lane;
        const uint32_t v = c < nchunks ? b.chunk_out[dy.chunk_first + c] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= uint32_t(o))
                incl += u;
        }
        if (c < nchunks)
            b.chunk_out[dy.chunk_first + c] = carry + incl - v;
        carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
    if (lane == 0)
        b.img_out_bytes[img] = carry;
}

__global__ void __launch_bounds__(kEncStuffThreads) k_enc_stuff_write(EncBatchDev b)
{
    const int img = find_image_warp(b.cta_map_stuff, b.num_images, blockIdx.x);
    const EncImageDesc& im = b.images[img];
    const EncImageDyn& dy = b.dyn[img];
    const uint32_t chunk = blockIdx.x - dy.chunk_first;
    const uint32_t lo = chunk * kEncStuffChunk, hi = min(lo + kEncStuffChunk, dy.stream_bytes);
    const uint32_t off = lo + threadIdx.x * 16;
    const uint32_t* so = b.seg_byte_off + im.seg_first;
    uint8_t bytes[16];
    uint32_t valid = 0, n = 0, seg_next = 0;
    if (off < hi) {
        const uint4 v = *reinterpret_cast<const uint4*>(b.stream + dy.stream_off + off);
        *reinterpret_cast<uint4*>(bytes) = v;
        valid = min(16u, hi - off);
        seg_next = enc_seg_lower_bound(so, im.num_segments, max(off, 1u)); // first segment start >= off (never segment 0)
        uint32_t sn = seg_next;
        for (uint32_t i = 0; i < valid; ++i) {
            if (sn < im.num_segments && so[sn] == off + i) {
                n += 2;
                ++sn;
            }
            n += bytes[i] == 0xFF ? 2u : 1u;
        }
    }
    // exclusive scan of n over the CTA
    __shared__ uint32_t s_warp[kEncStuffThreads / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= uint32_t(o))
            incl += u;
    }
    if (lane == 31)
        s_warp[warp] = incl;
    __syncthreads();
    uint32_t pre = 0;
    for (uint32_t i = 0; i < warp; ++i)
        pre += s_warp[i];
    uint8_t* dst = b.out + dy.out_off + b.chunk_out[blockIdx.x] + pre + incl - n;
    for (uint32_t i = 0; i < valid; ++i) {
        if (seg_next < im.num_segments && so[seg_next] == off + i) {
            *dst++ = 0xFF;
            *dst++ = uint8_t(0xD0 + ((seg_next - 1) & 7u));
            ++seg_next;
        }
        *dst++ = bytes[i];
        if (bytes[i] == 0xFF)
            *dst++ = 0;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Launchers
// ------------------------------------------------------------------------------------------------------------------
int launch_enc_dct(const EncBatchDev& b, const EncLaunchDims& d, cudaStream_t st)
{
    k_enc_color<<<d.color_ctas, kEncColorCta, 0, st>>>(b);
    k_enc_fdct<<<d.emit_ctas, kEncCta, 0, st>>>(b);
    return 2;
}

int launch_enc_scan(const EncBatchDev& b, const EncLaunchDims& d, cudaStream_t st)
{
    k_enc_bits<<<d.emit_ctas, kEncCta, 0, st>>>(b);
    k_enc_tilescan<<<d.num_images, 32, 0, st>>>(b);
    k_enc_segscan<<<d.num_images, kEncSegScanThreads, 0, st>>>(b);
    return 3;
}

int launch_enc_emit(const EncBatchDev& b, const EncLaunchDims& d, cudaStream_t st)
{
    k_enc_zero_edges<<<(d.emit_ctas + 255) / 256, 256, 0, st>>>(b, d.emit_ctas);
    k_enc_emit<<<d.emit_ctas, kEncCta, 0, st>>>(b);
    return 2;
}

int launch_enc_stuff(const EncBatchDev& b, const EncLaunchDims& d, cudaStream_t st)
{
    if (d.stuff_ctas == 0)
        return 0;
    k_enc_stuff_count<<<d.stuff_ctas, kEncStuffThreads, 0, st>>>(b);
    k_enc_stuff_scan<<<d.num_images, 32, 0, st>>>(b);
    k_enc_stuff_write<<<d.stuff_ctas, kEncStuffThreads, 0, st>>>(b);
    return 3;
}

} // namespace b200jpeg
