// encode_kernels.cu -- sm_100a kernels of the b200jpeg encode path.
//
//   E1 k_enc_dct          per CTA a run of MCUs of one MCU row: RGB -> YCbCr (or gray) into shared memory with the
//                         reference's edge replication, chroma box downsampling, then one 8x8 block per thread: level
//                         shift + islow FDCT + reciprocal quantisation -> zigzag int16 coefficients (coalesced through a
//                         shared-memory staging area), the block's DC and the entropy-coded length of its AC part
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

namespace b200jpeg {

__device__ __forceinline__ int enc_find_image(const uint32_t* __restrict__ map, int n, uint32_t v)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (map[mid] <= v)
            lo = mid;
        else
            hi = mid - 1;
    }
    return lo;
}

// ------------------------------------------------------------------------------------------------------------------
// E1: colour conversion + downsampling + FDCT + quantisation
// ------------------------------------------------------------------------------------------------------------------
constexpr int kEncTilePixels = 8192; // gray: 128 MCUs x 8 x 8; colour tiles have at most 4096 pixels

__global__ void __launch_bounds__(kEncCta) k_enc_dct(EncBatchDev b)
{
    __shared__ __align__(16) uint8_t s_y[kEncTilePixels];  // luma tile; for colour only the first 4096 bytes are used ...
    __shared__ __align__(16) uint8_t s_c[2][2048];          // downsampled (or 4:4:4) chroma, pitch = tile width / hs
    __shared__ __align__(16) uint32_t s_coef[kEncCta * kEncCoefPitch];
    __shared__ EncQuantTable s_q[2];
    __shared__ uint8_t s_aclen[2][256];
    // ... colour: full-resolution chroma lives in the second half of s_y's 8 KB while it waits to be downsampled
    const uint32_t cta_abs = blockIdx.x + b.cta_base_dct;
    const int img = enc_find_image(b.cta_map_dct, b.num_images, cta_abs);
    const EncImageDesc& im = b.images[img];
    const uint32_t cta = cta_abs - b.cta_map_dct[img];
    const int my = int(cta / uint32_t(im.tiles_per_row)), tile = int(cta - uint32_t(my) * uint32_t(im.tiles_per_row));
    const int mx0 = tile * im.mcus_per_tile;
    const int nm = min(im.mcus_per_tile, im.mcus_x - mx0); // MCUs in this tile
    const int hs = im.hs, vs = im.vs, ncomp = im.ncomp;
    const int mcu_w = 8 * hs, mcu_h = 8 * vs;
    const int tw = im.mcus_per_tile * mcu_w, th = mcu_h; // tile size in pixels (tw is a multiple of 64)
    const int W = im.width, H = im.height;
    uint8_t* s_cfull0 = s_y + 4096; // [th x tw] full-resolution Cb (<= 2048 ... 4096 bytes), Cr follows in s_coef's space
    uint8_t* s_cfull1 = reinterpret_cast<uint8_t*>(s_coef);

    { // tables
        const uint32_t* src = reinterpret_cast<const uint32_t*>(b.quant + im.qt);
        uint32_t* dst = reinterpret_cast<uint32_t*>(s_q);
        for (int i = threadIdx.x; i < int(2 * sizeof(EncQuantTable) / 4); i += kEncCta)
            dst[i] = src[i];
        for (int i = threadIdx.x; i < 512; i += kEncCta)
            s_aclen[i >> 8][i & 255] = uint8_t(b.huff->ac[i >> 8][i & 255] & 255u);
    }

    // ---- phase 1: pixels -> Y (and full-resolution Cb, Cr), 4 pixels per thread and step ----
    const int x0 = mx0 * mcu_w, y0 = my * mcu_h;
    const int used_w = nm * mcu_w; // columns of the tile that belong to MCUs of the image
    const int groups_x = used_w >> 2;
    const bool interleaved = im.in_fmt == 4; // B200JPEG_FMT_I_RGB
    const int64_t plane_stride = im.stride * int64_t(H);
    const bool aligned4 = ((reinterpret_cast<uintptr_t>(im.pixels) | uint64_t(im.stride)) & 3u) == 0;
    for (int g = threadIdx.x; g < groups_x * th; g += kEncCta) {
        const int ty = g / groups_x, tx = (g - ty * groups_x) << 2;
        const int sy = min(y0 + ty, H - 1);
        const int sx = x0 + tx;
        uint32_t yv = 0, cbv = 0, crv = 0;
        if (ncomp == 1) {
            const uint8_t* row = im.pixels + int64_t(sy) * im.stride;
            if (sx + 3 < W && aligned4)
                yv = *reinterpret_cast<const uint32_t*>(row + sx);
            else
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    yv |= uint32_t(row[min(sx + i, W - 1)]) << (8 * i);
        } else {
            uint32_t r[4], gg[4], bb[4];
            if (interleaved) {
                const uint8_t* row = im.pixels + int64_t(sy) * im.stride;
                if (sx + 3 < W && aligned4) {
                    const uint32_t* p = reinterpret_cast<const uint32_t*>(row + sx * 3);
                    const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
                    r[0] = w0 & 255u; gg[0] = (w0 >> 8) & 255u; bb[0] = (w0 >> 16) & 255u;
                    r[1] = w0 >> 24; gg[1] = w1 & 255u; bb[1] = (w1 >> 8) & 255u;
                    r[2] = (w1 >> 16) & 255u; gg[2] = w1 >> 24; bb[2] = w2 & 255u;
                    r[3] = (w2 >> 8) & 255u; gg[3] = (w2 >> 16) & 255u; bb[3] = w2 >> 24;
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const uint8_t* p = row + min(sx + i, W - 1) * 3;
                        r[i] = p[0]; gg[i] = p[1]; bb[i] = p[2];
                    }
                }
            } else {
                const uint8_t* row = im.pixels + int64_t(sy) * im.stride;
                if (sx + 3 < W && aligned4) {
                    const uint32_t wr = *reinterpret_cast<const uint32_t*>(row + sx);
                    const uint32_t wg = *reinterpret_cast<const uint32_t*>(row + plane_stride + sx);
                    const uint32_t wb = *reinterpret_cast<const uint32_t*>(row + 2 * plane_stride + sx);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        r[i] = (wr >> (8 * i)) & 255u; gg[i] = (wg >> (8 * i)) & 255u; bb[i] = (wb >> (8 * i)) & 255u;
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int x = min(sx + i, W - 1);
                        r[i] = row[x]; gg[i] = row[plane_stride + x]; bb[i] = row[2 * plane_stride + x];
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                yv |= enc_luma(r[i], gg[i], bb[i]) << (8 * i);
                cbv |= enc_cb(int(r[i]), int(gg[i]), int(bb[i])) << (8 * i);
                crv |= enc_cr(int(r[i]), int(gg[i]), int(bb[i])) << (8 * i);
            }
        }
        *reinterpret_cast<uint32_t*>(s_y + ty * tw + tx) = yv;
        if (ncomp == 3) {
            if (hs * vs == 1) {
                *reinterpret_cast<uint32_t*>(s_c[0] + ty * tw + tx) = cbv;
                *reinterpret_cast<uint32_t*>(s_c[1] + ty * tw + tx) = crv;
            } else {
                *reinterpret_cast<uint32_t*>(s_cfull0 + ty * tw + tx) = cbv;
                *reinterpret_cast<uint32_t*>(s_cfull1 + ty * tw + tx) = crv;
            }
        }
    }
    __syncthreads();

    // ---- phase 2: chroma box filter (4 outputs per thread and step) ----
    const int cw = tw / hs; // chroma pitch
    if (ncomp == 3 && hs * vs > 1) {
        // chroma rows past the last real row group repeat that group's result (expand_bottom_edge after downsampling)
        const int rows_ds = (H + vs - 1) / vs;
        const int cgroups_x = (used_w / hs) >> 2;
        for (int g = threadIdx.x; g < cgroups_x * 8 * 2; g += kEncCta) {
            const int c = g / (cgroups_x * 8), gi = g - c * (cgroups_x * 8);
            const int ry = gi / cgroups_x, cx = (gi - ry * cgroups_x) << 2;
            const int ryl = min(my * 8 + ry, rows_ds - 1) - my * 8; // source row group inside this tile
            const uint8_t* full = c ? s_cfull1 : s_cfull0;
            uint32_t out = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int x = cx + i;
                uint32_t v;
                if (hs == 2 && vs == 2) {
                    const uint8_t* p = full + (2 * ryl) * tw + 2 * x;
                    v = (uint32_t(p[0]) + p[1] + p[tw] + p[tw + 1] + ((x & 1) ? 2u : 1u)) >> 2;
                } else if (hs == 2) {
                    const uint8_t* p = full + ryl * tw + 2 * x;
                    v = (uint32_t(p[0]) + p[1] + uint32_t(x & 1)) >> 1;
                } else {
                    const uint8_t* p = full + (2 * ryl) * tw + x;
                    v = (uint32_t(p[0]) + p[tw] + 1u) >> 1;
                }
                out |= v << (8 * i);
            }
            *reinterpret_cast<uint32_t*>(s_c[c] + ry * cw + cx) = out;
        }
    }
    __syncthreads(); // s_coef (which held the full-resolution Cr) is free from here on

    // ---- phase 3: one block per thread ----
    const int t = threadIdx.x;
    const int nblocks = nm * im.bpm;
    const int m = t / im.bpm, j = t - m * im.bpm;
    int comp = 0, first_j = 0, count = 1;
    enc_block_comp(im, j, comp, first_j, count);
    const int mw = comp == 0 && ncomp == 3 ? hs : 1, mh = comp == 0 && ncomp == 3 ? vs : 1;
    const int jj = j - first_j, by = jj / mw, bx = jj - by * mw;
    const bool active = t < nblocks;
    const bool real = active && (mx0 + m) * mw + bx < im.wib[comp] && my * mh + by < im.hib[comp];
    uint32_t* my_coef = s_coef + t * kEncCoefPitch;
    uint32_t acbits = 0;
    if (real) {
        uint32_t rows[16];
        const uint8_t* src = comp == 0 ? s_y + (by * 8) * tw + (m * mw + bx) * 8 : s_c[comp - 1] + m * 8;
        const int pitch = comp == 0 ? tw : cw;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const uint2 v = *reinterpret_cast<const uint2*>(src + r * pitch);
            rows[2 * r] = v.x;
            rows[2 * r + 1] = v.y;
        }
        int16_t zz[64];
        enc_fdct_quant(rows, s_q[comp ? 1 : 0], zz);
        // entropy-coded length of the AC part (same walk as the emit kernel, lengths only)
        const uint8_t* aclen = s_aclen[comp ? 1 : 0];
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
            }
        }
        if (run)
            acbits += aclen[0];
#pragma unroll
        for (int k = 0; k < 32; ++k)
            my_coef[k] = uint32_t(uint16_t(zz[2 * k])) | (uint32_t(uint16_t(zz[2 * k + 1])) << 16);
    } else if (active) {
#pragma unroll
        for (int k = 0; k < 32; ++k)
            my_coef[k] = 0;
        acbits = s_aclen[comp ? 1 : 0][0]; // EOB only
    }
    __syncthreads();
    if (active && !real) {
        // dummy block (ljt/jccoefct.c:176-196): DC of the block before it in the MCU; a whole dummy row copies the block
        // before the row.  The source may itself be a dummy of the first kind; blocks at (0, 0) of a component are real.
        int sj = jj;
        for (;;) {
            const int sby = sj / mw, sbx = sj - sby * mw;
            const bool row_dummy = my * mh + sby >= im.hib[comp];
            const bool col_dummy = (mx0 + m) * mw + sbx >= im.wib[comp];
            if (!row_dummy && !col_dummy)
                break;
            sj = row_dummy ? sby * mw - 1 : sj - 1;
        }
        const uint32_t dcv = s_coef[(m * im.bpm + first_j + sj) * kEncCoefPitch] & 0xFFFFu;
        my_coef[0] = dcv;
    }
    __syncthreads();
    // ---- coalesced copy-out ----
    const uint64_t blk0 = im.blk_first + (uint64_t(my) * uint64_t(im.mcus_x) + uint64_t(mx0)) * uint64_t(im.bpm);
    uint32_t* gcoef = reinterpret_cast<uint32_t*>(b.coef + blk0 * 64);
    for (int i = threadIdx.x; i < nblocks * 32; i += kEncCta)
        gcoef[i] = s_coef[(i >> 5) * kEncCoefPitch + (i & 31)];
    if (active) {
        b.dc[blk0 + t] = int16_t(my_coef[0] & 0xFFFFu);
        b.acbits[blk0 + t] = uint16_t(acbits);
    }
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
    const int img = enc_find_image(b.cta_map_emit, b.num_images, cta_abs);
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
constexpr int kEncWindowWords = (kEncCta * kEncMaxBlockBytes) / 4 + 8;

struct CoefRegs
{
    uint32_t w[32];
    __device__ __forceinline__ int operator()(int k) const { return int(int16_t((k & 1) ? (w[k >> 1] >> 16) : (w[k >> 1] & 0xFFFFu))); }
};

__global__ void __launch_bounds__(kEncCta) k_enc_emit(EncBatchDev b)
{
    __shared__ uint32_t s_win[kEncWindowWords];
    __shared__ uint32_t s_ac[2][256];
    __shared__ uint32_t s_dc[2][16];
    const uint32_t cta_abs = blockIdx.x + b.cta_base_emit;
    const int img = enc_find_image(b.cta_map_emit, b.num_images, cta_abs);
    const EncImageDesc& im = b.images[img];
    const EncImageDyn& dy = b.dyn[img];
    const uint32_t cta = cta_abs - b.cta_map_emit[img];
    const uint32_t blk0 = cta * kEncCta, blk = blk0 + threadIdx.x;
    const uint32_t blk_end = min(blk0 + kEncCta, im.total_blocks);
    for (int i = threadIdx.x; i < 512; i += kEncCta)
        s_ac[i >> 8][i & 255] = b.huff->ac[i >> 8][i & 255];
    if (threadIdx.x < 32)
        s_dc[threadIdx.x >> 4][threadIdx.x & 15] = b.huff->dc[threadIdx.x >> 4][threadIdx.x & 15];
    // absolute bit position of a block inside the image's unstuffed stream
    auto bitpos = [&](uint32_t k) -> uint64_t { return enc_block_bitpos(b, im, img, k); };
    const uint64_t win_lo_bit = bitpos(blk0) & ~uint64_t(31);
    const uint64_t win_hi_bit = blk_end == im.total_blocks ? uint64_t(dy.stream_bytes) * 8u : bitpos(blk_end);
    const uint32_t win_words = uint32_t((win_hi_bit - win_lo_bit + 31u) >> 5);
    for (uint32_t i = threadIdx.x; i < win_words; i += kEncCta)
        s_win[i] = 0;
    __syncthreads();
    if (blk < blk_end) {
        CoefRegs c;
        const uint4* src = reinterpret_cast<const uint4*>(b.coef + (im.blk_first + blk) * 64);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint4 v = src[i];
            c.w[4 * i] = v.x; c.w[4 * i + 1] = v.y; c.w[4 * i + 2] = v.z; c.w[4 * i + 3] = v.w;
        }
        const int j = int(blk % uint32_t(im.bpm));
        int comp, first_j, count;
        enc_block_comp(im, j, comp, first_j, count);
        const int tsel = comp ? 1 : 0;
        EncWordSink sink;
        const uint64_t pos = bitpos(blk);
        sink.start(s_win, uint32_t(pos - win_lo_bit));
        enc_put_dc(enc_dc_diff(b, im, blk), s_dc[tsel], sink);
        enc_put_ac(c, s_ac[tsel], sink);
        // last block of a restart segment (or of the image): fill the byte with 1-bits (ljt/jchuff.c flush_bits:480)
        const bool seg_last = (blk + 1 == im.total_blocks) || ((blk + 1) % im.seg_blocks == 0);
        if (seg_last) {
            const int used = sink.n & 7; // the window starts on a 32-bit boundary of a byte-aligned stream
            if (used)
                sink.put((1u << (8 - used)) - 1u, 8 - used);
        }
        sink.finish();
    }
    __syncthreads();
    // window -> global stream (bytes in stream order): interior words plain, the two edge words may be shared with the
    // neighbouring CTAs' windows -> atomicOr on memory the zeroing kernel cleared
    uint32_t* gw = reinterpret_cast<uint32_t*>(b.stream + dy.stream_off) + (win_lo_bit >> 5);
    for (uint32_t i = threadIdx.x; i < win_words; i += kEncCta) {
        const uint32_t v = __byte_perm(s_win[i], 0, 0x0123);
        if (i == 0 || (i + 1 == win_words && (win_hi_bit & 31u)))
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
    const int img = enc_find_image(b.cta_map_emit, b.num_images, c);
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
    const int img = enc_find_image(b.cta_map_stuff, b.num_images, blockIdx.x);
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
    const int img = enc_find_image(b.cta_map_stuff, b.num_images, blockIdx.x);
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
    k_enc_dct<<<d.dct_ctas, kEncCta, 0, st>>>(b);
    return 1;
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
