// tests/emu/emu_encode.cpp -- TEST INFRASTRUCTURE.
// CPU emulation of the b200jpeg encode pipeline: the same per-thread bodies (nvimagecodec_b200/csrc/enc_math.cuh), the same
// planner / header writer (enc_plan.cpp) and the same data layout (tiles, zigzag coefficient array, per-block bit
// offsets, per-CTA stream windows with OR-ed words, stuffing chunks) as encode_kernels.cu, executed by plain loops over
// "CTAs" and "threads".  Lets the "not gpu" tests check the encoder logic byte-for-byte against the oracle.
#include <algorithm>
#include <cstring>
#include <vector>
#include "../../nvimagecodec_b200/csrc/enc_math.cuh"
#include "../../nvimagecodec_b200/csrc/enc_plan.h"
#include "../../include/b200jpeg.h"

using namespace b200jpeg;

namespace {
struct Coef64
{
    const int16_t* p;
    int operator()(int k) const { return p[k]; }
};

// sample of component `comp` at (x, y) of its (downsampled) plane, with the reference's edge rules
uint32_t sample_at(const EncImageDesc& im, int comp, int x, int y)
{
    const int W = im.width, H = im.height;
    auto px = [&](int sx, int sy, int c) -> int {
        sx = std::min(sx, W - 1);
        sy = std::min(sy, H - 1);
        if (im.in_fmt == B200JPEG_FMT_I_RGB)
            return im.pixels[int64_t(sy) * im.stride + sx * 3 + c];
        return im.pixels[int64_t(c) * im.stride * H + int64_t(sy) * im.stride + sx];
    };
    auto conv = [&](int sx, int sy) -> uint32_t {
        if (im.ncomp == 1)
            return uint32_t(px(sx, sy, 0));
        const int r = px(sx, sy, 0), g = px(sx, sy, 1), b = px(sx, sy, 2);
        return comp == 0 ? enc_luma(r, g, b) : (comp == 1 ? enc_cb(r, g, b) : enc_cr(r, g, b));
    };
    if (comp == 0 || im.hs * im.vs == 1)
        return conv(x, y);
    const int rows_ds = (H + im.vs - 1) / im.vs;
    const int yd = std::min(y, rows_ds - 1);
    if (im.hs == 2 && im.vs == 2)
        return (conv(2 * x, 2 * yd) + conv(2 * x + 1, 2 * yd) + conv(2 * x, 2 * yd + 1) + conv(2 * x + 1, 2 * yd + 1) + ((x & 1) ? 2u : 1u)) >> 2;
    if (im.hs == 2)
        return (conv(2 * x, yd) + conv(2 * x + 1, yd) + uint32_t(x & 1)) >> 1;
    return (conv(x, 2 * yd) + conv(x, 2 * yd + 1) + 1u) >> 1;
}
} // namespace

extern "C" long emu_encode(const uint8_t* pixels, int width, int height, int format, int64_t stride, int subsampling, int quality,
    int restart_interval, uint8_t* out, size_t cap)
{
    EncImageDesc im;
    if (!enc_plan_image(width, height, format, subsampling, restart_interval, im))
        return -3;
    im.pixels = pixels;
    im.stride = stride;
    EncQuantTable qt[2];
    enc_build_quant(quality, 0, qt[0]);
    enc_build_quant(quality, 1, qt[1]);
    EncHuffTables huff;
    enc_build_huff(huff);
    std::vector<uint8_t> file;
    enc_write_headers(im, qt[0], qt[1], file);

    // ---- E1: per tile, per block: samples -> FDCT -> quantise; dummy blocks; AC bit lengths ----
    const uint32_t nb = im.total_blocks;
    std::vector<int16_t> coef(size_t(nb) * 64, 0x5A5A), dc(nb);
    std::vector<uint16_t> acbits(nb);
    for (int my = 0; my < im.mcus_y; ++my)
        for (int tile = 0; tile < im.tiles_per_row; ++tile) {
            const int mx0 = tile * im.mcus_per_tile, nm = std::min(im.mcus_per_tile, im.mcus_x - mx0);
            const uint64_t blk0 = (uint64_t(my) * im.mcus_x + mx0) * im.bpm;
            for (int pass = 0; pass < 2; ++pass) // pass 0: real blocks, pass 1: dummy blocks (after the "__syncthreads")
                for (int t = 0; t < nm * im.bpm; ++t) {
                    const int m = t / im.bpm, j = t - m * im.bpm;
                    int comp, first_j, count;
                    enc_block_comp(im, j, comp, first_j, count);
                    const int mw = comp == 0 && im.ncomp == 3 ? im.hs : 1, mh = comp == 0 && im.ncomp == 3 ? im.vs : 1;
                    const int jj = j - first_j, by = jj / mw, bx = jj - by * mw;
                    const bool real = (mx0 + m) * mw + bx < im.wib[comp] && my * mh + by < im.hib[comp];
                    int16_t* zz = coef.data() + (blk0 + t) * 64;
                    if (pass == 0 && real) {
                        uint32_t rows[16];
                        for (int r = 0; r < 8; ++r) {
                            uint32_t lo = 0, hi = 0;
                            for (int c = 0; c < 4; ++c) {
                                lo |= sample_at(im, comp, ((mx0 + m) * mw + bx) * 8 + c, (my * mh + by) * 8 + r) << (8 * c);
                                hi |= sample_at(im, comp, ((mx0 + m) * mw + bx) * 8 + 4 + c, (my * mh + by) * 8 + r) << (8 * c);
                            }
                            rows[2 * r] = lo;
                            rows[2 * r + 1] = hi;
                        }
                        enc_fdct_quant(rows, qt[comp ? 1 : 0], zz);
                        EncBitCounter cnt;
                        enc_put_ac(Coef64{zz}, huff.ac[comp ? 1 : 0], cnt);
                        acbits[blk0 + t] = uint16_t(cnt.bits);
                    } else if (pass == 0) {
                        std::memset(zz, 0, 128);
                        acbits[blk0 + t] = uint16_t(huff.ac[comp ? 1 : 0][0] & 255u);
                    } else if (!real) {
                        int sj = jj;
                        for (;;) {
                            const int sby = sj / mw, sbx = sj - sby * mw;
                            const bool row_dummy = my * mh + sby >= im.hib[comp], col_dummy = (mx0 + m) * mw + sbx >= im.wib[comp];
                            if (!row_dummy && !col_dummy)
                                break;
                            sj = row_dummy ? sby * mw - 1 : sj - 1;
                        }
                        zz[0] = coef[(blk0 + m * im.bpm + first_j + sj) * 64];
                    }
                }
            for (int t = 0; t < nm * im.bpm; ++t)
                dc[blk0 + t] = coef[(blk0 + t) * 64];
        }

    // ---- E2/E3: bit offsets per block inside its segment, segment byte offsets ----
    auto dc_diff = [&](uint32_t blk) {
        const int64_t p = enc_dc_pred_block(im, blk);
        return p < 0 ? int(dc[blk]) : int(dc[blk]) - int(dc[size_t(p)]);
    };
    std::vector<uint32_t> blk_off(nb), seg_off(im.num_segments);
    uint32_t stream_bytes = 0;
    for (uint32_t seg = 0; seg < im.num_segments; ++seg) {
        const uint32_t first = seg * im.seg_blocks, last = std::min(first + im.seg_blocks, nb);
        uint32_t bits = 0;
        for (uint32_t blk = first; blk < last; ++blk) {
            int comp, first_j, count;
            enc_block_comp(im, int(blk % uint32_t(im.bpm)), comp, first_j, count);
            const int n = enc_category(dc_diff(blk));
            blk_off[blk] = bits;
            bits += acbits[blk] + (huff.dc[comp ? 1 : 0][n] & 255u) + uint32_t(n);
        }
        seg_off[seg] = stream_bytes;
        stream_bytes += (bits + 7) >> 3;
    }

    // ---- E4: emit, one window of OR-ed words per CTA of kEncCta blocks ----
    std::vector<uint32_t> stream_words((stream_bytes + 3) / 4 + 2, 0);
    auto bitpos = [&](uint32_t k) { return uint64_t(seg_off[k / im.seg_blocks]) * 8u + blk_off[k]; };
    for (uint32_t blk0 = 0; blk0 < nb; blk0 += kEncCta) {
        const uint32_t blk_end = std::min(blk0 + uint32_t(kEncCta), nb);
        const uint64_t lo = bitpos(blk0) & ~uint64_t(31), hi = blk_end == nb ? uint64_t(stream_bytes) * 8u : bitpos(blk_end);
        std::vector<uint32_t> win((hi - lo + 31) >> 5, 0);
        if (win.size() > size_t(kEncCta * kEncMaxBlockBytes / 4 + 8))
            return -4;
        for (uint32_t blk = blk0; blk < blk_end; ++blk) {
            int comp, first_j, count;
            enc_block_comp(im, int(blk % uint32_t(im.bpm)), comp, first_j, count);
            EncWordSink sink;
            sink.start(win.data(), uint32_t(bitpos(blk) - lo));
            enc_put_dc(dc_diff(blk), huff.dc[comp ? 1 : 0], sink);
            {
                const Coef64 get{coef.data() + size_t(blk) * 64};
                uint64_t mask = 0;
                for (int k = 1; k < 64; ++k)
                    if (get(k))
                        mask |= uint64_t(1) << k;
                if (blk & 1) // both statements of the AC walk: the one that tests every position and the mask-driven one
                    enc_put_ac(get, huff.ac[comp ? 1 : 0], sink);
                else
                    enc_put_ac_sparse(mask, get, huff.ac[comp ? 1 : 0], sink);
            }
            if (blk + 1 == nb || (blk + 1) % im.seg_blocks == 0) {
                const int used = sink.n & 7;
                if (used)
                    sink.put((1u << (8 - used)) - 1u, 8 - used);
            }
            sink.finish();
            if (sink.widx > win.size())
                return -5;
        }
        for (size_t i = 0; i < win.size(); ++i)
            stream_words[(lo >> 5) + i] |= __builtin_bswap32(win[i]);
    }
    const uint8_t* stream = reinterpret_cast<const uint8_t*>(stream_words.data());

    // ---- E5: stuffing + restart markers ----
    uint32_t seg_next = 1;
    for (uint32_t i = 0; i < stream_bytes; ++i) {
        if (seg_next < im.num_segments && seg_off[seg_next] == i) {
            file.push_back(0xFF);
            file.push_back(uint8_t(0xD0 + ((seg_next - 1) & 7u)));
            ++seg_next;
        }
        file.push_back(stream[i]);
        if (stream[i] == 0xFF)
            file.push_back(0);
    }
    file.push_back(0xFF);
    file.push_back(0xD9);
    if (file.size() > cap)
        return -2;
    std::memcpy(out, file.data(), file.size());
    return long(file.size());
}
