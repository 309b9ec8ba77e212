// enc_plan.cpp -- see enc_plan.h
#include "enc_plan.h"
#include <cstring>
#include "../../include/b200jpeg.h"
#include "enc_math.cuh"

namespace b200jpeg {

namespace {
// ITU-T T.81 Annex K.1 quantisation tables (natural order)
const uint8_t kBaseQuant[2][64] = {
    {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57, 69, 56, 14, 17, 22, 29, 51, 87, 80, 62,
        18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99},
    {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99, 99, 99, 47, 66, 99, 99, 99, 99, 99, 99,
        99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99}};

// ITU-T T.81 Annex K.3 Huffman tables: 16 length counts, then the symbols in code order.  Index: 0 DC luma, 1 DC chroma,
// 2 AC luma, 3 AC chroma.
struct StdHuff
{
    uint8_t counts[16];
    std::vector<uint8_t> symbols;
};

std::vector<uint8_t> ac_symbols(const uint8_t* head, int nhead)
{
    // After the irregular head both AC tables list the remaining run/size symbols in increasing numeric order.
    std::vector<uint8_t> v(head, head + nhead);
    bool seen[256] = {};
    for (uint8_t s : v)
        seen[s] = true;
    for (int s = 0; s < 256; ++s) {
        const int size = s & 15, run = s >> 4;
        const bool valid = (size >= 1 && size <= 10) || s == 0x00 || s == 0xF0;
        (void)run;
        if (valid && !seen[s])
            v.push_back(uint8_t(s));
    }
    return v;
}

const StdHuff& std_huff(int which)
{
    static StdHuff t[4];
    static bool init = false;
    if (!init) {
        const uint8_t c0[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
        const uint8_t c1[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
        const uint8_t c2[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
        const uint8_t c3[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
        std::memcpy(t[0].counts, c0, 16);
        std::memcpy(t[1].counts, c1, 16);
        std::memcpy(t[2].counts, c2, 16);
        std::memcpy(t[3].counts, c3, 16);
        for (int i = 0; i < 12; ++i) {
            t[0].symbols.push_back(uint8_t(i));
            t[1].symbols.push_back(uint8_t(i));
        }
        // the symbols with codes shorter than 16 bits (K.3.3.2): 37 for luminance, 43 for chrominance
        const uint8_t h2[] = {0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71, 0x14,
            0x32, 0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72, 0x82};
        const uint8_t h3[] = {0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22, 0x32,
            0x81, 0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16, 0x24, 0x34, 0xe1,
            0x25, 0xf1};
        t[2].symbols = ac_symbols(h2, int(sizeof(h2)));
        t[3].symbols = ac_symbols(h3, int(sizeof(h3)));
        init = true;
    }
    return t[which];
}

void put16(std::vector<uint8_t>& o, int v)
{
    o.push_back(uint8_t(v >> 8));
    o.push_back(uint8_t(v));
}
} // namespace

void enc_build_quant(int quality, int which, EncQuantTable& out)
{
    const int q = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
    const long scale = q < 50 ? 5000 / q : 200 - 2 * q;
    for (int i = 0; i < 64; ++i) {
        long v = (long(kBaseQuant[which][i]) * scale + 50) / 100;
        v = v < 1 ? 1 : (v > 255 ? 255 : v);
        out.qval[i] = uint8_t(v);
        // divisor = 8 * quant value (the forward DCT leaves the results scaled by 8)
        const uint32_t d = uint32_t(v) << 3;
        int msb = 31;
        while (!(d >> msb))
            --msb;
        int r = 32 + msb;
        uint64_t fq = (uint64_t(1) << r) / d;
        const uint64_t fr = (uint64_t(1) << r) % d;
        uint32_t c = d / 2;
        if (fr == 0) { // power of two: the reciprocal would not fit, halve it
            fq >>= 1;
            --r;
        } else if (fr <= d / 2) {
            ++c;
        } else {
            ++fq;
        }
        out.recip[i] = uint32_t(fq);
        out.corr[i] = uint16_t(c);
        out.shift[i] = uint8_t(r - 32);
    }
}

void enc_build_huff(EncHuffTables& out)
{
    std::memset(&out, 0, sizeof(out));
    for (int which = 0; which < 4; ++which) {
        const StdHuff& h = std_huff(which);
        uint32_t code = 0;
        size_t k = 0;
        for (int len = 1; len <= 16; ++len) {
            for (int i = 0; i < h.counts[len - 1]; ++i, ++k) {
                const uint8_t sym = h.symbols[k];
                const uint32_t e = (code++ << 8) | uint32_t(len);
                if (which < 2)
                    out.dc[which][sym & 15] = e;
                else
                    out.ac[which - 2][sym] = e;
            }
            code <<= 1;
        }
    }
}

bool enc_plan_image(int width, int height, int format, int subsampling, int restart_interval, EncImageDesc& im)
{
    std::memset(&im, 0, sizeof(im));
    if (width < 1 || height < 1 || width > 65535 || height > 65535 || restart_interval < 0 || restart_interval > 65535)
        return false;
    int hs = 1, vs = 1, ncomp = 3;
    switch (subsampling) {
    case B200JPEG_CSS_444: break;
    case B200JPEG_CSS_422: hs = 2; break;
    case B200JPEG_CSS_420: hs = 2; vs = 2; break;
    case B200JPEG_CSS_440: vs = 2; break;
    case B200JPEG_CSS_GRAY: ncomp = 1; break;
    default: return false;
    }
    if (format == B200JPEG_FMT_P_Y) {
        if (ncomp != 1)
            return false;
    } else if (format == B200JPEG_FMT_I_RGB || format == B200JPEG_FMT_P_RGB) {
        if (ncomp != 3)
            return false; // RGB -> gray conversion is not offered
    } else {
        return false;
    }
    im.width = width;
    im.height = height;
    im.ncomp = ncomp;
    im.hs = hs;
    im.vs = vs;
    im.in_fmt = format;
    im.mcus_x = (width + 8 * hs - 1) / (8 * hs);
    im.mcus_y = (height + 8 * vs - 1) / (8 * vs);
    im.bpm = ncomp == 1 ? 1 : hs * vs + 2;
    for (int c = 0; c < ncomp; ++c) {
        const int ch = c ? 1 : hs, cv = c ? 1 : vs;
        im.wib[c] = (width * ch + hs * 8 - 1) / (hs * 8);
        im.hib[c] = (height * cv + vs * 8 - 1) / (vs * 8);
    }
    im.restart_interval = restart_interval;
    im.total_blocks = uint32_t(im.mcus_x) * uint32_t(im.mcus_y) * uint32_t(im.bpm);
    const uint32_t mcus = uint32_t(im.mcus_x) * uint32_t(im.mcus_y);
    if (restart_interval) {
        im.seg_blocks = uint32_t(restart_interval) * uint32_t(im.bpm);
        im.num_segments = (mcus + uint32_t(restart_interval) - 1) / uint32_t(restart_interval);
    } else {
        im.seg_blocks = im.total_blocks;
        im.num_segments = 1;
    }
    // component planes between the colour-conversion and the FDCT kernels: whole MCUs, rows of a multiple of 16 bytes
    {
        const int pc = (im.mcus_x * 8 + 15) / 16 * 16;
        im.plane_pitch[0] = (ncomp == 3 && hs == 2) ? 2 * pc : pc;
        im.plane_pitch[1] = im.plane_pitch[2] = ncomp == 3 ? pc : 0;
        im.plane_off[0] = im.plane_off[1] = im.plane_off[2] = 0; // assigned by the encoder when it lays out the batch
    }
    // (CPU emulation of the tests) a run of MCUs of one MCU row handled together: as many as give at most kEncCta blocks, at most 1024 pixels wide
    int mpt = kEncCta / im.bpm;
    if (ncomp == 3 && hs * vs == 4)
        mpt = 16;
    else if (ncomp == 3)
        mpt = 32;
    im.mcus_per_tile = mpt;
    im.tiles_per_row = (im.mcus_x + mpt - 1) / mpt;
    return true;
}

void enc_write_headers(const EncImageDesc& im, const EncQuantTable& luma, const EncQuantTable& chroma, std::vector<uint8_t>& o)
{
    o.clear();
    put16(o, 0xFFD8);
    put16(o, 0xFFE0); // JFIF 1.01, no density, no thumbnail
    put16(o, 16);
    const uint8_t jfif[14] = {'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};
    o.insert(o.end(), jfif, jfif + 14);
    const int nsets = im.ncomp == 3 ? 2 : 1;
    for (int t = 0; t < nsets; ++t) {
        const EncQuantTable& q = t ? chroma : luma;
        put16(o, 0xFFDB);
        put16(o, 67);
        o.push_back(uint8_t(t));
        for (int k = 0; k < 64; ++k)
            o.push_back(q.qval[enc_zigzag(k)]);
    }
    put16(o, 0xFFC0);
    put16(o, 8 + 3 * im.ncomp);
    o.push_back(8);
    put16(o, im.height);
    put16(o, im.width);
    o.push_back(uint8_t(im.ncomp));
    for (int c = 0; c < im.ncomp; ++c) {
        o.push_back(uint8_t(c + 1));
        o.push_back(uint8_t(c ? 0x11 : (im.hs << 4 | im.vs)));
        o.push_back(uint8_t(c ? 1 : 0));
    }
    for (int t = 0; t < nsets; ++t)
        for (int ac = 0; ac < 2; ++ac) {
            const StdHuff& h = std_huff(t + 2 * ac);
            put16(o, 0xFFC4);
            put16(o, int(2 + 1 + 16 + h.symbols.size()));
            o.push_back(uint8_t(t | (ac << 4)));
            o.insert(o.end(), h.counts, h.counts + 16);
            o.insert(o.end(), h.symbols.begin(), h.symbols.end());
        }
    if (im.restart_interval) {
        put16(o, 0xFFDD);
        put16(o, 4);
        put16(o, im.restart_interval);
    }
    put16(o, 0xFFDA);
    put16(o, 6 + 2 * im.ncomp);
    o.push_back(uint8_t(im.ncomp));
    for (int c = 0; c < im.ncomp; ++c) {
        o.push_back(uint8_t(c + 1));
        o.push_back(uint8_t(c ? 0x11 : 0x00));
    }
    o.push_back(0);
    o.push_back(63);
    o.push_back(0);
}

} // namespace b200jpeg
