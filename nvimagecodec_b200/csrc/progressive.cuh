// progressive.cuh -- decode of the scans of a progressive JPEG (SOF2) into dense quantised coefficients, as host + device
// inline code (the CUDA kernel k_prog_decode wraps it; compiled with g++ the same body runs in the CPU emulation of the tests).
//
// Reference arithmetic restated here ("ljt/" = external/libjpeg-turbo in the reference tree):
//   ljt/jdphuff.c  decode_mcu_DC_first:289-356  decode_mcu_AC_first:364-441  decode_mcu_DC_refine:451-498
//                  decode_mcu_AC_refine:506-655  process_restart:236-273  start_pass_phuff_decoder:96-211
//   ljt/jdhuff.c   jpeg_fill_bit_buffer:287-391 (FF00 stuffing, stop at a marker)  jpeg_huff_decode:398-450
//   ljt/jdinput.c  per_scan_setup:156-250 (MCU geometry of interleaved and single-component scans)
//   ljt/jdcoefct.c consume_data:193-262 (where the blocks of a scan live in the coefficient arrays)
// A progressive stream is a sequence of scans that each refine the coefficient arrays: a scan is walked serially (the
// refinement scans depend, symbol by symbol, on which coefficients are already non-zero), one thread per image; the images of
// a batch run side by side.  The coefficients land in the batch's dense coefficient buffer (MCU scan order, natural order
// inside a block, DC in slot 0) and the reconstruction kernel takes them from there instead of decoding AC symbols.
#pragma once
#include "device_types.h"
#if !defined(__CUDACC__)
#include "host_shim.h"
#endif

namespace b200jpeg {

// Bit reader over the stuffed bytes of one scan (jpeg_fill_bit_buffer): FF00 is a data byte FF, FF FF.. is fill, any other
// FF xx is a marker -- the data ends there and zeros are supplied (the caller treats running into a marker as an error,
// except for the RSTn it expects).
struct ProgBits
{
    const uint8_t* d;
    uint32_t pos, end; // next byte, end of the scan's bytes
    uint64_t buf;      // bits, left aligned; zeros behind the valid ones
    int32_t n;         // valid (real) bits in buf
    bool at_marker;    // pos sits on a marker (or the end): no more data
    bool starved;      // bits were consumed that the stream did not have

    B2J_HD void init(const uint8_t* data, uint32_t from, uint32_t to)
    {
        d = data;
        pos = from;
        end = to;
        buf = 0;
        n = 0;
        at_marker = false;
        starved = false;
    }
    B2J_HD void fill()
    {
        // four bytes at once when none of them is an FF (255 of 256 are not): the loads do not depend on each other, and
        // no byte has to be looked at twice; anything with an FF in it goes through the byte loop below
        while (n <= 32 && !at_marker && pos + 4u <= end) {
            const uint32_t c0 = d[pos], c1 = d[pos + 1], c2 = d[pos + 2], c3 = d[pos + 3];
            if (c0 == 0xFF || c1 == 0xFF || c2 == 0xFF || c3 == 0xFF)
                break;
            buf |= uint64_t((c0 << 24) | (c1 << 16) | (c2 << 8) | c3) << (32 - n);
            n += 32;
            pos += 4;
        }
        while (n <= 56 && !at_marker) {
            if (pos >= end) {
                at_marker = true;
                break;
            }
            uint32_t c = d[pos];
            if (c == 0xFF) {
                uint32_t q = pos + 1;
                while (q < end && d[q] == 0xFF)
                    ++q; // fill bytes
                if (q < end && d[q] == 0x00) {
                    pos = q + 1; // stuffed FF
                } else {
                    at_marker = true; // stay in front of the marker
                    break;
                }
            } else {
                ++pos;
            }
            buf |= uint64_t(c) << (56 - n);
            n += 8;
        }
    }
    B2J_HD uint32_t peek16()
    {
        if (n < 16)
            fill();
        return uint32_t(buf >> 48);
    }
    B2J_HD void skip(int k)
    {
        buf <<= k;
        n -= k;
        if (n < 0) {
            starved = true;
            n = 0;
        }
    }
    B2J_HD uint32_t get(int k) // k in [0, 16]
    {
        if (k == 0)
            return 0;
        if (n < k)
            fill();
        const uint32_t v = uint32_t(buf >> (64 - k));
        skip(k);
        return v;
    }
    // process_restart (ljt/jdphuff.c:236-273): drop what is left of the current byte, expect the marker RSTn next
    B2J_HD bool restart(int expected)
    {
        if (n >= 8)
            return false; // whole data bytes in front of the marker: the segment is longer than its MCUs
        buf = 0;
        n = 0;
        at_marker = false;
        while (pos + 1 < end && d[pos] == 0xFF && d[pos + 1] == 0xFF)
            ++pos;
        if (pos + 1 < end && d[pos] == 0xFF && d[pos + 1] == uint32_t(0xD0 + (expected & 7))) {
            pos += 2;
            return true;
        }
        return false;
    }
};

// Look-ahead table of a Huffman table for codes of up to kProgFastBits bits: entry = length << 8 | symbol, 0 = longer code.
// The baseline kernels' LUTs (DevHuffTable::lut) do not serve here: they fold the run length of the end-of-band symbols away,
// which a progressive scan needs (EOBn, ljt/jdphuff.c:395-404).  Entry i is independent of the others, so a warp fills the
// table of the scan at hand side by side (k_prog_decode); the CPU emulation fills it in a loop.
constexpr int kProgFastBits = 8;
B2J_HD uint16_t prog_fast_entry(const DevHuffTable* t, uint32_t prefix /* kProgFastBits bits */)
{
    for (int l = 1; l <= kProgFastBits; ++l) {
        const int code = int(prefix >> (kProgFastBits - l));
        if (code <= t->maxcode[l])
            return uint16_t((l << 8) | t->vals[(code + t->valoffset[l]) & 255]);
    }
    return 0;
}

// One Huffman symbol (jpeg_huff_decode, ljt/jdhuff.c:398-450): look-ahead table first (fast may be null), then the
// canonical walk over the longer lengths.  -1 for a code that is not in the table.
B2J_HD int prog_huff_symbol(ProgBits& br, const DevHuffTable* t, const uint16_t* fast = nullptr)
{
    const uint32_t w = br.peek16();
    int l = 1;
    if (fast) {
        const uint32_t e = fast[w >> (16 - kProgFastBits)];
        if (e) {
            br.skip(int(e >> 8));
            return int(e & 255u);
        }
        l = kProgFastBits + 1;
    }
    for (; l <= 16; ++l) {
        const int code = int(w >> (16 - l));
        if (code <= t->maxcode[l]) {
            br.skip(l);
            return t->vals[(code + t->valoffset[l]) & 255];
        }
    }
    return -1;
}

B2J_HD int prog_extend(uint32_t r, int s) { return r < (1u << (s - 1)) ? int(r) - (1 << s) + 1 : int(r); } // HUFF_EXTEND

// Index (in the image's coefficient array, blocks in MCU scan order) of block (bx, by) of component c.
B2J_HD uint32_t prog_block_index(const ImageDesc& im, int c, int bx, int by)
{
    const int hs = im.hs[c], vs = im.vs[c], lh = hs >> 1, lv = vs >> 1; // sampling factors are 1, 2 or 4: log2 = f >> 1
    const uint32_t mcu = uint32_t(by >> lv) * uint32_t(im.mcus_x) + uint32_t(bx >> lh);
    return mcu * uint32_t(im.bpm) + uint32_t(im.blk_first[c] + (by & (vs - 1)) * hs + (bx & (hs - 1)));
}

// The AC part of one block in a refinement scan (decode_mcu_AC_refine).  zz = zigzag position -> natural index.
B2J_HD bool prog_ac_refine_block(ProgBits& br, const DevHuffTable* tab, int16_t* blk, int ss, int se, int al, uint32_t& eobrun,
    const uint8_t* zz, const uint16_t* fast = nullptr)
{
    const int p1 = 1 << al, m1 = -(1 << al);
    int k = ss;
    if (eobrun == 0) {
        for (; k <= se; ++k) {
            int s = prog_huff_symbol(br, tab, fast);
            if (s < 0)
                return false;
            int r = s >> 4;
            s &= 15;
            if (s) { // a newly non-zero coefficient: its size is always 1, the next bit is its sign
                s = br.get(1) ? p1 : m1;
            } else if (r != 15) {
                eobrun = 1u << r; // EOBr: the run covers this block as well
                if (r)
                    eobrun += br.get(r);
                break;
            }
            // go over the coefficients that are already non-zero (each takes a correction bit) and r that are still zero
            do {
                int16_t* c = blk + zz[k];
                if (*c != 0) {
                    if (br.get(1) && (*c & p1) == 0)
                        *c = int16_t(*c + (*c >= 0 ? p1 : m1));
                } else if (--r < 0) {
                    break;
                }
                ++k;
            } while (k <= se);
            if (s && k <= se)
                blk[zz[k]] = int16_t(s);
        }
    }
    if (eobrun > 0) { // what is left of the band: correction bits only
        for (; k <= se; ++k) {
            int16_t* c = blk + zz[k];
            if (*c != 0 && br.get(1) && (*c & p1) == 0)
                *c = int16_t(*c + (*c >= 0 ? p1 : m1));
        }
        --eobrun;
    }
    return true;
}

// The AC part of one block in a first scan of its band (decode_mcu_AC_first).
B2J_HD bool prog_ac_first_block(ProgBits& br, const DevHuffTable* tab, int16_t* blk, int ss, int se, int al, uint32_t& eobrun, const uint8_t* zz,
    const uint16_t* fast = nullptr)
{
    if (eobrun > 0) {
        --eobrun;
        return true;
    }
    for (int k = ss; k <= se; ++k) {
        int s = prog_huff_symbol(br, tab, fast);
        if (s < 0)
            return false;
        const int r = s >> 4;
        s &= 15;
        if (s) {
            k += r;
            const int v = prog_extend(br.get(s), s);
            if (k > 63)
                return false;
            blk[zz[k]] = int16_t(v * (1 << al));
        } else if (r == 15) {
            k += 15;
        } else {
            eobrun = 1u << r;
            if (r)
                eobrun += br.get(r);
            --eobrun; // this block is the first of the run
            break;
        }
    }
    return true;
}

// Decode one scan of image im into its coefficients (coef = the image's first block).  raw = the image's bytes in the raw
// buffer (ScanDesc::off is relative to it).  Returns 0 or kErrProgressive.
// fast: optional look-ahead tables (prog_fast_entry), 256 entries per table of the scan: its AC table for an AC scan, else
// one per component of the scan.
B2J_HD uint32_t prog_decode_scan(const ImageDesc& im, const ScanDesc& sc, const uint8_t* raw, const DevHuffTable* pool, int16_t* coef,
    const uint8_t* zz, const uint16_t* fast = nullptr)
{
    ProgBits br;
    br.init(raw, sc.off, sc.off + sc.len);
    const bool dc = sc.ss == 0;
    int last_dc[3] = {0, 0, 0};
    uint32_t eobrun = 0;
    uint32_t to_go = sc.restart_interval ? uint32_t(sc.restart_interval) : 0xFFFFFFFFu;
    int next_rst = 0;
    // MCU geometry (per_scan_setup): one component -> its own blocks in raster order, the blocks that cover the image only;
    // several -> the frame's MCUs with the scan's components' blocks
    const bool single = sc.ncomp == 1;
    const int c0 = sc.comp[0];
    const int mcus_x = single ? (im.ds_w[c0] + 7) / 8 : im.mcus_x;
    const int mcus_y = single ? (im.ds_h[c0] + 7) / 8 : im.mcus_y;
    for (int my = 0; my < mcus_y; ++my) {
        for (int mx = 0; mx < mcus_x; ++mx) {
            if (to_go == 0) { // process_restart
                if (!br.restart(next_rst))
                    return kErrProgressive;
                next_rst = (next_rst + 1) & 7;
                last_dc[0] = last_dc[1] = last_dc[2] = 0;
                eobrun = 0;
                to_go = uint32_t(sc.restart_interval);
            }
            for (int i = 0; i < sc.ncomp; ++i) {
                const int c = sc.comp[i];
                const int nh = single ? 1 : im.hs[c], nv = single ? 1 : im.vs[c];
                for (int v = 0; v < nv; ++v)
                    for (int h = 0; h < nh; ++h) {
                        const int bx = single ? mx : mx * nh + h, by = single ? my : my * nv + v;
                        int16_t* blk = coef + uint64_t(prog_block_index(im, c, bx, by)) * 64;
                        if (dc) {
                            if (sc.ah == 0) { // decode_mcu_DC_first
                                int s = prog_huff_symbol(br, pool + sc.dc_tab[i], fast ? fast + (i << kProgFastBits) : nullptr);
                                if (s < 0 || s > 15)
                                    return kErrProgressive;
                                if (s)
                                    s = prog_extend(br.get(s), s);
                                last_dc[i] += s;
                                blk[0] = int16_t(last_dc[i] * (1 << sc.al));
                            } else if (br.get(1)) { // decode_mcu_DC_refine
                                blk[0] = int16_t(blk[0] | (1 << sc.al));
                            }
                        } else {
                            const bool ok = sc.ah == 0 ? prog_ac_first_block(br, pool + sc.ac_tab, blk, sc.ss, sc.se, sc.al, eobrun, zz, fast)
                                                       : prog_ac_refine_block(br, pool + sc.ac_tab, blk, sc.ss, sc.se, sc.al, eobrun, zz, fast);
                            if (!ok)
                                return kErrProgressive;
                        }
                    }
            }
            --to_go;
            if (br.starved)
                return kErrProgressive; // the scan ended before its blocks did
        }
    }
    return 0;
}

} // namespace b200jpeg
