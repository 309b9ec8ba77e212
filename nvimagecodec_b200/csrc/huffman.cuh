// huffman.cuh -- the per-thread Huffman decode loop of b200jpeg (host + device inline, so the CPU emulation used by the
// "not gpu" tests runs the very same code).
//
// Reference arithmetic restated here ("ljt/" = external/libjpeg-turbo in the reference tree):
//   ljt/jdhuff.c decode_mcu_slow:552-660, HUFF_EXTEND:483-488, jpeg_huff_decode:398-450, process_restart:512-542
//   zigzag: jpeg_natural_order, ljt/jutils.c:59
// The reference walks the stream serially.  Here the unstuffed stream of an image is cut into subsequences of
// `subseq_bits` bits; thread k decodes every symbol that *starts* inside subsequence k.  A thread that does not know the
// decoder state at its first bit starts from a guess and the guesses are repaired by the synchronisation rounds
// (decode_kernels.cu); Huffman streams self-synchronise, so the repair converges after a few rounds.
#pragma once
#include "device_types.h"

namespace b200jpeg {

#if defined(__CUDA_ARCH__)
static __constant__ uint8_t c_zigzag[80] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,
    7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55,
    62, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};
#define B2J_ZZ(i) c_zigzag[i]
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

B2J_HD uint32_t funnel_left(uint32_t hi, uint32_t lo, uint32_t sh) // top 32 bits of (hi:lo) << sh, sh in [0, 31]
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, sh);
#else
    return sh ? (hi << sh) | (lo >> (32 - sh)) : hi;
#endif
}

// Long-code walk (code length > kLutBits), ljt/jdhuff.c jpeg_huff_decode:398-450.  Returns a LUT-format entry; an
// invalid code comes back as length 16, size 0, zadv 1 with ok = false.
B2J_HD uint32_t huff_long_code(const DevHuffTable* t, uint32_t win, bool& ok)
{
    for (int l = kLutBits + 1; l <= 16; ++l) {
        const int code = int(win >> (32 - l));
        if (code <= t->maxcode[l])
            return huff_entry(l, t->vals[(code + t->valoffset[l]) & 255], t->is_dc != 0);
    }
    ok = false;
    return 16u | (1u << 9);
}

// Everything a decoding thread needs to know about its image; the pointers may be shared or global memory.
struct HuffEnv
{
    const uint32_t* words;   // unstuffed stream of the image (4-byte aligned, readable 16 bytes past the end); with
                             // PRESWAPPED the words were byte-swapped to big-endian bit order when they were staged
    const uint16_t* lut;     // LUT base the TabSel offsets refer to
    const TabSel* tabsel;    // [bpm]
    const DevHuffTable* pool; // global table pool (long-code walk)
    const uint32_t* seg;     // segment starts in unstuffed bytes, [num_segments + 1]
    int32_t num_segments;
    uint32_t seg_blocks;
    uint32_t total_blocks;
    uint32_t clean_bits;
    int32_t bpm;
};

struct HuffState
{
    uint32_t p;      // bit position of the next symbol
    uint32_t z;      // zigzag index inside the current block (0 = the DC symbol is next)
    uint32_t blk;    // block inside the MCU
    uint32_t nblk;   // blocks completed: absolute block index of the current block if abs, else relative to the entry
    bool abs;
};

// Decode every symbol that starts in [s.p, end) (segment changes included), beginning in state s.
// WRITE = false: synchronisation pass, nothing is stored.
// WRITE = true : s.abs must hold.  Coefficients of blocks >= first_block are stored: AC de-zigzagged into coef (64 int16
//                per block, absolute block index), the DC *difference* into dc[block].  With extend = true the thread
//                keeps going past `end` until the block it is in is complete (the last subsequence of a CTA finishes
//                the block that straddles into the next CTA's range).  Returns error flags.
template <bool WRITE, bool PRESWAPPED = false>
B2J_HD uint32_t decode_span(const HuffEnv& env, HuffState& s, uint32_t end, bool extend, uint32_t first_block, int16_t* coef, int16_t* dc)
{
    if (end > env.clean_bits)
        end = env.clean_bits;
    uint32_t err = 0;
    if (WRITE && (s.nblk >= env.total_blocks || (extend && s.p >= end && s.z == 0)))
        return 0; // nothing left to decode (trailing bytes after the last block / the block ended with the subsequence)
    const int nseg = env.num_segments;
    int g = 0;
    if (nseg > 1) { // locate the segment containing p
        int lo = 0, hi = nseg - 1;
        const uint32_t pb = s.p >> 3;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (env.seg[mid] <= pb)
                lo = mid;
            else
                hi = mid - 1;
        }
        g = lo;
        if (!s.abs && s.p == env.seg[g] * 8u) { // exactly at a segment start: the state is known
            s.abs = true;
            s.nblk = uint32_t(g) * env.seg_blocks;
            s.z = 0;
            s.blk = 0;
        }
    }
    uint32_t p = s.p, z = s.z, blk = s.blk, nblk = s.nblk;
    bool stop = false;
    while (!stop) {
        const bool last = (g == nseg - 1);
        const uint32_t B = last ? env.clean_bits : env.seg[g + 1] * 8u; // end of this segment's bits
        const uint32_t lim = extend ? B : (end < B ? end : B);
        const uint32_t seg_end = last ? env.total_blocks : uint32_t(g + 1) * env.seg_blocks; // used when s.abs
        TabSel ts = env.tabsel[blk];
        uint32_t widx = p >> 5;
        uint32_t w0 = env.words[widx], w1 = env.words[widx + 1], w2 = env.words[widx + 2];
        if (!PRESWAPPED) {
            w0 = bswap32(w0);
            w1 = bswap32(w1);
            w2 = bswap32(w2);
        }
        bool cross = false;
        while (p < lim) {
            const uint32_t win = funnel_left(w0, w1, p & 31u);
            const uint32_t base = z ? ts.ac_off : ts.dc_off;
            uint32_t e = env.lut[base + (win >> (32 - kLutBits))];
            if ((e & 31u) == 0) { // code longer than kLutBits
                bool ok = true;
                if (e) {
                    const uint32_t n = (e >> 5) & 7u;
                    e = env.lut[base + uint32_t(kLutSize) + (e >> 8) * 2u + ((win >> (32 - kLutBits - n)) & ((1u << n) - 1u))];
                    if (e == 0) { // unused code space below that prefix
                        ok = false;
                        e = 16u | (1u << 9);
                    }
                } else {
                    e = huff_long_code(env.pool + (z ? ts.ac_idx : ts.dc_idx), win, ok);
                }
                if (WRITE && !ok)
                    err |= kErrBadCode;
            }
            const uint32_t len = e & 31u, sz = (e >> 5) & 15u, zadv = e >> 9;
            const uint32_t newp = p + len + sz;
            if (newp > B) { // the "symbol" runs over the segment end: it was padding
                cross = true;
                break;
            }
            if (WRITE) {
                int val = 0;
                if (sz) {
                    const uint32_t bits = (win << len) >> (32u - sz);
                    val = int(bits);
                    if (!(bits >> (sz - 1u)))
                        val -= (1 << sz) - 1; // HUFF_EXTEND
                }
                if (nblk - first_block < env.total_blocks - first_block) { // first_block <= nblk < total_blocks
                    if (z == 0)
                        dc[nblk] = int16_t(val);
                    else if (sz) {
                        uint32_t zt = z + zadv - 1u;
                        if (zt > 63u) {
                            zt = 63u;
                            err |= kErrBadCode;
                        }
                        coef[uint64_t(nblk) * 64u + B2J_ZZ(zt)] = int16_t(val);
                    }
                }
            }
            z += zadv;
            p = newp;
            if ((p >> 5) != widx) {
                ++widx;
                w0 = w1;
                w1 = w2;
                w2 = PRESWAPPED ? env.words[widx + 2] : bswap32(env.words[widx + 2]);
            }
            if (z >= 64u) { // block complete
                z = 0;
                ++nblk;
                ++blk;
                if (blk == uint32_t(env.bpm))
                    blk = 0;
                ts = env.tabsel[blk];
                if (WRITE && nblk >= seg_end) { // all blocks of the segment are there: what follows is padding
                    cross = true;
                    break;
                }
                if (extend && p >= end) {
                    stop = true;
                    break;
                }
            }
        }
        if (stop)
            break;
        if (cross || p >= B) {
            if (last) {
                if (cross) {
                    p = B; // trailing pad bits
                    if (WRITE && z != 0)
                        err |= kErrOverrun;
                }
                break;
            }
            if (WRITE && (z != 0 || nblk != uint32_t(g + 1) * env.seg_blocks))
                err |= kErrOverrun; // the segment's data ended before (or after) its blocks were complete
            ++g; // next restart segment: byte aligned, predictors and position reset (process_restart)
            p = env.seg[g] * 8u;
            z = 0;
            blk = 0;
            nblk = uint32_t(g) * env.seg_blocks;
            s.abs = true;
            if (extend && p >= end)
                break; // the straddling block ended with its segment
        } else {
            break; // p >= end
        }
    }
    s.p = p;
    s.z = z;
    s.blk = blk;
    s.nblk = nblk;
    return err;
}

} // namespace b200jpeg
