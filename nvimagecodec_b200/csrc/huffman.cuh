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
#if !defined(__CUDACC__)
#include "host_shim.h"
#endif

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

B2J_HD uint32_t funnel_left(uint32_t hi, uint32_t lo, uint32_t sh) // top 32 bits of (hi:lo) << (sh mod 32)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, sh); // the hardware shift wraps at 32: callers pass the bit position itself
#else
    sh &= 31u;
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

// Entry of a code longer than kLutBits (first-level entry e0 with a zero length field): second-level table, or the
// maxcode walk when the table was too irregular for it.  Invalid codes come back as "16 bits, no value, advance 1".
B2J_HD uint32_t huff_long_entry(const uint16_t* lut, uint32_t base, uint32_t e0, const DevHuffTable* tab, uint32_t win, bool& ok)
{
    if (e0) {
        const uint32_t n = (e0 >> 5) & 7u;
        const uint32_t e = lut[base + uint32_t(kLutSize) + (e0 >> 8) * 2u + ((win >> (32 - kLutBits - n)) & ((1u << n) - 1u))];
        if (e)
            return e;
        ok = false; // unused code space below that prefix
        return 16u | (1u << 9);
    }
    return huff_long_code(tab, win, ok);
}

struct HuffEnv;
B2J_HD uint32_t staged_word(const HuffEnv& env, uint32_t wi);

// Everything a decoding thread needs to know about its image; the pointers may be shared or global memory.
struct HuffEnv
{
    const uint32_t* words;   // unstuffed stream of the image (4-byte aligned, readable 16 bytes past the end); with
                             // PRESWAPPED the words were byte-swapped to big-endian bit order when they were staged
    uint32_t words_s;        // device, PRESWAPPED only: shared-window address of words[0] (the fast loop reads with ld.shared)
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

B2J_HD uint32_t staged_word(const HuffEnv& env, uint32_t wi) // word wi of the staged (byte-swapped) stream
{
#if defined(__CUDA_ARCH__)
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(env.words_s + wi * 4u));
    return v;
#else
    return env.words[wi];
#endif
}

struct HuffState
{
    uint32_t p;      // bit position of the next symbol
    uint32_t z;      // zigzag index inside the current block (0 = the DC symbol is next)
    uint32_t blk;    // block inside the MCU
    uint32_t nblk;   // blocks completed: absolute block index of the current block if abs, else relative to the entry
    bool abs;
};

// Where the block pass puts what it finds (pointers are per image): for every block the DC difference and the bit position
// of its first AC symbol.  The AC symbols themselves are decoded by the thread that reconstructs the block
// (decode_block_ac below, called from the IDCT kernel): coefficients never travel through memory.
struct BlkOut
{
    int16_t* dc;        // per block: DC difference
    uint32_t* pos;      // per block: bit position (in the image's unstuffed stream) right behind the DC symbol
};

B2J_HD uint32_t shr_upto32(uint32_t x, uint32_t n) // x >> n for n in [0, 32]
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_rc(x, 0u, n);
#else
    return n >= 32u ? 0u : x >> n;
#endif
}

// value bits -> coefficient (HUFF_EXTEND, ljt/jdhuff.c:483-488) without branches.  t = stream window shifted so that the
// first value bit is bit 31; sz = number of value bits.  A first value bit of 1 means a positive coefficient equal to the sz
// bits; a 0 means the sz bits minus (2^sz - 1).  sz == 0 yields 0.
B2J_HD int huff_extend(uint32_t t, uint32_t sz)
{
    const uint32_t v = shr_upto32(t, 32u - sz);
    const uint32_t neg = ~uint32_t(int(t) >> 31); // all ones when the first value bit is 0
    return int(v) - int(neg & ((1u << sz) - 1u));
}

// Decode every symbol that starts in [s.p, end) (segment changes included), beginning in state s.
// WRITE = false: synchronisation pass, nothing is stored.
// WRITE = true : block pass, s.abs must hold.  Every DC symbol stores the DC difference of its block and the bit position
//                behind it (where the block's AC symbols start).  Returns error flags.
// DIRECT selects the inner loop used well inside a segment: true = both stream words fetched from shared memory per
// symbol (fewest instructions: the throughput kernels, whose warps hide the extra latency), false = sliding window in
// registers (shortest dependency chain: the latency-bound work-list rounds).
template <bool WRITE, bool PRESWAPPED = false, bool DIRECT = true>
B2J_HD uint32_t decode_span(const HuffEnv& env, HuffState& s, uint32_t end, BlkOut* out)
{
    if (end > env.clean_bits)
        end = env.clean_bits;
    uint32_t err = 0;
    if (WRITE && s.nblk >= env.total_blocks)
        return 0; // trailing bytes after the last block
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
    for (;;) {
        const bool last = (g == nseg - 1);
        const uint32_t B = last ? env.clean_bits : env.seg[g + 1] * 8u; // end of this segment's bits
        const uint32_t lim = end < B ? end : B;
        const uint32_t seg_end = last ? env.total_blocks : uint32_t(g + 1) * env.seg_blocks; // used when s.abs
        TabSel ts = env.tabsel[blk];
        bool cross = false;
        if (PRESWAPPED && DIRECT && p < lim && lim + 32u <= B && (!WRITE || nblk + (lim - p) / 2u + 16u < seg_end)) {
            // Well inside a segment (the common case): no symbol can run over the segment end (and, writing, the blocks
            // of the segment cannot run out: a block takes at least two bits), so the loop carries nothing but the
            // decoder state.  Both stream words are fetched per symbol (shared memory): cheaper than keeping a sliding
            // window up to date in a loop whose lanes all diverge.
            do {
                const uint32_t wi = p >> 5;
                const uint32_t win = funnel_left(staged_word(env, wi), staged_word(env, wi + 1), p);
                const uint32_t base = z ? ts.ac_off : ts.dc_off;
                uint32_t e = env.lut[base + (win >> (32 - kLutBits))];
                if ((e & 31u) == 0) {
                    bool ok = true;
                    e = huff_long_entry(env.lut, base, e, env.pool + (z ? ts.ac_idx : ts.dc_idx), win, ok);
                    if (WRITE && !ok)
                        err |= kErrBadCode;
                }
                const uint32_t tot = e & 31u, zadv = e >> 9;
                p += tot;
                if (WRITE && z == 0) {
                    const uint32_t sz = (e >> 5) & 15u;
                    out->dc[nblk] = int16_t(huff_extend(win << (tot - sz), sz));
                    out->pos[nblk] = p;
                }
                z += zadv;
                if (z >= 64u) {
                    z = 0;
                    ++nblk;
                    ++blk;
                    if (blk == uint32_t(env.bpm))
                        blk = 0;
                    ts = env.tabsel[blk];
                }
            } while (p < lim);
            break; // p >= end, still inside the segment
        }
        uint32_t widx = p >> 5;
        uint32_t w0 = env.words[widx], w1 = env.words[widx + 1], w2 = env.words[widx + 2];
        if (!PRESWAPPED) {
            w0 = bswap32(w0);
            w1 = bswap32(w1);
            w2 = bswap32(w2);
        }
        while (p < lim) {
            const uint32_t win = funnel_left(w0, w1, p);
            const uint32_t base = z ? ts.ac_off : ts.dc_off;
            uint32_t e = env.lut[base + (win >> (32 - kLutBits))];
            bool bad = false;
            if ((e & 31u) == 0) { // code longer than kLutBits
                bool ok = true;
                e = huff_long_entry(env.lut, base, e, env.pool + (z ? ts.ac_idx : ts.dc_idx), win, ok);
                bad = !ok;
            }
            const uint32_t tot = e & 31u, sz = (e >> 5) & 15u, zadv = e >> 9;
            const uint32_t newp = p + tot;
            if (newp > B) { // the "symbol" runs over the segment end: it was padding (a thread may well start inside
                cross = true; // the pad bits of a segment whose blocks are complete; they never form a whole code)
                break;
            }
            if (WRITE && bad)
                err |= kErrBadCode;
            if (WRITE && z == 0) {
                out->dc[nblk] = int16_t(huff_extend(win << (tot - sz), sz));
                out->pos[nblk] = newp;
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
            }
        }
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
            if (p >= end)
                break;
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

// AC coefficients of one block straight from the stream (ljt/jdhuff.c decode_mcu_slow:598-640): p = bit position of the
// block's first AC symbol (BlkOut::pos).  words = the image's unstuffed stream as stored (byte order of the file), readable
// up to word index max_word; lut / ac_off / tab = the block's AC table as staged by the caller.  blk = kBlkSlots
// natural-order coefficients, zeroed by the caller (entry 0 is not touched: the DC value travels separately; entry 64 is a
// dump).  zzd = kZigzagDump entries: zigzag position -> natural index, 64 (the dump) for positions behind the block; on the
// device a shared-memory copy (the lanes of a warp look up different positions, which constant memory would serialise).
// Every symbol stores: a ZRL or end-of-block symbol writes a zero to a position that is zero anyway, or to the dump -- no
// branch in the loop but the one for codes longer than the first-level table and the one that refills the window.
// Returns error flags (invalid code, a run that leaves the block: corrupt data).  At most 63 symbols are read.
constexpr int kBlkSlots = 72;     // 64 coefficients + dump, padded to a 16-byte multiple
constexpr int kZigzagDump = 128;  // positions 1 .. 63 + 63 can be asked for
B2J_HD uint8_t zigzag_dump_entry(const uint8_t* zigzag64, int k) { return k < 64 ? zigzag64[k] : uint8_t(64); }

// STAGED: words are already in bit order and sit in fast memory (the caller staged the span its blocks need): both words of
// the window are read per symbol.  Otherwise words are as stored (file byte order) in global memory and a three-word window
// slides along in registers.
template <bool STAGED>
B2J_HD uint32_t decode_block_ac(const uint32_t* words, uint32_t max_word, uint32_t p, const uint16_t* lut, uint32_t ac_off,
    const DevHuffTable* tab, const uint8_t* zzd, int16_t* blk)
{
    uint32_t widx = p >> 5;
    if (widx > max_word)
        return kErrOverrun;
    auto load = [&](uint32_t i) {
        // STAGED: no clamp.  A block takes at most 63 symbols of at most 31 bits, so damaged data can carry the window 62 words
        // past the word its block starts in and no further; the caller keeps that much readable memory behind the staged
        // words (k_idct: the Huffman tables follow the stream in shared memory), and what is read there is garbage for garbage.
        if (!STAGED)
            i = i > max_word ? max_word : i;
#if defined(__CUDA_ARCH__)
        return STAGED ? words[i] : bswap32(__ldg(words + i));
#else
        return STAGED ? words[i] : bswap32(words[i]);
#endif
    };
    uint32_t w0 = 0, w1 = 0, w2 = 0;
    if (!STAGED) {
        w0 = load(widx);
        w1 = load(widx + 1);
        w2 = load(widx + 2);
    }
    uint32_t z = 1, err = 0, zadv;
    do {
        uint32_t win;
        if (STAGED) {
            const uint32_t wi = p >> 5;
            win = funnel_left(load(wi), load(wi + 1), p);
        } else {
            win = funnel_left(w0, w1, p);
        }
        uint32_t e = lut[ac_off + (win >> (32 - kLutBits))];
        if ((e & 31u) == 0) { // code longer than kLutBits
            bool ok = true;
            e = huff_long_entry(lut, ac_off, e, tab, win, ok);
            if (!ok)
                err |= kErrBadCode;
        }
        const uint32_t tot = e & 31u, sz = (e >> 5) & 15u;
        zadv = e >> 9;
        blk[zzd[z + zadv - 1u]] = int16_t(huff_extend(win << (tot - sz), sz)); // position of the coefficient: after zadv - 1 zeros
        z += zadv;
        p += tot;
        if (!STAGED && (p >> 5) != widx) {
            ++widx;
            w0 = w1;
            w1 = w2;
            w2 = load(widx + 2);
        }
    } while (z < 64u);
    // z == 64: the last coefficient sat at position 63; z >= 65 with zadv == 64: end of block; anything else ran over the block
    if (z > 64u && zadv != 64u)
        err |= kErrBadCode;
    return err;
}

// Decoder state at a subsequence boundary in 32 bits, as the fused kernel keeps it in shared memory and hands it from CTA to
// CTA: bit position relative to the start of subsequence k (an exit state lies at most one symbol behind the subsequence's
// end: S + 63 < 2^16 for subsequences up to 4 KB... the planner allows 256 bytes) | (block-in-MCU * 64 + zigzag index) << 16.
B2J_HD uint32_t pack_rel(uint32_t p, uint32_t k, uint32_t S, uint32_t slot) { return (p - k * S) | (slot << 16); }
B2J_HD uint32_t rel_pos(uint32_t packed, uint32_t k, uint32_t S) { return k * S + (packed & 0xFFFFu); }

// Run-ahead shortcut for periodic streams (flat image areas: every MCU has the same bits, so a decoder that starts at
// the wrong phase can stay there for ever and the repair travels one subsequence at a time).  If the decode of
// subsequence k went from state (p, slot) to (p + S, slot) and the stream words of subsequence k + 1 -- including the two
// words a last symbol or table lookup may reach into -- equal those of subsequence k, the decode of k + 1 from that exit
// state is the same decode shifted by S bits: nothing needs to be decoded.  `words` = the image's stream (either byte
// order), first = first word of subsequence k, spw = words per subsequence.
B2J_HD bool subseq_repeats(const uint32_t* words, uint32_t first, uint32_t spw)
{
    for (uint32_t j = 0; j < spw + 2u; ++j)
        if (words[first + j] != words[first + spw + j])
            return false;
    return true;
}

} // namespace b200jpeg
