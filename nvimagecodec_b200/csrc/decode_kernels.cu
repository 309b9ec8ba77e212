// decode_kernels.cu -- sm_100a kernels of the b200jpeg decode path and their launcher.
//
// Pipeline for one batch (all kernels are descriptor-driven: one launch covers every image of the batch, a CTA finds its
// image with a binary search over a per-kernel prefix map):
//   U1 k_unstuff_count   count kept bytes / RSTn markers per 16 KB chunk of entropy-coded data
//   U2 k_unstuff_scan    per image: exclusive scan of the chunk counts, unstuffed length, segment table init
//   U3 k_unstuff_write   compact: drop FF00 stuffing, FF fill, RSTn markers; record restart-segment starts
//   H1 k_huff_sync x 2   self-synchronising subsequence decode (no output): round 0 from a guessed state, round 1 from the
//                        predecessor's exit state; subsequences whose exit state changed feed a work list
//   H2 k_huff_sync_tail  cooperative kernel: Jacobi rounds over the (compacted) work list until it is empty
//   H3 k_subseq_scan     segmented scan of coefficient-slot counts -> absolute block position per subsequence (chunks of an image chained)
//   H4 k_huff_write      block pass: per block the DC difference and the bit position of its first AC symbol
//   D1 k_dc_scan         DC prediction = segmented (restart interval) prefix sum per component over dc[] (chunks of an image chained)
//   I1 k_idct            one 8x8 block per thread: Huffman-decode the block's AC symbols from the stream into a dense block
//                        in shared memory, dequantise + islow IDCT, write the component planes -- coefficients never leave
//                        the SM
//   C1 k_color           upsample (fancy / box) + YCbCr->RGB; a warp produces 512 pixels x (1|2) rows, staged in shared
//                        memory and written with fully coalesced 128-byte store instructions at any alignment
// The per-thread arithmetic lives in huffman.cuh / jpeg_math.cuh (shared with the CPU emulation used by the tests).
#include <cuda_runtime.h>
#include "decode_kernels.h"
#include "huffman.cuh"
#include "bulk_copy.cuh"
#include "jpeg_math.cuh"
#include "cta_lookup.cuh"

namespace b200jpeg {

// ------------------------------------------------------------------------------------------------------------------
// Unstuffing
// ------------------------------------------------------------------------------------------------------------------
struct ByteClass
{
    uint32_t keep;   // bit i: byte i is entropy-coded data
    uint32_t marker; // bit i: byte i is the FF of an RSTn marker
    uint32_t bad;    // bit i: byte i is the FF of any other marker (none may occur inside entropy-coded data)
};

// byte mask (0xFF / 0x00 per byte) -> 4 bits
__device__ __forceinline__ uint32_t mask4(uint32_t m) { return ((m & 0x01010101u) * 0x01020408u) >> 24; }

// Classify the 16 bytes at [base, base+16) of an image's raw scan data, four bytes per SIMD-in-register step.
//   removed: 00 after FF (stuffing), FF before FF (fill), FF Dn and the Dn after FF (restart marker); anything else after
//   an FF is a marker that must not occur inside entropy-coded data.
__device__ __forceinline__ ByteClass classify16(const uint8_t* __restrict__ raw, uint32_t raw_len, uint32_t base, uint32_t w[4], uint32_t& next)
{
    if (base + 16 <= raw_len) {
        const uint4 v = *reinterpret_cast<const uint4*>(raw + base);
        w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
    } else {
        w[0] = w[1] = w[2] = w[3] = 0;
        for (uint32_t i = 0; base + i < raw_len && i < 16; ++i)
            w[i >> 2] |= uint32_t(raw[base + i]) << (8 * (i & 3));
    }
    const uint32_t prev = (base > 0 && base <= raw_len) ? raw[base - 1] : 0;
    next = (base + 16 < raw_len) ? raw[base + 16] : 0;
    uint32_t remove = 0, marker = 0, bad = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t cur = w[j];
        const uint32_t pw = (cur << 8) | (j ? (w[j > 0 ? j - 1 : 0] >> 24) : prev);
        const uint32_t nw = (cur >> 8) | ((j < 3 ? w[j < 3 ? j + 1 : 3] : next) << 24);
        const uint32_t ff = __vcmpeq4(cur, 0xFFFFFFFFu), zz = __vcmpeq4(cur, 0u), rst = __vcmpeq4(cur & 0xF8F8F8F8u, 0xD0D0D0D0u);
        const uint32_t pff = __vcmpeq4(pw, 0xFFFFFFFFu);
        const uint32_t nff = __vcmpeq4(nw, 0xFFFFFFFFu), nz = __vcmpeq4(nw, 0u), nrst = __vcmpeq4(nw & 0xF8F8F8F8u, 0xD0D0D0D0u);
        remove |= mask4((zz & pff) | (ff & (nff | nrst)) | (rst & pff)) << (4 * j);
        marker |= mask4(ff & nrst) << (4 * j);
        bad |= mask4(ff & ~(nz | nff | nrst)) << (4 * j);
    }
    const uint32_t n_in = raw_len - base < 16u ? raw_len - base : 16u;
    const uint32_t in_mask = (1u << n_in) - 1u;
    ByteClass c;
    c.keep = ~remove & in_mask;
    c.marker = marker & in_mask;
    const uint32_t n_succ = raw_len - base - 1u < 16u ? raw_len - base - 1u : 16u; // bytes that have a successor inside the data
    c.bad = bad & ((1u << n_succ) - 1u);
    return c;
}

// Classification of the CTA's chunk: every thread has kUnstuffGranules consecutive 16-byte granules.  Only a granule that
// holds an FF (or follows one) can have anything removed, and only one byte in 256 is an FF -- but nearly every warp has a
// lane with such a granule, so a per-lane slow path would run in nearly every warp.  Instead the few granules that need a
// look are queued in shared memory and classified by the first threads of the CTA (dense warps); everybody else keeps all
// 16 bytes.  w = the thread's bytes (zero-filled behind the data).  Contains barriers: call with the whole CTA.
constexpr int kChunkGranules = kUnstuffThreads * kUnstuffGranules;
struct CtaClassify
{
    uint32_t n;
    uint16_t list[kChunkGranules], keep[kChunkGranules], marker[kChunkGranules], bad[kChunkGranules];
};

__device__ __forceinline__ void classify_cta(CtaClassify& sh, const uint8_t* __restrict__ raw, uint32_t len, uint32_t base,
    uint32_t w[kUnstuffGranules][4], ByteClass c[kUnstuffGranules])
{
    const uint32_t lane = threadIdx.x & 31u;
    if (threadIdx.x == 0)
        sh.n = 0;
    __syncthreads();
    uint32_t slow = 0; // bit q: granule q has to be looked at
#pragma unroll
    for (int q = 0; q < kUnstuffGranules; ++q) {
        const uint32_t gb = base + 16u * q;
        w[q][0] = w[q][1] = w[q][2] = w[q][3] = 0;
        c[q] = ByteClass{0, 0, 0};
        if (gb + 16 <= len) {
            const uint4 v = *reinterpret_cast<const uint4*>(raw + gb);
            w[q][0] = v.x; w[q][1] = v.y; w[q][2] = v.z; w[q][3] = v.w;
        } else if (gb < len) { // the last granule of the data
            for (uint32_t i = 0; gb + i < len; ++i)
                w[q][i >> 2] |= uint32_t(raw[gb + i]) << (8 * (i & 3));
            slow |= 1u << q;
        }
    }
    const uint32_t up = __shfl_up_sync(0xFFFFFFFFu, w[kUnstuffGranules - 1][3], 1);
#pragma unroll
    for (int q = 0; q < kUnstuffGranules; ++q) {
        const uint32_t gb = base + 16u * q;
        if (gb < len && !((slow >> q) & 1u)) {
            const uint32_t prev = q ? w[q > 0 ? q - 1 : 0][3] >> 24 : (lane ? up >> 24 : (gb ? raw[gb - 1] : 0u));
            const uint32_t ff = __vcmpeq4(w[q][0], 0xFFFFFFFFu) | __vcmpeq4(w[q][1], 0xFFFFFFFFu) | __vcmpeq4(w[q][2], 0xFFFFFFFFu) |
                                __vcmpeq4(w[q][3], 0xFFFFFFFFu);
            if (ff || prev == 0xFFu)
                slow |= 1u << q;
            else
                c[q].keep = 0xFFFFu;
        }
    }
    if (slow) {
        uint32_t at = atomicAdd(&sh.n, uint32_t(__popc(slow)));
#pragma unroll
        for (int q = 0; q < kUnstuffGranules; ++q)
            if ((slow >> q) & 1u)
                sh.list[at++] = uint16_t(threadIdx.x * kUnstuffGranules + q);
    }
    __syncthreads();
    const uint32_t n = sh.n, chunk_base = base - threadIdx.x * kUnstuffBytesPerThread;
    for (uint32_t j = threadIdx.x; j < n; j += kUnstuffThreads) {
        const uint32_t g = sh.list[j];
        uint32_t ww[4], nx;
        const ByteClass cc = classify16(raw, len, chunk_base + g * 16u, ww, nx);
        sh.keep[g] = uint16_t(cc.keep);
        sh.marker[g] = uint16_t(cc.marker);
        sh.bad[g] = uint16_t(cc.bad);
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < kUnstuffGranules; ++q)
        if ((slow >> q) & 1u) {
            const uint32_t g = threadIdx.x * kUnstuffGranules + q;
            c[q].keep = sh.keep[g];
            c[q].marker = sh.marker[g];
            c[q].bad = sh.bad[g];
        }
}

__global__ void __launch_bounds__(kUnstuffThreads) k_unstuff_count(BatchDev b)
{
    const uint32_t cta_abs = blockIdx.x + b.base_unstuff;
    const int img = find_image_warp(b.cta_map_unstuff, b.num_images, cta_abs);
    const ImageDesc& im = b.images[img];
    const uint32_t chunk = cta_abs - b.cta_map_unstuff[img];
    const uint32_t base = chunk * kChunkBytes + threadIdx.x * kUnstuffBytesPerThread;
    uint32_t w[kUnstuffGranules][4];
    ByteClass c[kUnstuffGranules];
    __shared__ CtaClassify s_cls;
    classify_cta(s_cls, b.raw + im.raw_off, im.raw_len, base, w, c);
    uint32_t kept = 0, markers = 0, bad_at = 0xFFFFFFFFu;
#pragma unroll
    for (int q = kUnstuffGranules - 1; q >= 0; --q) {
        kept += __popc(c[q].keep);
        markers += __popc(c[q].marker);
        if (c[q].bad)
            bad_at = base + 16u * q + uint32_t(__ffs(int(c[q].bad))) - 1u;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        kept += __shfl_down_sync(0xFFFFFFFFu, kept, o);
        markers += __shfl_down_sync(0xFFFFFFFFu, markers, o);
    }
    __shared__ uint32_t s_k[kUnstuffThreads / 32], s_m[kUnstuffThreads / 32];
    if ((threadIdx.x & 31) == 0) {
        s_k[threadIdx.x >> 5] = kept;
        s_m[threadIdx.x >> 5] = markers;
    }
    // A marker other than RSTn ends the entropy-coded data there, as it does for the reference (ljt/jdhuff.c
    // jpeg_fill_bit_buffer:287-391 stops feeding bytes at the first marker): files that carry more after their first EOI
    // (MPO / gain-map / motion-photo appendices, embedded thumbnails) decode like everywhere else.  The earliest such
    // marker wins: the per-image maximum of raw_len - position.
    if (bad_at != 0xFFFFFFFFu)
        atomicMax(&b.states[img].trunc, im.raw_len - bad_at);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t k = 0, m = 0;
#pragma unroll
        for (int i = 0; i < kUnstuffThreads / 32; ++i) {
            k += s_k[i];
            m += s_m[i];
        }
        b.chunk_counts[im.chunk_first + chunk] = uint64_t(k) | (uint64_t(m) << 32);
    }
}

// one warp per image
__global__ void __launch_bounds__(32) k_unstuff_scan(BatchDev b)
{
    const int img = blockIdx.x;
    const ImageDesc& im = b.images[img];
    const uint32_t lane = threadIdx.x;
    // entropy-coded bytes end at the first marker that is not RSTn (k_unstuff_count): chunks behind that point count
    // nothing, the chunk it falls into is counted again up to it
    const uint32_t trunc = b.states[img].trunc;
    const uint32_t eff_len = im.raw_len - trunc;
    const uint32_t cut = trunc ? eff_len / kChunkBytes : 0xFFFFFFFFu;
    uint64_t cut_count = 0;
    if (trunc) {
        uint32_t kept = 0, markers = 0;
        for (uint32_t off = lane * 16u; off < uint32_t(kChunkBytes); off += 32u * 16u) {
            const uint32_t base = cut * kChunkBytes + off;
            if (base < eff_len) {
                uint32_t w[4], next;
                const ByteClass c = classify16(b.raw + im.raw_off, eff_len, base, w, next);
                kept += __popc(c.keep);
                markers += __popc(c.marker);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            kept += __shfl_xor_sync(0xFFFFFFFFu, kept, o);
            markers += __shfl_xor_sync(0xFFFFFFFFu, markers, o);
        }
        cut_count = uint64_t(kept) | (uint64_t(markers) << 32);
    }
    uint64_t carry = 0;
    for (uint32_t c0 = 0; c0 < im.chunk_count; c0 += 32) {
        const uint32_t c = c0 + lane;
        uint64_t v = c < im.chunk_count ? b.chunk_counts[im.chunk_first + c] : 0;
        if (c >= cut)
            v = c == cut ? cut_count : 0;
        uint64_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint64_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= uint32_t(o))
                inc += t;
        }
        if (c < im.chunk_count)
            b.chunk_counts[im.chunk_first + c] = carry + inc - v;
        carry += __shfl_sync(0xFFFFFFFFu, inc, 31);
    }
    const uint32_t clean_len = uint32_t(carry), markers = uint32_t(carry >> 32);
    uint32_t* S = b.seg_start + im.seg_first;
    for (uint32_t g = lane; g <= uint32_t(im.num_segments); g += 32)
        S[g] = g == 0 ? 0u : clean_len; // U3 overwrites [1 .. markers]
    if (lane == 0) {
        b.states[img].clean_len = clean_len;
        b.states[img].raw_eff = eff_len;
        b.states[img].marker_count = markers;
        if (markers != uint32_t(im.num_segments - 1))
            atomicOr(&b.states[img].error, kErrBadMarker);
    }
}

__global__ void __launch_bounds__(kUnstuffThreads) k_unstuff_write(BatchDev b)
{
    const uint32_t cta_abs = blockIdx.x + b.base_unstuff;
    const int img = find_image_warp(b.cta_map_unstuff, b.num_images, cta_abs);
    const ImageDesc& im = b.images[img];
    const uint32_t chunk = cta_abs - b.cta_map_unstuff[img];
    const uint32_t base = chunk * kChunkBytes + threadIdx.x * kUnstuffBytesPerThread;
    const uint32_t eff_len = b.states[img].raw_eff; // raw_len, or up to the first marker that is not RSTn
    const uint64_t cb = b.chunk_counts[im.chunk_first + chunk]; // (needed after the scan: requested now)
    uint32_t w[kUnstuffGranules][4];
    ByteClass c[kUnstuffGranules];
    __shared__ CtaClassify s_cls;
    __shared__ uint32_t s_w[kUnstuffThreads / 32];
    __shared__ __align__(16) uint32_t s_out[kChunkBytes / 4 + 4];
    __shared__ uint32_t s_total;
    classify_cta(s_cls, b.raw + im.raw_off, eff_len, base, w, c);
    // block exclusive scan of (kept | markers << 16)
    uint32_t mine = 0;
#pragma unroll
    for (int q = 0; q < kUnstuffGranules; ++q)
        mine += uint32_t(__popc(c[q].keep)) | (uint32_t(__popc(c[q].marker)) << 16);
    uint32_t inc = mine;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= uint32_t(o))
            inc += t;
    }
#pragma unroll
    for (int q = 0; q < kUnstuffGranules; ++q) // the compaction below ORs into the output words
        reinterpret_cast<uint4*>(s_out)[threadIdx.x * kUnstuffGranules + q] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0)
        reinterpret_cast<uint4*>(s_out)[kChunkGranules] = make_uint4(0, 0, 0, 0);
    if (lane == 31)
        s_w[warp] = inc;
    __syncthreads();
    uint32_t wpre = 0;
#pragma unroll
    for (int i = 0; i < kUnstuffThreads / 32; ++i)
        if (i < int(warp))
            wpre += s_w[i];
    const uint32_t excl = wpre + inc - mine;
    // Kept bytes are compacted into shared memory first, then copied out with aligned 16-byte stores.  The compaction
    // happens in registers: a granule's removed bytes (usually none, or the one zero behind a stuffed FF) are dropped from
    // its 128-bit value by shifting what lies above them down, and the value, shifted to its byte offset, is ORed into the
    // (zeroed) output words -- five shared-memory atomics whatever the alignment, instead of a store per byte.
    uint8_t* sb = reinterpret_cast<uint8_t*>(s_out);
    uint32_t lpos = excl & 0xFFFFu;                    // position inside the chunk's output
    const uint32_t gbase = uint32_t(cb);               // position of the chunk's output in the image's clean stream
    uint32_t ord = uint32_t(cb >> 32) + (excl >> 16);
    uint32_t bad = 0;
#pragma unroll
    for (int q = 0; q < kUnstuffGranules; ++q) {
        const uint32_t keep = c[q].keep, cnt = uint32_t(__popc(keep));
        if (c[q].marker) { // restart markers: the next segment starts at the clean position of the marker
            uint32_t* S = b.seg_start + im.seg_first;
            for (uint32_t mk = c[q].marker; mk; mk &= mk - 1u) {
                const int i = __ffs(int(mk)) - 1;
                const uint32_t at = base + 16u * q + uint32_t(i) + 1u; // the byte behind the FF
                const uint32_t nx = at < eff_len ? b.raw[im.raw_off + at] : 0u;
                if (ord + 1 < uint32_t(im.num_segments)) {
                    S[ord + 1] = gbase + lpos + uint32_t(__popc(keep & ((1u << i) - 1u)));
                    if (nx != 0xD0u + (ord & 7u))
                        bad = 1;
                } else
                    bad = 1;
                ++ord;
            }
        }
        if (cnt) {
            uint32_t v0 = w[q][0], v1 = w[q][1], v2 = w[q][2], v3 = w[q][3];
            uint32_t rem = ~keep & ((2u << (31 - __clz(int(keep)))) - 1u); // removed bytes below the last kept one
            while (rem) {
                const uint32_t r = uint32_t(__ffs(int(rem))) - 1u, wi = r >> 2;
                const uint32_t low = (1u << ((r & 3u) * 8u)) - 1u; // the bytes of word wi below byte r stay
                const uint32_t d0 = __funnelshift_r(v0, v1, 8), d1 = __funnelshift_r(v1, v2, 8), d2 = __funnelshift_r(v2, v3, 8), d3 = v3 >> 8;
                v0 = wi == 0 ? (v0 & low) | (d0 & ~low) : v0;
                v1 = wi == 1 ? (v1 & low) | (d1 & ~low) : (wi < 1 ? d1 : v1);
                v2 = wi == 2 ? (v2 & low) | (d2 & ~low) : (wi < 2 ? d2 : v2);
                v3 = wi == 3 ? (v3 & low) | (d3 & ~low) : (wi < 3 ? d3 : v3);
                rem = (rem >> (r + 1u)) << r; // what lay above moved down by one byte
            }
            if (cnt < 16u) { // bytes from cnt on must not reach the neighbour's words
                const uint32_t kw = cnt >> 2, km = (1u << ((cnt & 3u) * 8u)) - 1u;
                v0 = kw == 0 ? v0 & km : v0;
                v1 = kw == 1 ? v1 & km : (kw < 1 ? 0u : v1);
                v2 = kw == 2 ? v2 & km : (kw < 2 ? 0u : v2);
                v3 = kw == 3 ? v3 & km : (kw < 3 ? 0u : v3);
            }
            const uint32_t sh = (lpos & 3u) * 8u;
            uint32_t* o = s_out + (lpos >> 2);
            atomicOr(o, v0 << sh);
            atomicOr(o + 1, __funnelshift_l(v0, v1, sh));
            atomicOr(o + 2, __funnelshift_l(v1, v2, sh));
            atomicOr(o + 3, __funnelshift_l(v2, v3, sh));
            if (sh)
                atomicOr(o + 4, v3 >> (32u - sh));
            lpos += cnt;
        }
    }
    if (threadIdx.x == kUnstuffThreads - 1)
        s_total = lpos;
    if (bad)
        atomicOr(&b.states[img].error, kErrBadMarker);
    __syncthreads();
    // copy out: bytes up to the first 16-byte boundary of the destination, then 16-byte stores, then the last bytes
    const uint32_t K = s_total;
    uint8_t* d = b.clean + im.clean_off + gbase;
    uint32_t head = uint32_t(16u - (reinterpret_cast<uintptr_t>(d) & 15u)) & 15u;
    if (head > K)
        head = K;
    if (threadIdx.x < head)
        d[threadIdx.x] = sb[threadIdx.x];
    const uint32_t nv = (K - head) >> 4;
    uint4* d16 = reinterpret_cast<uint4*>(d + head);
    const uint32_t o0 = head >> 2, sh = (head & 3u) * 8u;
    for (uint32_t i = threadIdx.x; i < nv; i += kUnstuffThreads) {
        const uint32_t* sw = s_out + o0 + 4u * i;
        const uint32_t a0 = sw[0], a1 = sw[1], a2 = sw[2], a3 = sw[3], a4 = sw[4];
        d16[i] = make_uint4(__funnelshift_r(a0, a1, sh), __funnelshift_r(a1, a2, sh), __funnelshift_r(a2, a3, sh), __funnelshift_r(a3, a4, sh));
    }
    const uint32_t done = head + (nv << 4);
    if (threadIdx.x < K - done)
        d[done + threadIdx.x] = sb[done + threadIdx.x];
}

// ------------------------------------------------------------------------------------------------------------------
// DC prediction: per image, segmented (restart interval) inclusive scan of the DC differences per component.
// ------------------------------------------------------------------------------------------------------------------
// An image is cut into chunks of kDcChunkMcus MCUs, one CTA each.  Inside the CTA every warp takes a contiguous range of 256
// MCUs: it issues the loads of all its eight iterations first (the values go back into the array the differences come from,
// so the compiler cannot move a load across a store), scans 32 MCUs at a time with warp shuffles, carrying the totals along;
// one barrier combines the warps.  The chunks of an image form a chain: a CTA publishes what the image's MCUs up to its end
// leave in the three predictors (two 64-bit words, dc_rec), its successor waits for that (CTAs take their place from a ticket
// counter, so a CTA only waits for CTAs started before it).  (One CTA per image took 45 us for a 1080p image and 0.47 ms for a
// 4K 4:4:4 one, on 32 .. 64 SMs, on the critical path of its group.)
struct DcTotals
{
    int v0, v1, v2, flag; // sums per component since the last restart boundary; flag: a boundary was passed
};
__device__ __forceinline__ DcTotals dc_combine(const DcTotals& a, const DcTotals& x) // a = earlier, x = later
{
    return x.flag ? x : DcTotals{a.v0 + x.v0, a.v1 + x.v1, a.v2 + x.v2, a.flag};
}

__global__ void __launch_bounds__(kDcCta, 2) k_dc_scan(BatchDev b)
{
    constexpr int NW = kDcCta / 32;
    __shared__ DcTotals s_w[NW];
    __shared__ DcTotals s_pred;
    __shared__ uint32_t s_ticket;
    if (threadIdx.x == 0)
        s_ticket = atomicAdd(&b.sync->ticket_dc, 1u);
    __syncthreads();
    const uint32_t cta_abs = s_ticket + b.base_dc;
    const int img = find_image_warp(b.cta_map_dc, b.num_images, cta_abs);
    const ImageDesc& im = b.images[img];
    const uint32_t k = cta_abs - b.cta_map_dc[img];
    const uint32_t mcus = uint32_t(im.mcus_x) * uint32_t(im.mcus_y);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int16_t* dc = b.dc + im.coef_off / 64;
    const uint32_t ri = im.restart_interval ? uint32_t(im.restart_interval) : 0xFFFFFFFFu;
    const uint32_t bpm = uint32_t(im.bpm);
    const uint32_t w_lo = min(k * uint32_t(kDcChunkMcus) + warp * 32u * kDcBatch, mcus), w_hi = min(w_lo + 32u * kDcBatch, mcus);
    uint32_t comp_of = 0; // 2 bits per block of the MCU
    for (uint32_t j = 0; j < bpm; ++j)
        comp_of |= uint32_t(im.blk_comp[j]) << (2 * j);
    int s0[kDcBatch], s1[kDcBatch], s2[kDcBatch];
#pragma unroll
    for (int it = 0; it < kDcBatch; ++it) {
        const uint32_t m = w_lo + uint32_t(it) * 32u + lane;
        s0[it] = s1[it] = s2[it] = 0;
        if (m < w_hi) {
            const int16_t* pm = dc + uint64_t(m) * bpm;
            for (uint32_t j = 0; j < bpm; ++j) {
                const int dv = pm[j];
                const uint32_t c = (comp_of >> (2 * j)) & 3u;
                s0[it] += c == 0 ? dv : 0;
                s1[it] += c == 1 ? dv : 0;
                s2[it] += c == 2 ? dv : 0;
            }
        }
    }
    // inclusive totals through the lane's MCU of each iteration, relative to the start of the warp's range
    int i0[kDcBatch], i1[kDcBatch], i2[kDcBatch];
    uint32_t fmask = 0; // bit it: a restart boundary lies between the start of the warp's range and this MCU (inclusive)
    DcTotals carry{0, 0, 0, 0};
#pragma unroll
    for (int it = 0; it < kDcBatch; ++it) {
        const uint32_t m = w_lo + uint32_t(it) * 32u + lane;
        int v0 = s0[it], v1 = s1[it], v2 = s2[it];
        int flag = (m < w_hi && m % ri == 0) ? 1 : 0; // (m == 0 included: the image starts with every predictor at zero)
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t0 = __shfl_up_sync(0xFFFFFFFFu, v0, o), t1 = __shfl_up_sync(0xFFFFFFFFu, v1, o), t2 = __shfl_up_sync(0xFFFFFFFFu, v2, o),
                      tf = __shfl_up_sync(0xFFFFFFFFu, flag, o);
            if (lane >= uint32_t(o)) {
                if (!flag) {
                    v0 += t0;
                    v1 += t1;
                    v2 += t2;
                }
                flag |= tf;
            }
        }
        if (!flag) { // nothing started anew in this iteration up to this lane: the range's totals so far carry on
            v0 += carry.v0;
            v1 += carry.v1;
            v2 += carry.v2;
        }
        flag |= carry.flag;
        i0[it] = v0;
        i1[it] = v1;
        i2[it] = v2;
        fmask |= uint32_t(flag) << it;
        carry.v0 = __shfl_sync(0xFFFFFFFFu, v0, 31);
        carry.v1 = __shfl_sync(0xFFFFFFFFu, v1, 31);
        carry.v2 = __shfl_sync(0xFFFFFFFFu, v2, 31);
        carry.flag = __shfl_sync(0xFFFFFFFFu, flag, 31);
    }
    if (lane == 0)
        s_w[warp] = carry;
    __syncthreads();
    // the chain: what the image's MCUs in front of this chunk leave behind
    if (threadIdx.x == 0) {
        DcTotals pred{0, 0, 0, 1}; // the image starts with every predictor at zero
        if (k) {
            const volatile unsigned long long* src = b.dc_rec + 2 * size_t(cta_abs - 1u);
            unsigned long long w1;
            uint32_t spins = 0;
            while (((w1 = src[1]) >> 63) == 0) {
                __nanosleep(32);
                if (++spins > (1u << 24)) { // the predecessor is gone (it cannot be): fail the image rather than hang the device
                    atomicOr(&b.states[img].error, kErrNotConverged);
                    break;
                }
            }
            __threadfence();
            const unsigned long long w0 = src[0];
            pred = DcTotals{int(uint32_t(w0)), int(uint32_t(w0 >> 32)), int(uint32_t(w1)), int((w1 >> 32) & 1u)};
        }
        DcTotals tot = pred;
        for (int i = 0; i < NW; ++i)
            tot = dc_combine(tot, s_w[i]);
        if (k + 1 < im.dc_cta_count) {
            unsigned long long* dst = b.dc_rec + 2 * size_t(cta_abs);
            dst[0] = (unsigned long long)uint32_t(tot.v0) | ((unsigned long long)uint32_t(tot.v1) << 32);
            __threadfence();
            dst[1] = (unsigned long long)uint32_t(tot.v2) | ((unsigned long long)(0x80000000u | uint32_t(tot.flag ? 1 : 0)) << 32);
        }
        s_pred = pred;
    }
    __syncthreads();
    DcTotals outer = s_pred; // everything in front of this warp's range
    for (uint32_t i = 0; i < warp; ++i)
        outer = dc_combine(outer, s_w[i]);
#pragma unroll
    for (int it = 0; it < kDcBatch; ++it) {
        const uint32_t m = w_lo + uint32_t(it) * 32u + lane;
        if (m < w_hi) {
            const bool fresh = (fmask >> it) & 1u; // a boundary between the range's start and this MCU: nothing from outside reaches it
            int p0 = i0[it] - s0[it] + (fresh ? 0 : outer.v0), p1 = i1[it] - s1[it] + (fresh ? 0 : outer.v1), p2 = i2[it] - s2[it] + (fresh ? 0 : outer.v2);
            int16_t* pm = dc + uint64_t(m) * bpm;
            for (uint32_t j = 0; j < bpm; ++j) {
                const uint32_t c = (comp_of >> (2 * j)) & 3u;
                const int val = (c == 0 ? p0 : (c == 1 ? p1 : p2)) + pm[j];
                p0 = c == 0 ? val : p0;
                p1 = c == 1 ? val : p1;
                p2 = c == 2 ? val : p2;
                pm[j] = int16_t(val);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------------------
// IDCT: one block per thread, a CTA takes kIdctCta consecutive blocks of the stream (MCU scan order).  The block pass left,
// per block, the bit position of its first AC symbol; the CTA stages the piece of the stream its blocks occupy (a few KB,
// contiguous) in shared memory in bit order, and every thread decodes the AC symbols of one block (huffman.cuh
// decode_block_ac: the image's AC tables staged in shared memory too) straight into that block's dense 8x8 slot (row pitch
// 144 bytes: conflict free for 16-byte accesses).  The slots are then read back as the IDCT input.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kIdctPitch = 144;
#ifndef B2J_IDCT_STREAM_WORDS
#define B2J_IDCT_STREAM_WORDS (16 * B2J_IDCT_CTA)
#endif
constexpr int kIdctStreamWords = B2J_IDCT_STREAM_WORDS; // staged stream per CTA: 64 bytes per block by default; longer pieces are read from global memory
#ifndef B2J_IDCT_MIN_CTAS
#define B2J_IDCT_MIN_CTAS 3 // resident CTAs per SM the register allocation aims at
#endif
__global__ void __launch_bounds__(kIdctCta, B2J_IDCT_MIN_CTAS) k_idct(BatchDev b)
{
    extern __shared__ __align__(16) uint8_t s_dyn[];
    uint8_t* s_blk = s_dyn;                                                                // kIdctCta * kIdctPitch
    uint16_t* s_q = reinterpret_cast<uint16_t*>(s_dyn + kIdctCta * kIdctPitch);            // 3 * 64
    uint8_t* s_zz = reinterpret_cast<uint8_t*>(s_q + 3 * 64);                              // kZigzagDump
    uint16_t* s_order = reinterpret_cast<uint16_t*>(s_zz + kZigzagDump);                   // kIdctCta: decode order
    uint32_t* s_job = reinterpret_cast<uint32_t*>(s_order + kIdctCta);                     // kIdctCta x {bit position, table}
    uint32_t* s_stream = s_job + 2 * kIdctCta;                                             // kIdctStreamWords
    uint16_t* s_lut = reinterpret_cast<uint16_t*>(s_stream + kIdctStreamWords);            // the image's AC tables (b.lut_ac_slots)
    __shared__ uint32_t s_cnt[8];
    __shared__ __align__(8) uint64_t s_bar;
    if (threadIdx.x == 0)
        mbar_init(&s_bar, 1);
    __syncthreads();
    const uint32_t cta_abs = blockIdx.x + b.base_idct;
    const int img = find_image_warp(b.cta_map_idct, b.num_images, cta_abs);
    const ImageDesc& im = b.images[img];
    for (int i = threadIdx.x; i < im.ncomp_out * 64; i += kIdctCta)
        s_q[i] = b.quant[im.qt[i >> 6] * 64 + (i & 63)];
    if (threadIdx.x < kZigzagDump)
        s_zz[threadIdx.x] = zigzag_dump_entry(c_zigzag, threadIdx.x);
    if (threadIdx.x < 8)
        s_cnt[threadIdx.x] = 0;
    const uint32_t u0 = (cta_abs - b.cta_map_idct[img]) * kIdctCta, u = u0 + threadIdx.x;
    const uint32_t* words = reinterpret_cast<const uint32_t*>(b.clean + im.clean_off);
    const uint32_t max_word = (im.raw_len + 12u) >> 2;
    const uint32_t clean_bits = b.states[img].clean_len * 8u;
    const bool complete = b.states[img].done_blocks == im.total_blocks; // else the image is reported failed: nothing is decoded
    uint32_t w_first = 0, n_words = 0;
    const bool dense = b.dense_coef || im.progressive; // multi-scan streams: the scans were accumulated into dense coefficients
    const bool decode = !dense && complete;
    if (decode) {
        if (threadIdx.x == 0) { // the AC tables arrive as bulk copies while the threads stage the stream
            mbar_expect_tx(&s_bar, uint32_t(im.nac) * kLutEntries * 2u);
            for (int s = 0; s < im.nac; ++s)
                bulk_copy_g2s(s_lut + s * kLutEntries, b.huff[im.ac_pool[s]].lut, kLutEntries * 2u, &s_bar);
        }
        // the piece of the stream between the first block's AC symbols and the block after the CTA's last one
        const uint32_t* pos = b.blk_pos + im.coef_off / 64;
        const uint32_t e = min(u0 + uint32_t(kIdctCta), im.total_blocks);
        const uint32_t p_first = min(pos[u0], clean_bits), p_end = e < im.total_blocks ? min(pos[e], clean_bits) : clean_bits;
        w_first = p_first >> 5;
        n_words = p_end >= p_first ? ((p_end + 31u) >> 5) - w_first + 3u : 0u; // a window may look two words past its symbol
        if (n_words <= uint32_t(kIdctStreamWords))
            for (uint32_t i = threadIdx.x; i < n_words; i += kIdctCta)
                s_stream[i] = __byte_perm(__ldg(words + min(w_first + i, max_word)), 0, 0x0123);
    }
    IdctBlockRef ref = idct_locate(im, u);
    bool wanted = ref.comp >= 0;
    if (wanted) { // region of interest: blocks none of whose samples can reach a pixel of the region (8 pixels of margin cover
                  // the neighbours the upsampling filters look at) are not needed
        const int sx = (8 * im.hmax) >> (im.hs[ref.comp] >> 1), sy = (8 * im.vmax) >> (im.vs[ref.comp] >> 1); // block size in image pixels
        if ((ref.bx + 1) * sx <= im.roi_x - 8 || ref.bx * sx >= im.roi_x + im.roi_w + 8 || (ref.by + 1) * sy <= im.roi_y - 8 ||
            ref.by * sy >= im.roi_y + im.roi_h + 8)
            wanted = false;
    }
    uint8_t* slot = s_blk + threadIdx.x * kIdctPitch;
    if (decode)
        mbar_wait(&s_bar, 0);
    __syncthreads();
    if (dense) {
        if (!complete)
            wanted = false; // a scan failed: the image is reported failed
        if (wanted) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
                reinterpret_cast<uint4*>(slot)[i] = reinterpret_cast<const uint4*>(b.coef + im.coef_off + uint64_t(ref.src_blk) * 64)[i];
        }
    } else {
        // Every thread zeroes the slot of its own block and posts the block as a job: where its AC symbols start and which
        // table they use.  The jobs are then dealt out again by length (a counting sort over eight classes of the block's
        // bit count, longest first): the lanes of a warp decode blocks of similar length instead of waiting for the longest
        // of 32 arbitrary ones.  The thread that decoded a block goes straight on to reconstruct it: no barrier between
        // the decode and the IDCT (the warps with the short blocks would spend it waiting for the warp with the long ones).
#pragma unroll
        for (int i = 0; i < 9; ++i)
            reinterpret_cast<uint4*>(slot)[i] = make_uint4(0, 0, 0, 0);
        uint32_t cls = 0, rank = 0;
        if (wanted) {
            const uint32_t* pos = b.blk_pos + im.coef_off / 64 + ref.src_blk;
            const uint32_t p = pos[0];
            if (complete && p < clean_bits) {
                const uint32_t nxt = ref.src_blk + 1u < im.total_blocks ? pos[1] : clean_bits;
                const uint32_t bits = nxt > p ? nxt - p : 0u;
                cls = 7u - min(7u, uint32_t(32 - __clz(bits >> 4))); // 0: 1024 bits and more ... 7: under 16 bits
                rank = atomicAdd(&s_cnt[cls], 1u);
                s_job[2 * threadIdx.x] = p;
                s_job[2 * threadIdx.x + 1] = uint32_t(im.ac_slot[ref.comp]);
            } else {
                wanted = false; // a block the block pass never reached: its image is reported failed
                atomicOr(&b.states[img].error, kErrOverrun);
            }
        }
        __syncthreads();
        uint32_t total = 0;
        if (wanted) {
            uint32_t before = 0;
#pragma unroll
            for (uint32_t c = 0; c < 8; ++c) {
                before += c < cls ? s_cnt[c] : 0u;
                total += s_cnt[c];
            }
            s_order[before + rank] = uint16_t(threadIdx.x);
        } else {
#pragma unroll
            for (uint32_t c = 0; c < 8; ++c)
                total += s_cnt[c];
        }
        __syncthreads();
        wanted = threadIdx.x < total;
        if (wanted) {
            const uint32_t j = s_order[threadIdx.x];
            const uint32_t sl = s_job[2 * j + 1];
            slot = s_blk + j * kIdctPitch;
            int16_t* dst = reinterpret_cast<int16_t*>(slot);
            uint32_t err;
            if (n_words <= uint32_t(kIdctStreamWords))
                err = decode_block_ac<true>(s_stream - w_first, w_first + n_words - 1u, s_job[2 * j], s_lut, sl * kLutEntries, b.huff + im.ac_pool[sl],
                    s_zz, dst);
            else
                err = decode_block_ac<false>(words, max_word, s_job[2 * j], s_lut, sl * kLutEntries, b.huff + im.ac_pool[sl], s_zz, dst);
            if (err)
                atomicOr(&b.states[img].error, err);
            ref = idct_locate(im, u0 + j); // from here on this thread's block is the one it decoded
        }
    }
    if (!wanted)
        return;
    alignas(16) int16_t blk[64];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        reinterpret_cast<uint4*>(blk)[i] = reinterpret_cast<const uint4*>(slot)[i];
    if (b.dump_coef && !dense) { // debug tap: the dense coefficients as the reference holds them (DC slot unused)
#pragma unroll
        for (int i = 0; i < 8; ++i)
            reinterpret_cast<uint4*>(b.coef + im.coef_off + uint64_t(ref.src_blk) * 64)[i] = reinterpret_cast<const uint4*>(blk)[i];
    }
    const int dcval = dense ? int(blk[0]) : int(b.dc[im.coef_off / 64 + ref.src_blk]);
    uint32_t out[16];
    idct_block(blk, dcval, s_q + ref.comp * 64, out);
    // Horizontally subsampled planes: the colour kernel reads column ds_w as the right neighbour of the last real column
    // ds_w - 1 (the reference's edge rule is "the neighbour of the last column is the column itself", jdsample.c:276-411),
    // so when both sit in this block (ds_w % 8 != 0) column ds_w gets a copy of column ds_w - 1.
    const int dsw = im.ds_w[ref.comp];
    if (im.hs[ref.comp] < im.hmax && (dsw & 7) != 0) {
        if (ref.bx == (dsw >> 3)) {
            const int e = dsw & 7; // 1..7: copy sample e-1 to sample e in every row
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const unsigned long long v = (unsigned long long)out[2 * r] | ((unsigned long long)out[2 * r + 1] << 32);
                const unsigned long long src = (v >> (8 * (e - 1))) & 0xFFull;
                const unsigned long long w = (v & ~(0xFFull << (8 * e))) | (src << (8 * e));
                out[2 * r] = uint32_t(w);
                out[2 * r + 1] = uint32_t(w >> 32);
            }
        }
    }
    uint8_t* dst = b.planes + ref.dst;
#pragma unroll
    for (int r = 0; r < 8; ++r)
        *reinterpret_cast<uint2*>(dst + r * ref.pitch) = make_uint2(out[2 * r], out[2 * r + 1]);
}

// ------------------------------------------------------------------------------------------------------------------
// Launcher
// ------------------------------------------------------------------------------------------------------------------
int launch_unstuff(const BatchDev& b, const LaunchDims& d, cudaStream_t st)
{
    k_unstuff_count<<<d.unstuff_ctas, kUnstuffThreads, 0, st>>>(b);
    k_unstuff_scan<<<d.num_images, 32, 0, st>>>(b);
    k_unstuff_write<<<d.unstuff_ctas, kUnstuffThreads, 0, st>>>(b);
    return 3;
}

int launch_dc(const BatchDev& b, const LaunchDims& d, cudaStream_t st)
{
    k_dc_scan<<<d.dc_ctas, kDcCta, 0, st>>>(b);
    return 1;
}

static size_t idct_smem_bytes(int lut_ac_slots)
{
    return size_t(kIdctCta) * kIdctPitch + 3 * 64 * 2 + kZigzagDump + size_t(kIdctCta) * (2 + 8) + size_t(kIdctStreamWords) * 4 +
           size_t(lut_ac_slots) * kLutEntries * 2;
}

int idct_configure()
{
    return cudaFuncSetAttribute(k_idct, cudaFuncAttributeMaxDynamicSharedMemorySize, int(idct_smem_bytes(3))) == cudaSuccess ? 0 : -1;
}

int launch_idct(const BatchDev& b, const LaunchDims& d, cudaStream_t st)
{
    k_idct<<<d.idct_ctas, kIdctCta, idct_smem_bytes(b.lut_ac_slots), st>>>(b);
    return 1;
}

} // namespace b200jpeg
