// progressive_kernel.cu -- P1 k_prog_decode: the scans of the progressive images of a batch -> dense coefficients.
// One CTA of kProgWarps warps per image.  A scan is serial by nature (a refinement scan depends symbol by symbol on what is
// non-zero already), so one warp takes one scan: lane 0 walks its symbols (progressive.cuh), the other lanes move the
// blocks between global and shared memory.  But the scans of an image are not all in sequence: scans that share no
// component, or whose bands do not overlap, touch different 16-bit coefficients (ScanDesc::deps, batch_plan.cpp), so the
// warps of the CTA take the scans in file order from a counter and each waits only for the scans its own depends on
// (libjpeg's usual script: the five first scans side by side, then four refinements, then the last one).
// The images of the batch run side by side; the reconstruction kernel then reads the coefficients (ImageDesc::progressive).
#include <cuda_runtime.h>
#include "decode_kernels.h"
#include "huffman.cuh"
#include "progressive.cuh"

namespace b200jpeg {

// decode_mcu_AC_refine (ljt/jdphuff.c:506-655) on a block in shared memory, given which positions of the band are non-zero
// already (nz: bit k = zigzag position k, from a ballot of the lanes): zeros are stepped over with bit operations, so the cost
// follows the symbols and the non-zero coefficients (each takes one correction bit), not the 63 positions.  Same result as
// progressive.cuh prog_ac_refine_block, which the CPU emulation runs.
__device__ __forceinline__ bool prog_ac_refine_block_mask(ProgBits& br, const DevHuffTable* tab, int16_t* blk, unsigned long long nz, int ss,
    int se, int al, uint32_t& eobrun, const uint8_t* zz, const uint16_t* fast)
{
    const int p1 = 1 << al, m1 = -(1 << al);
    auto correct = [&](int k) { // a correction bit for the non-zero coefficient at position k
        if (br.get(1)) {
            int16_t* c = blk + zz[k];
            if ((*c & p1) == 0)
                *c = int16_t(*c + (*c >= 0 ? p1 : m1));
        }
    };
    int k = ss;
    if (eobrun == 0) {
        while (k <= se) {
            int s = prog_huff_symbol(br, tab, fast);
            if (s < 0)
                return false;
            int r = s >> 4;
            s &= 15;
            if (s) {
                s = br.get(1) ? p1 : m1;
            } else if (r != 15) {
                eobrun = 1u << r;
                if (r)
                    eobrun += br.get(r);
                break;
            }
            while (k <= se) { // over the non-zero coefficients (a correction bit each) up to the (r + 1)-th zero
                const unsigned long long rest = nz >> k;
                const int t = rest ? __ffsll((long long)rest) - 1 : se - k + 1; // zeros in front of the next non-zero one
                if (r < t) {
                    k += r;
                    r = -1;
                    break;
                }
                k += t;
                r -= t;
                if (k > se)
                    break;
                correct(k);
                ++k;
            }
            if (s && k <= se)
                blk[zz[k]] = int16_t(s);
            ++k;
        }
    }
    if (eobrun > 0) {
        unsigned long long rest = k <= 63 ? (nz >> k) << k : 0ull;
        while (rest) {
            const int pos = __ffsll((long long)rest) - 1;
            rest &= rest - 1;
            correct(pos);
        }
        --eobrun;
    }
    return true;
}

// An AC scan (one component, its blocks in raster order) by the whole warp: the 32 lanes move the block between global and
// shared memory (128 bytes, one word per lane, the next block's load in flight while this one is decoded), lane 0 walks
// the block's symbols on the shared copy (progressive.cuh: the same block functions the CPU emulation runs).  Blocks inside
// an end-of-band run of a first scan are skipped without touching memory.
__device__ uint32_t prog_ac_scan_warp(const ImageDesc& im, const ScanDesc& sc, const uint8_t* raw, const DevHuffTable* tab, int16_t* coef,
    int16_t* s_blk, const uint8_t* zz, const uint16_t* fast)
{
    const uint32_t lane = threadIdx.x & 31u;
    const int c0 = sc.comp[0];
    const uint32_t bw = uint32_t(im.ds_w[c0] + 7) / 8, bh = uint32_t(im.ds_h[c0] + 7) / 8, nblocks = bw * bh;
    ProgBits br;
    br.init(raw, sc.off, sc.off + sc.len);
    uint32_t eobrun = 0, to_go = sc.restart_interval ? uint32_t(sc.restart_interval) : 0xFFFFFFFFu;
    int next_rst = 0;
    uint32_t* s_w = reinterpret_cast<uint32_t*>(s_blk);
    uint32_t bx_n = 0, by_n = 0; // coordinates of the block after the one at hand (the blocks go in raster order)
    auto next_block_words = [&]() {
        uint32_t* g = reinterpret_cast<uint32_t*>(coef + uint64_t(prog_block_index(im, c0, int(bx_n), int(by_n))) * 64);
        if (++bx_n == bw) {
            bx_n = 0;
            ++by_n;
        }
        return g;
    };
    uint32_t* g_next = nblocks ? next_block_words() : nullptr;
    uint32_t v_next = nblocks ? g_next[lane] : 0u;
    for (uint32_t n = 0; n < nblocks; ++n) {
        uint32_t* g = g_next;
        const uint32_t v_cur = v_next;
        if (n + 1 < nblocks) {
            g_next = next_block_words();
            v_next = g_next[lane];
        }
        int action = 1; // 0: nothing to do for this block, 1: decode it, 2: error
        if (lane == 0) {
            if (to_go == 0) { // process_restart
                if (!br.restart(next_rst))
                    action = 2;
                next_rst = (next_rst + 1) & 7;
                eobrun = 0;
                to_go = uint32_t(sc.restart_interval);
            }
            --to_go;
            if (action == 1 && sc.ah == 0 && eobrun > 0) {
                --eobrun;
                action = 0;
            }
        }
        action = __shfl_sync(0xFFFFFFFFu, action, 0);
        if (action == 2)
            return kErrProgressive;
        if (action == 0)
            continue;
        s_w[lane] = v_cur;
        __syncwarp();
        unsigned long long nz = 0;
        if (sc.ah != 0) { // which positions of the band are non-zero: a ballot over the block in zigzag order
            const uint32_t lo = __ballot_sync(0xFFFFFFFFu, s_blk[zz[lane]] != 0), hi = __ballot_sync(0xFFFFFFFFu, s_blk[zz[lane + 32]] != 0);
            nz = ((unsigned long long)hi << 32) | lo;
            const unsigned long long band = (sc.se >= 63 ? ~0ull : ((1ull << (sc.se + 1)) - 1ull)) & ~((1ull << sc.ss) - 1ull);
            nz &= band;
        }
        int ok = 1;
        if (lane == 0) {
            ok = (sc.ah == 0 ? prog_ac_first_block(br, tab, s_blk, sc.ss, sc.se, sc.al, eobrun, zz, fast)
                             : prog_ac_refine_block_mask(br, tab, s_blk, nz, sc.ss, sc.se, sc.al, eobrun, zz, fast)) &&
                 !br.starved;
        }
        __syncwarp();
        // write back what changed, 16 bits at a time: the other half of a word may belong to the band of a scan that another
        // warp is decoding right now
        const uint32_t v_new = s_w[lane], diff = v_new ^ v_cur;
        if (diff & 0xFFFFu)
            reinterpret_cast<uint16_t*>(g)[2 * lane] = uint16_t(v_new);
        if (diff >> 16)
            reinterpret_cast<uint16_t*>(g)[2 * lane + 1] = uint16_t(v_new >> 16);
        if (!__shfl_sync(0xFFFFFFFFu, ok, 0))
            return kErrProgressive;
    }
    return 0;
}

// A DC refinement scan without restart markers is one bit per block in block order: lane 0 reads 32 bits, every lane sets the
// bit in its own block (the read-modify-writes of 32 blocks overlap instead of following each other).
__device__ uint32_t prog_dc_refine_warp(const ImageDesc& im, const ScanDesc& sc, const uint8_t* raw, int16_t* coef)
{
    const uint32_t lane = threadIdx.x & 31u;
    ProgBits br;
    br.init(raw, sc.off, sc.off + sc.len);
    const bool single = sc.ncomp == 1;
    const int c0 = sc.comp[0];
    const uint32_t bw = uint32_t(im.ds_w[c0] + 7) / 8, bh = uint32_t(im.ds_h[c0] + 7) / 8;
    const uint32_t nblocks = single ? bw * bh : im.total_blocks; // an interleaved scan of all components visits the blocks in MCU order
    for (uint32_t n0 = 0; n0 < nblocks; n0 += 32) {
        const uint32_t cnt = min(32u, nblocks - n0);
        uint32_t bits = 0;
        int ok = 1;
        if (lane == 0) {
            const uint32_t hi = br.get(int(min(16u, cnt)));
            const uint32_t lo = cnt > 16 ? br.get(int(cnt - 16)) : 0u;
            bits = cnt > 16 ? (hi << (cnt - 16)) | lo : hi; // the first block's bit is the most significant of the cnt bits
            ok = !br.starved;
        }
        bits = __shfl_sync(0xFFFFFFFFu, bits, 0);
        if (!__shfl_sync(0xFFFFFFFFu, ok, 0))
            return kErrProgressive;
        if (lane < cnt && ((bits >> (cnt - 1 - lane)) & 1u)) {
            const uint32_t n = n0 + lane;
            uint32_t idx = n;
            if (single) {
                const uint32_t by = n / bw, bx = n - by * bw;
                idx = prog_block_index(im, c0, int(bx), int(by));
            }
            int16_t* dc = coef + uint64_t(idx) * 64;
            *dc = int16_t(*dc | (1 << sc.al));
        }
    }
    return 0;
}

constexpr int kProgWarps = 4;       // warps per image when the batch has few progressive images (latency: 56 -> 41 ms for 64 .. 256 images)
constexpr int kProgManyImages = 512;  // per launch (a group of the batch); from this many on one warp per image: the images alone fill the device (4096 images in four groups: 93 ms against 156 ms)
struct ProgWarpSmem // what one warp needs for the scan it is decoding
{
    DevHuffTable tab;                      // the AC table of an AC scan
    uint16_t fast[3 << kProgFastBits];     // look-ahead tables: the AC table, or the DC tables of a first DC scan's components
    int16_t blk[64];
};

template <int NW>
__global__ void __launch_bounds__(32 * NW) k_prog_decode(BatchDev b, uint32_t first, uint32_t count)
{
    const uint32_t j = blockIdx.x;
    if (j >= count)
        return;
    const int img = int(b.prog_images[first + j]) - b.image_base;
    const ImageDesc& im = b.images[img];
    int16_t* coef = b.coef + im.coef_off;
    { // zero the image's coefficients: 128 bytes per block, 16 bytes per thread per step
        uint4* p = reinterpret_cast<uint4*>(coef);
        const uint64_t n16 = uint64_t(im.total_blocks) * 8;
        for (uint64_t i = threadIdx.x; i < n16; i += 32 * NW)
            p[i] = make_uint4(0, 0, 0, 0);
    }
    __shared__ __align__(16) ProgWarpSmem s_warp[NW];
    __shared__ uint8_t s_zz[64];
    __shared__ uint32_t s_next, s_done, s_ndone, s_err;
    for (uint32_t i = threadIdx.x; i < 64; i += 32 * NW) // (one warp per image when the batch holds many: NW == 1)
        s_zz[i] = c_zigzag[i];
    if (threadIdx.x == 0)
        s_next = s_done = s_ndone = s_err = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u;
    ProgWarpSmem& w = s_warp[threadIdx.x >> 5];
    const uint8_t* raw = b.raw + im.raw_off;
    for (;;) {
        // the next scan of the file nobody has taken yet; everything it depends on was taken before it, so waiting cannot dead-lock
        uint32_t s = 0;
        if (lane == 0)
            s = atomicAdd(&s_next, 1u);
        s = __shfl_sync(0xFFFFFFFFu, s, 0);
        if (s >= im.scan_count)
            break;
        const ScanDesc& sc = b.scans[im.scan_first + s];
        uint32_t stop = 0;
        if (lane == 0) {
            const uint32_t deps = sc.deps;
            const volatile uint32_t *done = &s_done, *ndone = &s_ndone, *failed = &s_err;
            if (im.scan_count > 32) {
                while (*ndone < s && !*failed)
                    __nanosleep(200);
            } else {
                while ((*done & deps) != deps && !*failed)
                    __nanosleep(200);
            }
            stop = *failed;
        }
        if (__shfl_sync(0xFFFFFFFFu, stop, 0))
            break;
        __threadfence(); // the coefficients the scans before this one wrote
        uint32_t err = 0;
        if (sc.ss != 0) {
            const uint4* src = reinterpret_cast<const uint4*>(b.huff + sc.ac_tab);
            for (uint32_t i = lane; i < sizeof(DevHuffTable) / 16; i += 32)
                reinterpret_cast<uint4*>(&w.tab)[i] = __ldg(src + i);
            __syncwarp();
            for (uint32_t e = lane; e < (1u << kProgFastBits); e += 32)
                w.fast[e] = prog_fast_entry(&w.tab, e);
            __syncwarp();
            err = prog_ac_scan_warp(im, sc, raw, &w.tab, coef, w.blk, s_zz, w.fast);
        } else if (sc.ah != 0 && sc.restart_interval == 0 && (sc.ncomp == 1 || sc.ncomp == im.ncomp)) {
            err = prog_dc_refine_warp(im, sc, raw, coef); // (every component of the frame: the scan's block order is the buffer's)
        } else {
            const int ntab = sc.ah == 0 ? sc.ncomp : 0; // a first DC scan: one DC table per component
            for (int t = 0; t < ntab; ++t)
                for (uint32_t e = lane; e < (1u << kProgFastBits); e += 32)
                    w.fast[(t << kProgFastBits) + e] = prog_fast_entry(b.huff + sc.dc_tab[t], e);
            __syncwarp();
            if (lane == 0) // first DC scans only write; the rest is rare
                err = prog_decode_scan(im, sc, raw, b.huff, coef, s_zz, ntab ? w.fast : nullptr);
            err = __shfl_sync(0xFFFFFFFFu, err, 0);
        }
        __syncwarp();
        __threadfence(); // this scan's coefficients, before it is reported complete
        if (lane == 0) {
            if (err)
                atomicOr(&s_err, err);
            atomicOr(&s_done, 1u << (s & 31u));
            atomicAdd(&s_ndone, 1u);
        }
    }
    __syncthreads();
    if (threadIdx.x != 0)
        return;
    if (s_err)
        atomicOr(&b.states[img].error, s_err);
    else
        b.states[img].done_blocks = im.total_blocks;
}

int launch_progressive(const BatchDev& b, uint32_t first, uint32_t count, cudaStream_t st)
{
    if (count == 0)
        return 0;
    if (count >= uint32_t(kProgManyImages))
        k_prog_decode<1><<<count, 32, 0, st>>>(b, first, count);
    else
        k_prog_decode<kProgWarps><<<count, 32 * kProgWarps, 0, st>>>(b, first, count);
    return 1;
}

} // namespace b200jpeg
