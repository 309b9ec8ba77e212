// huffman_kernels.cu -- sm_100a kernels of the entropy-decode stage of b200jpeg.
//
//   H1 k_huff_sync x 2   self-synchronising subsequence decode (no output): round 0 from a guessed state, round 1 from the
//                        predecessor's exit state; subsequences whose exit state changed feed a work list
//   H2 k_huff_sync_tail  cooperative kernel: Jacobi rounds over the (compacted) work list until it is empty
//   H3 k_subseq_scan     segmented scan of coefficient-slot counts -> absolute block position per subsequence (an image's chunks
//                        of 2048 subsequences hand each other their totals)
//   H* k_huff_fused      (option fused=1) H1 .. H4 in one launch; see its header below
//   H4 k_huff_write      block pass: a last decode from the agreed entry states that stores, for every block, the DC
//                        difference and the bit position of its first AC symbol (the AC symbols are decoded by the thread
//                        that reconstructs the block, decode_kernels.cu k_idct)
// Shared memory layout of the decoding kernels (dynamic): [LUTs: lut_slots x 5 KB][TabSel x 8][stream staging].
// The unstuffed stream of the CTA's subsequences is staged (byte-swapped to bit order) with coalesced loads; thread t
// owns the words [t * W, (t + 1) * W) where W = stream_words is odd, so threads at the same position of their
// subsequence hit different banks; the last W - S/32 words of a region duplicate the start of the next one.
// The per-thread decode loop lives in huffman.cuh (shared with the CPU emulation used by the tests).
#include <algorithm>
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include "decode_kernels.h"
#include "huffman.cuh"
#include "bulk_copy.cuh"

namespace cg = cooperative_groups;
#include "cta_lookup.cuh"

namespace b200jpeg {

struct HuffSmem
{
    uint16_t* lut;
    TabSel* tabsel;
    uint32_t* stream;
};

__device__ __forceinline__ HuffSmem carve_smem(const BatchDev& b, unsigned char* base, int tabsel_entries)
{
    HuffSmem s;
    s.lut = reinterpret_cast<uint16_t*>(base);
    s.tabsel = reinterpret_cast<TabSel*>(base + size_t(b.lut_slots) * kLutEntries * 2);
    s.stream = reinterpret_cast<uint32_t*>(s.tabsel + tabsel_entries);
    return s;
}

// Stage the image's distinct Huffman LUTs, its per-block table selectors and (if enabled) the stream of the CTA's
// subsequences in shared memory.
__device__ __forceinline__ void stage_cta(const BatchDev& b, const ImageDesc& im, const HuffSmem& sm, uint32_t k0)
{
    // the tables travel as bulk copies (one thread issues them, the TMA engine moves them) while all threads stage the stream
    __shared__ __align__(8) uint64_t s_bar;
    if (threadIdx.x == 0)
        mbar_init(&s_bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(&s_bar, uint32_t(im.ntabs) * kLutEntries * 2u);
        for (int s = 0; s < im.ntabs; ++s)
            bulk_copy_g2s(sm.lut + s * kLutEntries, b.huff[im.tab_pool[s]].lut, kLutEntries * 2u, &s_bar);
    }
    if (threadIdx.x < kMaxBlocksPerMcu)
        sm.tabsel[threadIdx.x] = im.tabsel_s[threadIdx.x];
    if (b.stream_words) {
        const uint32_t spw = uint32_t(b.subseq_bits) >> 5, W = uint32_t(b.stream_words), extra = W - spw;
        const uint32_t* words = reinterpret_cast<const uint32_t*>(b.clean + im.clean_off);
        const uint32_t first = k0 * spw, max_words = (im.raw_len + 16u) >> 2; // words of this image that exist
        const uint32_t total = kSubseqCta * spw + extra;
        for (uint32_t g = threadIdx.x; g < total; g += kSubseqCta) {
            const uint32_t a = first + g;
            const uint32_t v = a < max_words ? __byte_perm(words[a], 0, 0x0123) : 0u;
            const uint32_t t = g / spw, j = g - t * spw;
            if (t < kSubseqCta)
                sm.stream[t * W + j] = v;
            if (j < extra && t > 0)
                sm.stream[(t - 1) * W + spw + j] = v;
        }
    }
    mbar_wait(&s_bar, 0);
    __syncthreads();
}

// words_s: shared-window address matching words[0] when the stream is staged in shared memory (32-bit arithmetic, may
// wrap below the window for the per-thread "virtual base"), 0 otherwise
__device__ __forceinline__ HuffEnv make_env(const BatchDev& b, const ImageDesc& im, uint32_t clean_len, const uint32_t* words,
    const uint16_t* lut, const TabSel* tabsel, uint32_t words_s = 0)
{
    HuffEnv env;
    env.words = words;
    env.words_s = words_s;
    env.lut = lut;
    env.tabsel = tabsel;
    env.pool = b.huff;
    env.seg = b.seg_start + im.seg_first;
    env.num_segments = im.num_segments;
    env.seg_blocks = im.seg_blocks;
    env.total_blocks = im.total_blocks;
    env.clean_bits = clean_len * 8u;
    env.bpm = im.bpm;
    return env;
}

// Decode subsequence k of an image from entry state E in synchronisation mode and publish the result.
// Returns true if the exit state differs from the one recorded before.
template <bool PRESWAPPED, bool DIRECT = true>
__device__ __forceinline__ bool sync_decode(const BatchDev& b, const HuffEnv& env, uint32_t t, uint32_t k, unsigned long long E, bool exact,
    unsigned long long* exit_state = nullptr)
{
    HuffState s;
    s.p = uint32_t(E);
    const uint32_t slot = uint32_t(E >> 32);
    s.z = slot & 63u;
    s.blk = slot >> 6;
    s.nblk = 0;
    s.abs = exact;
    const uint32_t z_entry = s.z;
    decode_span<false, PRESWAPPED, DIRECT>(env, s, (k + 1u) * uint32_t(b.subseq_bits), nullptr);
    const unsigned long long X = pack_state(s.p, s.blk * 64u + s.z);
    const bool changed = (X != __ldcg(b.sub_exit + t));
    if (exit_state)
        *exit_state = X;
    b.sub_exit[t] = X;
    b.sub_entry[t] = E;
    // slots consumed: absolute (from the image start) once a known state was passed, else relative to the entry
    const uint32_t n = s.abs ? s.nblk * 64u + s.z : s.nblk * 64u + s.z - z_entry;
    b.sub_rec[t] = (n & 0x7FFFFFFFu) | (s.abs ? 0x80000000u : 0u);
    return changed;
}

extern __shared__ __align__(16) unsigned char g_dyn_smem[];

__global__ void __launch_bounds__(kSubseqCta) k_huff_sync(BatchDev b, int round)
{
    const uint32_t cta_abs = blockIdx.x + b.base_subseq;
    const int img = find_image_warp(b.cta_map_subseq, b.num_images, cta_abs);
    const ImageDesc& im = b.images[img];
    const HuffSmem sm = carve_smem(b, g_dyn_smem, kMaxBlocksPerMcu);
    const uint32_t k0 = (cta_abs - b.cta_map_subseq[img]) * kSubseqCta;
    stage_cta(b, im, sm, k0);
    const uint32_t k = k0 + threadIdx.x;
    const uint32_t t = im.subseq_first + k;
    const uint32_t clean_len = b.states[img].clean_len;
    const uint32_t clean_bits = clean_len * 8u;
    const uint32_t S = uint32_t(b.subseq_bits);
    const bool idle = uint64_t(k) * S >= uint64_t(clean_bits);
    const bool staged = b.stream_words != 0;
    const uint32_t* words = staged ? sm.stream + threadIdx.x * uint32_t(b.stream_words) - k * (S >> 5)
                                   : reinterpret_cast<const uint32_t*>(b.clean + im.clean_off);
    const uint32_t words_s = uint32_t(__cvta_generic_to_shared(sm.stream)) + (threadIdx.x * uint32_t(b.stream_words) - k * (S >> 5)) * 4u;
    const HuffEnv env = make_env(b, im, clean_len, words, sm.lut, sm.tabsel, words_s);
    if (round == 0) {
        b.sub_stamp[t] = 0; // not an item of any work-list round
        if (idle) {
            b.sub_exit[t] = b.sub_entry[t] = pack_state(clean_bits, 0);
            b.sub_rec[t] = 0;
            return;
        }
        b.sub_exit[t] = ~0ull;
        const unsigned long long E = pack_state(k * S, 0);
        if (staged)
            sync_decode<true>(b, env, t, k, E, k == 0);
        else
            sync_decode<false>(b, env, t, k, E, k == 0);
    } else {
        // Round 1: every subsequence again, from its predecessor's exit state.  Where that changes the exit state, the
        // successor is due once more: successors inside this CTA (whose stream is staged here already) are collected in
        // shared memory and done right away by the first threads of the CTA -- a second, compacted round without another
        // launch and another staging of the stream; only what changes then, and the successors in the next CTA, go to the
        // global work list.
        __shared__ uint32_t s_n;
        __shared__ uint16_t s_list[kSubseqCta];
        if (threadIdx.x == 0)
            s_n = 0;
        __syncthreads();
        auto push_global = [&](uint32_t tt) {
            const uint32_t i = atomicAdd(&b.sync->count[0], 1u);
            b.worklist[0][i] = tt;
            b.sub_stamp[tt] = 1u; // item of work-list round 0
        };
        if (!idle && k != 0) {
            const unsigned long long E = b.sub_exit[t - 1];
            if (E != b.sub_entry[t]) {
                const bool changed = staged ? sync_decode<true>(b, env, t, k, E, false) : sync_decode<false>(b, env, t, k, E, false);
                if (changed && uint64_t(k + 1) * S < uint64_t(clean_bits)) {
                    if (staged && threadIdx.x + 1 < kSubseqCta)
                        s_list[atomicAdd(&s_n, 1u)] = uint16_t(threadIdx.x + 1);
                    else
                        push_global(t + 1);
                }
            }
        }
        __syncthreads(); // the exit states written above are visible to the whole CTA
        if (threadIdx.x < s_n) {
            const uint32_t li = s_list[threadIdx.x], k2 = k0 + li, t2 = im.subseq_first + k2;
            const unsigned long long E2 = b.sub_exit[t2 - 1];
            if (E2 != b.sub_entry[t2]) {
                const uint32_t* w2 = sm.stream + li * uint32_t(b.stream_words) - k2 * (S >> 5);
                const uint32_t w2_s = uint32_t(__cvta_generic_to_shared(sm.stream)) + (li * uint32_t(b.stream_words) - k2 * (S >> 5)) * 4u;
                const HuffEnv env2 = make_env(b, im, clean_len, w2, sm.lut, sm.tabsel, w2_s);
                if (sync_decode<true>(b, env2, t2, k2, E2, false) && uint64_t(k2 + 1) * S < uint64_t(clean_bits))
                    push_global(t2 + 1);
            }
        }
    }
}

// Jacobi rounds over the work list until no subsequence changes any more.  Cooperative launch: every CTA is resident,
// grid.sync() separates the rounds.  Round r reads list r & 1 (length count[r % 3]) and appends to list (r + 1) & 1
// (length count[(r + 1) % 3]); count[(r + 2) % 3] is cleared for the round after.
// These rounds are latency bound (a chain of n unsynchronised subsequences takes n rounds, each as long as one thread
// needs for one subsequence), so everything the decode loop touches is staged in shared memory first: the whole LUT pool
// (when it has at most lut_slots tables), and per thread its image's table selectors and its subsequence's words.
// One Jacobi round over work list r & 1; returns false when the list was empty (nothing left to do).
__device__ __forceinline__ bool tail_round(const BatchDev& b, const HuffSmem& sm, bool pool_staged, int r, uint32_t tid, uint32_t nthreads,
    uint32_t& redone)
{
    const uint32_t S = uint32_t(b.subseq_bits), spw = S >> 5, W = uint32_t(b.stream_words);
    TabSel* my_tabsel = sm.tabsel + threadIdx.x * kMaxBlocksPerMcu;
    uint32_t* my_words = sm.stream + threadIdx.x * W;
    const uint32_t n = __ldcg(&b.sync->count[r % 3]);
    if (n == 0)
        return false;
    if (tid == 0)
        b.sync->count[(r + 2) % 3] = 0;
    const uint32_t* in = b.worklist[r & 1];
    uint32_t* out = b.worklist[(r + 1) & 1];
    const bool run_ahead = r >= kRunAheadFrom;
    for (uint32_t i = tid; i < n; i += nthreads) {
        uint32_t t = __ldcg(in + i);
        // Run-ahead rounds: only the head of a run of consecutive listed subsequences is worked on -- by one thread that walks
        // the whole run (every item of it is due) and on into the subsequences behind it for as long as exit states keep
        // changing.  The items behind the head are left to that thread.
        if (run_ahead && __ldcg(b.sub_stamp + t - 1) >= uint32_t(r) + 1u)
            continue;
        const int img = int(__ldg(b.subseq_cta_img + t / kSubseqCta)) - b.image_base;
        const ImageDesc& im = b.images[img];
        const uint32_t clean_len = b.states[img].clean_len;
        const uint32_t* gwords = reinterpret_cast<const uint32_t*>(b.clean + im.clean_off);
        const uint16_t* lut = reinterpret_cast<const uint16_t*>(b.huff);
        const TabSel* tabsel = im.tabsel_g;
        if (pool_staged) { // selectors relative to the staged pool: table i sits at i * kLutEntries
            for (int j = 0; j < im.bpm; ++j) {
                const TabSel g = im.tabsel_g[j];
                my_tabsel[j] = TabSel{g.dc_idx * kLutEntries, g.ac_idx * kLutEntries, g.dc_idx, g.ac_idx};
            }
            lut = sm.lut;
            tabsel = my_tabsel;
        }
        unsigned long long E = __ldcg(b.sub_exit + t - 1); // written by other SMs in earlier rounds: bypass L1
        bool prev_shift = false; // the subsequence before t was done by this thread and moved the state by exactly S bits
        bool in_run = true;      // t belongs to the run of listed items this thread started with
        for (;;) {
            const uint32_t k = t - im.subseq_first;
            bool changed = false;
            unsigned long long X;
            if (E == __ldcg(b.sub_entry + t)) {
                X = __ldcg(b.sub_exit + t); // already decoded from this state
            } else if (prev_shift && im.num_segments == 1 && uint64_t(k + 1) * S + 64u <= uint64_t(clean_len) * 8u &&
                       subseq_repeats(gwords, (k - 1u) * spw, spw)) {
                // periodic stretch (huffman.cuh subseq_repeats): the same decode as the one before, S bits further on
                X = E + S;
                changed = (X != __ldcg(b.sub_exit + t));
                b.sub_exit[t] = X;
                b.sub_entry[t] = E;
                b.sub_rec[t] = b.sub_rec[t - 1];
                ++redone;
            } else if (W) {
                // the subsequence starts on a 16-byte boundary of the image's stream (which is readable 16 bytes past its
                // end): fetch it with independent 16-byte loads first, then byte-swap into the thread's staging words
                const uint32_t first = k * spw;
                const uint4* src = reinterpret_cast<const uint4*>(gwords + first);
                constexpr int kMaxVec = 17; // (256 B / 4 + 3 | 1) words
                uint4 v[kMaxVec];
                const int nvec = int(W + 3u) >> 2;
                const int avail = int((((im.raw_len + 16u) >> 2) - first + 3u) >> 2); // 16-byte groups inside the stream buffer
#pragma unroll
                for (int j = 0; j < kMaxVec; ++j)
                    if (j < nvec)
                        v[j] = j < avail ? __ldcg(src + j) : make_uint4(0, 0, 0, 0);
#pragma unroll
                for (int j = 0; j < kMaxVec; ++j)
                    if (j < nvec) {
                        const uint32_t w4[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            if (uint32_t(4 * j + q) < W)
                                my_words[4 * j + q] = __byte_perm(w4[q], 0, 0x0123);
                    }
                const HuffEnv env = make_env(b, im, clean_len, my_words - first, lut, tabsel,
                    uint32_t(__cvta_generic_to_shared(my_words)) - first * 4u);
                changed = sync_decode<true, true>(b, env, t, k, E, false, &X);
                ++redone;
            } else {
                const HuffEnv env = make_env(b, im, clean_len, gwords, lut, tabsel);
                changed = sync_decode<false>(b, env, t, k, E, false, &X);
                ++redone;
            }
            if (uint64_t(k + 1) * S >= uint64_t(clean_len) * 8u)
                break; // no subsequence behind this one
            if (!run_ahead) { // Jacobi: a changed exit state makes the successor an item of the next round
                if (changed) {
                    out[atomicAdd(&b.sync->count[(r + 1) % 3], 1u)] = t + 1;
                    b.sub_stamp[t + 1] = uint32_t(r) + 2u;
                }
                break;
            }
            const bool next_listed = __ldcg(b.sub_stamp + t + 1) >= uint32_t(r) + 1u;
            if (next_listed && !in_run) { // the head of another run: its thread may have read the old exit state
                if (changed) {
                    out[atomicAdd(&b.sync->count[(r + 1) % 3], 1u)] = t + 1;
                    b.sub_stamp[t + 1] = uint32_t(r) + 2u;
                }
                break;
            }
            if (!next_listed) {
                if (!changed)
                    break;
                in_run = false;
            }
            prev_shift = (X == E + S);
            E = X;
            ++t;
        }
    }
    return true;
}

// Work-list position a thread takes first.  The rounds are latency bound and their lists shrink geometrically, so the
// items are dealt lane-major (tail_mode bit 0): item i goes to lane i / (warps of the grid), the warps taken CTA after
// CTA.  A list shorter than the grid's warp count then leaves one decoding lane per warp, which does not wait for the
// block ends and long codes of diverging neighbours (measured: 324 -> 294 us for the 7 rounds of 64 1080p images).
__device__ __forceinline__ uint32_t tail_slot(const BatchDev& b)
{
    if (!(b.tail_mode & 1))
        return blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    return lane * (gridDim.x * (blockDim.x >> 5)) + warp * gridDim.x + blockIdx.x;
}

__device__ __forceinline__ bool tail_stage_pool(const BatchDev& b, const HuffSmem& sm)
{
    const bool pool_staged = b.num_tables <= b.lut_slots;
    __shared__ __align__(8) uint64_t s_bar;
    if (threadIdx.x == 0)
        mbar_init(&s_bar, 1);
    __syncthreads();
    if (pool_staged) {
        if (threadIdx.x == 0) {
            mbar_expect_tx(&s_bar, uint32_t(b.num_tables) * kLutEntries * 2u);
            for (int s = 0; s < b.num_tables; ++s)
                bulk_copy_g2s(sm.lut + s * kLutEntries, b.huff[s].lut, kLutEntries * 2u, &s_bar);
        }
        mbar_wait(&s_bar, 0);
    }
    __syncthreads();
    return pool_staged;
}

__global__ void __launch_bounds__(kTailCta) k_huff_sync_tail(BatchDev b, int first_round)
{
    cg::grid_group grid = cg::this_grid();
    if (__ldcg(&b.sync->count[first_round % 3]) == 0) // every thread reads the same value: nothing left, no barrier needed
        return;
    const HuffSmem sm = carve_smem(b, g_dyn_smem, kTailCta * kMaxBlocksPerMcu);
    const bool pool_staged = tail_stage_pool(b, sm);
    const uint32_t nthreads = gridDim.x * blockDim.x, tid = tail_slot(b);
    uint32_t redone = 0;
    int r = first_round;
    unsigned long long t0 = 0;
    if (tid == 0)
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;; ++r) { // until the list is empty: at most one round per subsequence of the longest unsynchronised chain
        if (tid == 0 && r < kTraceRounds)
            b.sync->round_items[r] = __ldcg(&b.sync->count[r % 3]);
        if (!tail_round(b, sm, pool_staged, r, tid, nthreads, redone))
            break;
        grid.sync();
        if (tid == 0 && r < kTraceRounds) {
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            b.sync->round_ns[r] = uint32_t(t1 - t0);
        }
    }
    if (redone)
        atomicAdd(&b.sync->redone, redone);
    if (tid == 0)
        b.sync->rounds = uint32_t(r);
}

// The same round as an ordinary launch (option tail_launches): no co-residency requirement, so the round's few CTAs
// interleave freely with the other groups' kernels; the kernel boundary is the barrier between rounds.
// Its grid is not bound by co-residency: it has a thread for every subsequence of the group, dealt thread-major, so a long
// list (the first one: a quarter of all subsequences) is done in one pass of dense warps instead of two passes of the
// cooperative grid; CTAs past the end of the list leave at once.
__global__ void __launch_bounds__(kTailCta) k_huff_tail_round(BatchDev b, int r)
{
    const uint32_t n = __ldcg(&b.sync->count[r % 3]);
    if (n == 0 && blockIdx.x == 0 && threadIdx.x == 0)
        b.sync->count[(r + 2) % 3] = 0; // keep the rotation going: a launch after the last round must not find an old length
    if (blockIdx.x * blockDim.x >= n)
        return;
    const HuffSmem sm = carve_smem(b, g_dyn_smem, kTailCta * kMaxBlocksPerMcu);
    const bool pool_staged = tail_stage_pool(b, sm);
    const uint32_t nthreads = gridDim.x * blockDim.x, tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t redone = 0;
    tail_round(b, sm, pool_staged, r, tid, nthreads, redone);
    if (redone)
        atomicAdd(&b.sync->redone, redone);
    if (tid == 0)
        b.sync->rounds = uint32_t(r + 1);
}

// segmented-scan combine: a = earlier prefix, x = later element; bit31 = "absolute" flag
__device__ __forceinline__ uint32_t seg_combine(uint32_t a, uint32_t x) { return (x & 0x80000000u) ? x : a + x; }

// An image's subsequences are cut into chunks of kScanChunk, one CTA each.  Every warp takes a contiguous range of 256: all
// its loads first, then a warp-level segmented scan 32 at a time with the totals carried along; one barrier combines the
// warps; the chunks of an image form a chain (scan_rec: one word per chunk, CTAs in ticket order, as in k_dc_scan).  (One CTA
// per image with a barrier-synchronised scan per 256 subsequences sat on the critical path of its group: 26 us per 1080p
// image, and in proportion more for big ones.)
constexpr int kScanThreads = 256;
constexpr int kScanBatch = kScanChunk / kScanThreads;
__global__ void __launch_bounds__(kScanThreads) k_subseq_scan(BatchDev b)
{
    constexpr uint32_t NW = kScanThreads / 32;
    __shared__ uint32_t s_w[NW];
    __shared__ uint32_t s_pred, s_ticket;
    if (threadIdx.x == 0)
        s_ticket = atomicAdd(&b.sync->ticket_scan, 1u);
    __syncthreads();
    const uint32_t cta_abs = s_ticket + b.base_scan;
    const int img = find_image_warp(b.cta_map_scan, b.num_images, cta_abs);
    const ImageDesc& im = b.images[img];
    const uint32_t k = cta_abs - b.cta_map_scan[img];
    const uint32_t count = im.subseq_cta_count * kSubseqCta;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lo = min(k * uint32_t(kScanChunk) + warp * 32u * kScanBatch, count), hi = min(lo + 32u * kScanBatch, count);
    const uint32_t* __restrict__ rec = b.sub_rec + im.subseq_first;
    const unsigned long long* __restrict__ entry = b.sub_entry + im.subseq_first;
    const unsigned long long* __restrict__ exitst = b.sub_exit + im.subseq_first;
    const uint64_t live_bits = uint64_t(b.states[img].clean_len) * 8u;
    uint32_t v[kScanBatch];
    bool bad = false;
#pragma unroll
    for (int it = 0; it < kScanBatch; ++it) {
        const uint32_t c = lo + uint32_t(it) * 32u + lane;
        v[it] = c < hi ? __ldg(rec + c) : 0u;
        if (c < hi && c > 0 && __ldg(entry + c) != __ldg(exitst + c - 1) && uint64_t(c) * uint32_t(b.subseq_bits) < live_bits)
            bad = true; // the chain of entry/exit states is not consistent: synchronisation did not finish
    }
    uint32_t carry = 0; // identity: no absolute flag, nothing consumed (warp-uniform)
#pragma unroll
    for (int it = 0; it < kScanBatch; ++it) {
        uint32_t x = v[it];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t tprev = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= uint32_t(o))
                x = seg_combine(tprev, x);
        }
        x = seg_combine(carry, x);
        v[it] = x; // relative to the start of the warp's range unless an absolute count was passed
        carry = __shfl_sync(0xFFFFFFFFu, x, 31);
    }
    if (lane == 0)
        s_w[warp] = carry;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t pred = 0x80000000u; // the image starts at an absolute zero
        if (k) {
            const volatile unsigned long long* src = b.scan_rec + (cta_abs - 1u);
            unsigned long long w;
            uint32_t spins = 0;
            while (((w = *src) >> 63) == 0) {
                __nanosleep(32);
                if (++spins > (1u << 24)) { // the predecessor is gone (it cannot be): fail the image rather than hang the device
                    atomicOr(&b.states[img].error, kErrNotConverged);
                    break;
                }
            }
            pred = uint32_t(w) | 0x80000000u;
        }
        uint32_t tot = pred;
        for (uint32_t i = 0; i < NW; ++i)
            tot = seg_combine(tot, s_w[i]);
        if (k + 1 < im.scan_cta_count)
            b.scan_rec[cta_abs] = (unsigned long long)tot | (1ull << 63);
        s_pred = pred;
    }
    __syncthreads();
    uint32_t pre = s_pred;
    for (uint32_t i = 0; i < warp; ++i)
        pre = seg_combine(pre, s_w[i]);
    uint32_t* __restrict__ abs_out = b.sub_abs + im.subseq_first;
#pragma unroll
    for (int it = 0; it < kScanBatch; ++it) {
        const uint32_t c = lo + uint32_t(it) * 32u + lane;
        if (c < hi)
            abs_out[c] = seg_combine(pre, v[it]);
    }
    if (bad)
        atomicOr(&b.states[img].error, kErrNotConverged);
}

__global__ void __launch_bounds__(kSubseqCta) k_huff_write(BatchDev b)
{
    const uint32_t cta_abs = blockIdx.x + b.base_subseq;
    const int img = find_image_warp(b.cta_map_subseq, b.num_images, cta_abs);
    const ImageDesc& im = b.images[img];
    const HuffSmem sm = carve_smem(b, g_dyn_smem, kMaxBlocksPerMcu);
    const uint32_t cta = cta_abs - b.cta_map_subseq[img];
    const uint32_t k0 = cta * kSubseqCta, k = k0 + threadIdx.x;
    stage_cta(b, im, sm, k0);
    const uint32_t t = im.subseq_first + k;
    const uint32_t clean_len = b.states[img].clean_len;
    const uint32_t clean_bits = clean_len * 8u;
    const uint32_t S = uint32_t(b.subseq_bits);
    if (uint64_t(k) * S >= uint64_t(clean_bits))
        return;
    HuffState s;
    if (k == 0) {
        s.p = 0;
        s.z = 0;
        s.blk = 0;
        s.nblk = 0;
    } else {
        const unsigned long long E = b.sub_exit[t - 1];
        const uint32_t slot = uint32_t(E >> 32);
        s.p = uint32_t(E);
        s.z = slot & 63u;
        s.blk = slot >> 6;
        s.nblk = (b.sub_abs[t - 1] & 0x7FFFFFFFu) >> 6;
    }
    s.abs = true;
    BlkOut out;
    out.dc = b.dc + im.coef_off / 64;
    out.pos = b.blk_pos + im.coef_off / 64;
    const uint32_t nblk_in = s.nblk;
    uint32_t err;
    if (b.stream_words) {
        const HuffEnv env = make_env(b, im, clean_len, sm.stream + threadIdx.x * uint32_t(b.stream_words) - k * (S >> 5), sm.lut, sm.tabsel,
            uint32_t(__cvta_generic_to_shared(sm.stream)) + (threadIdx.x * uint32_t(b.stream_words) - k * (S >> 5)) * 4u);
        err = decode_span<true, true>(env, s, (k + 1u) * S, &out);
    } else {
        const HuffEnv env = make_env(b, im, clean_len, reinterpret_cast<const uint32_t*>(b.clean + im.clean_off), sm.lut, sm.tabsel);
        err = decode_span<true, false>(env, s, (k + 1u) * S, &out);
    }
    if (s.nblk == im.total_blocks && s.z == 0 && nblk_in < im.total_blocks)
        b.states[img].done_blocks = s.nblk; // this thread completed the last block
    if (err)
        atomicOr(&b.states[img].error, err);
}

// ------------------------------------------------------------------------------------------------------------------
// H* k_huff_fused: the whole entropy-decode stage of a group of images in ONE launch (default; the kernels above are the
// `fused=0` path).  A CTA owns kFusedOwn consecutive subsequences of an image and also decodes the kFusedHalo subsequences
// before them (its halo).  It stages that piece of the stream once and then runs, entirely in shared memory,
//   round 0   every subsequence from a guessed state (thread per subsequence)
//   round 1   every subsequence from its predecessor's exit state
//   rounds 2+ the subsequences whose predecessor's exit state changed, compacted onto the first threads; a thread walks on
//             into the following subsequences for as long as exit states keep changing and no other thread of the round
//             starts there (a stretch whose words repeat the one before is not decoded at all: periodic streams of flat
//             image areas never lock on, huffman.cuh subseq_repeats), until no exit state changes any more.
// The halo makes the state at the start of the own range right without any word from the CTA before (an unsynchronised
// chain dies out geometrically: 8 subsequences leave about 3 of 10^5 CTAs wrong).  "About" is not "always", so the CTAs of an
// image form a chain: each publishes the exit state of its last own subsequence together with the number of coefficient
// slots consumed up to there (one 64-bit word, fused_rec), its successor compares that state with what its halo delivered
// and, if they differ, walks its own range again from the true state before it publishes in turn.  CTAs take their
// place in the chain from a ticket counter, so a CTA only ever waits for CTAs that were started before it.
// With the slot count of its predecessor a CTA knows the absolute block index at every subsequence start (segmented scan
// inside the CTA) and finishes with the block pass (DC difference and bit position of the first AC symbol per block): the
// per-subsequence arrays, the work lists, the cooperative kernel, k_subseq_scan and k_huff_write are not needed.
// Shared memory (dynamic): [LUTs][TabSel x 10][stream: 256 x W words][exit state x 256][slot count x 256][2 lists x 256 B].
// ------------------------------------------------------------------------------------------------------------------
struct FusedSmem
{
    uint16_t* lut;
    TabSel* tabsel;
    uint32_t* stream;
    uint32_t* exit;  // pack_rel per subsequence of the CTA
    uint32_t* rec;   // coefficient slots consumed by the last decode of that subsequence; bit 31: absolute
    uint8_t* list[2];
};

__host__ __device__ inline uint32_t fused_stream_words(int subseq_bits) { return (uint32_t(subseq_bits) / 32u + 1u) | 1u; }

static size_t fused_smem_bytes(const BatchDev& b)
{
    return size_t(b.lut_slots) * kLutEntries * 2 + kMaxBlocksPerMcu * sizeof(TabSel) + (size_t(kSubseqCta) * fused_stream_words(b.subseq_bits) + 4) * 4 +
           size_t(kSubseqCta) * (4 + 4 + 2);
}

#if defined(B2J_FUSED_TRACE) // developer build (tools/build_variant.sh): %globaltimer stamps of every CTA's phases in the (otherwise unused) sub_exit array
#define B2J_STAMP(j)                                                          \
    do {                                                                      \
        if (threadIdx.x == 0) {                                               \
            unsigned long long t__;                                           \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t__));           \
            b.sub_exit[size_t(trace_cta) * 8 + (j)] = t__;                    \
        }                                                                     \
    } while (0)
#else
#define B2J_STAMP(j) do { } while (0)
#endif

__global__ void __launch_bounds__(kSubseqCta, 4) k_huff_fused(BatchDev b)
{
    __shared__ uint32_t s_ticket, s_cnt[2], s_start[2][kSubseqCta / 32], s_warp[kSubseqCta / 32], s_agg, s_ok;
    __shared__ unsigned long long s_pred;
    __shared__ __align__(8) uint64_t s_bar;
    const uint32_t i = threadIdx.x, lane = i & 31u, warp = i >> 5;
    if (i == 0) {
        s_ticket = atomicAdd(&b.sync->ticket, 1u);
        s_cnt[0] = s_cnt[1] = 0;
        mbar_init(&s_bar, 1);
    }
    if (i < 2 * (kSubseqCta / 32))
        (&s_start[0][0])[i] = 0;
    __syncthreads();
    const uint32_t cta_abs = s_ticket + b.base_fused;
#if defined(B2J_FUSED_TRACE)
    const uint32_t trace_cta = cta_abs;
#endif
    B2J_STAMP(0);
    const int img = find_image_warp(b.cta_map_fused, b.num_images, cta_abs);
    const ImageDesc& im = b.images[img];
    const uint32_t c = cta_abs - b.cta_map_fused[img];
    const uint32_t h = c ? uint32_t(kFusedHalo) : 0u;
    const uint32_t k0 = c * uint32_t(kFusedOwn) - h;
    const uint32_t clean_len = b.states[img].clean_len, clean_bits = clean_len * 8u;
    const uint32_t S = uint32_t(b.subseq_bits), spw = S >> 5, W = fused_stream_words(b.subseq_bits);
    const uint32_t live = uint32_t((uint64_t(clean_bits) + S - 1u) / S); // subsequences of the image that hold data
    const uint32_t nv = live > k0 ? min(uint32_t(kFusedOwn) + h, live - k0) : 0u; // threads of this CTA with a subsequence
    if (nv <= h)
        return; // nothing of its own (the plan counts subsequences of the stuffed length); the same holds for every later CTA of the image

    FusedSmem sm;
    {
        unsigned char* base = g_dyn_smem;
        sm.lut = reinterpret_cast<uint16_t*>(base);
        sm.tabsel = reinterpret_cast<TabSel*>(base + size_t(b.lut_slots) * kLutEntries * 2);
        sm.stream = reinterpret_cast<uint32_t*>(sm.tabsel + kMaxBlocksPerMcu);
        sm.exit = sm.stream + size_t(kSubseqCta) * W + 4;
        sm.rec = sm.exit + kSubseqCta;
        sm.list[0] = reinterpret_cast<uint8_t*>(sm.rec + kSubseqCta);
        sm.list[1] = sm.list[0] + kSubseqCta;
    }
    // ---- staging: tables by bulk copy, the stream of subsequences [k0, k0 + 256) byte-swapped to bit order; thread t's region
    // holds its subsequence and the first word(s) of the next (a symbol that starts in the last word reads one word on)
    if (i == 0) {
        mbar_expect_tx(&s_bar, uint32_t(im.ntabs) * kLutEntries * 2u);
        for (int s = 0; s < im.ntabs; ++s)
            bulk_copy_g2s(sm.lut + s * kLutEntries, b.huff[im.tab_pool[s]].lut, kLutEntries * 2u, &s_bar);
    }
    if (i < kMaxBlocksPerMcu)
        sm.tabsel[i] = im.tabsel_s[i];
    {
        const uint32_t extra = W - spw;
        const uint32_t* words = reinterpret_cast<const uint32_t*>(b.clean + im.clean_off);
        const uint32_t first = k0 * spw, max_words = (im.raw_len + 16u) >> 2;
        const uint32_t total = kSubseqCta * spw + extra;
        for (uint32_t g = i; g < total; g += kSubseqCta) {
            const uint32_t a = first + g;
            const uint32_t v = a < max_words ? __byte_perm(words[a], 0, 0x0123) : 0u;
            const uint32_t t = g / spw, j = g - t * spw;
            if (t < kSubseqCta)
                sm.stream[t * W + j] = v;
            if (j < extra && t > 0)
                sm.stream[(t - 1) * W + spw + j] = v;
        }
        if (i < 4)
            sm.stream[size_t(kSubseqCta) * W + i] = 0;
    }
    sm.exit[i] = 0xFFFFFFFFu;
    sm.rec[i] = 0;
    mbar_wait(&s_bar, 0);
    __syncthreads();
    B2J_STAMP(1);

    const uint32_t stream_s = uint32_t(__cvta_generic_to_shared(sm.stream));
    HuffEnv env = make_env(b, im, clean_len, nullptr, sm.lut, sm.tabsel, 0);
    auto point_env = [&](uint32_t li) { // the staged region of list index li, addressed as if the whole stream were there
        const uint32_t off = li * W - (k0 + li) * spw;
        env.words = sm.stream + off;
        env.words_s = stream_s + off * 4u;
    };
    // word j (0 .. spw + 1) of the subsequence at list index li, li <= 254
    auto word_of = [&](uint32_t li, uint32_t j) { return j <= spw ? sm.stream[li * W + j] : sm.stream[(li + 1u) * W + j - spw]; };

    // ---- synchronisation rounds ----
    uint32_t n_cur = 0, cur = 0, redone = 0, round = 0, inc = 0, exc = 0;
    const bool own = i >= h && i < nv;
    unsigned long long pred = 0;
    for (;; ++round) {
        int li = -1;
        if (round == 0)
            li = i < nv ? int(i) : -1;
        else if (round == 1)
            li = (i < nv && i > 0) ? int(i) : -1; // i == 0: the image's first subsequence (state known) or the first of the halo (keeps its guess)
        else if (i < n_cur)
            li = sm.list[cur][i];
        uint32_t E = 0;
        if (li > 0)
            E = sm.exit[li - 1];
        if (round >= 2 && li >= 0)
            atomicOr(&s_start[round & 1][li >> 5], 1u << (li & 31));
        __syncthreads(); // every entry state is read before any exit state changes: the rounds are deterministic
        // Late rounds: when a whole run of consecutive subsequences is still listed, decoding its members side by side is
        // wasted (each will change again when the state from the run's head arrives): only the head's thread works and
        // walks the run, the others leave their items to it.
        const bool run_heads = round >= uint32_t(kFusedRunHeads);
        if (run_heads && li > 0 && ((s_start[round & 1][(li - 1) >> 5] >> ((li - 1) & 31)) & 1u))
            li = -1;
        if (li >= 0) {
            bool prev_shift = false, in_run = true;
            for (;;) {
                const uint32_t kk = k0 + uint32_t(li);
                uint32_t X;
                if (prev_shift && im.num_segments == 1 && uint32_t(li) + 2u < uint32_t(kSubseqCta) && uint64_t(kk + 1) * S + 64u <= uint64_t(clean_bits)) {
                    bool same = true;
                    for (uint32_t j = 0; j < spw + 2u && same; ++j)
                        same = word_of(uint32_t(li) - 1u, j) == word_of(uint32_t(li), j);
                    prev_shift = same;
                }
                else
                    prev_shift = false;
                bool changed;
                if (prev_shift) { // the same words from the same relative state: the same decode, S bits further on
                    X = E;
                    changed = X != sm.exit[li];
                    sm.exit[li] = X;
                    sm.rec[li] = sm.rec[li - 1];
                } else {
                    HuffState s;
                    if (round == 0) {
                        s.p = kk * S;
                        s.z = 0;
                        s.blk = 0;
                        s.abs = kk == 0;
                    } else {
                        s.p = rel_pos(E, kk - 1u, S);
                        s.z = (E >> 16) & 63u;
                        s.blk = E >> 22;
                        s.abs = false;
                    }
                    s.nblk = 0;
                    const uint32_t z_entry = s.z;
                    point_env(uint32_t(li));
                    if (round < 2) // every lane of every warp decodes: fewest instructions (two stream words fetched per symbol)
                        decode_span<false, true, true>(env, s, (kk + 1u) * S, nullptr);
                    else // a few lanes decode and everything else waits for them: shortest dependency chain (sliding window)
                        decode_span<false, true, false>(env, s, (kk + 1u) * S, nullptr);
                    X = pack_rel(s.p, kk, S, s.blk * 64u + s.z);
                    changed = X != sm.exit[li];
                    sm.exit[li] = X;
                    const uint32_t n = s.abs ? s.nblk * 64u + s.z : s.nblk * 64u + s.z - z_entry;
                    sm.rec[li] = (n & 0x7FFFFFFFu) | (s.abs ? 0x80000000u : 0u);
                }
                if (round == 0)
                    break;
                if (round >= 2)
                    ++redone;
                if (uint32_t(li) + 1u >= nv)
                    break;
                const bool next_listed = round >= 2 && ((s_start[round & 1][(li + 1) >> 5] >> ((li + 1) & 31)) & 1u);
                bool push = false, stop;
                if (round == 1) { // a plain Jacobi round
                    push = changed;
                    stop = true;
                } else if (!run_heads) { // a walk ends where nothing changes any more, or where another thread's item begins
                    push = changed && next_listed;
                    stop = !changed || next_listed;
                } else if (next_listed && !in_run) { // the head of another run: its thread read the old exit state
                    push = changed;
                    stop = true;
                } else {
                    stop = !next_listed && !changed;
                    if (!next_listed)
                        in_run = false;
                }
                if (push)
                    sm.list[cur ^ 1u][atomicAdd(&s_cnt[cur ^ 1u], 1u)] = uint8_t(li + 1); // an item of the next round
                if (stop)
                    break;
                prev_shift = (X == E); // moved by exactly one subsequence length
                E = X;
                ++li;
            }
        }
        __syncthreads();
        if (round < 3)
            B2J_STAMP(2 + round);
        const uint32_t n_next = s_cnt[cur ^ 1u];
        __syncthreads();
        if (i == 0)
            s_cnt[cur] = 0;
        if (i < kSubseqCta / 32)
            s_start[round & 1][i] = 0;
        cur ^= 1u;
        n_cur = n_next;
        if (round >= 1 && n_cur == 0) {
            // ---- nothing changes any more.  Slots consumed up to every subsequence start of the own range: segmented
            // inclusive scan, first without the predecessor's count so that the link to the next CTA is short ----
            const uint32_t mine = own ? sm.rec[i] : 0u;
            inc = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t tprev = __shfl_up_sync(0xFFFFFFFFu, inc, o);
                if (lane >= uint32_t(o))
                    inc = seg_combine(tprev, inc);
            }
            if (lane == 31)
                s_warp[warp] = inc;
            __syncthreads();
            uint32_t pre = 0;
            for (uint32_t w = 0; w < warp; ++w)
                pre = seg_combine(pre, s_warp[w]);
            inc = seg_combine(pre, inc);
            exc = __shfl_up_sync(0xFFFFFFFFu, inc, 1);
            if (lane == 0)
                exc = pre;
            if (i == nv - 1u)
                s_agg = inc;
            __syncthreads();
            B2J_STAMP(5);
            // ---- the link: thread 0 waits for the word of the CTA before this one in the image, compares that CTA's last
            // exit state with what the halo delivered and, if they agree, publishes this CTA's word at once ----
            if (i == 0) {
                unsigned long long v = 1ull << 63; // the first CTA of an image starts at an absolute zero
                bool ok = true;
                if (c) {
                    const volatile unsigned long long* src = b.fused_rec + (cta_abs - 1u);
                    uint32_t spins = 0;
                    while (((v = *src) >> 63) == 0) {
                        __nanosleep(32);
                        if (++spins > (1u << 24)) { // about a second: the predecessor is gone (it cannot be, but never hang a device over it)
                            atomicOr(&b.states[img].error, kErrNotConverged);
                            v = (1ull << 63) | sm.exit[h - 1u];
                            break;
                        }
                    }
                    ok = uint32_t(v) == sm.exit[h - 1u];
                }
                if (ok)
                    b.fused_rec[cta_abs] = (unsigned long long)sm.exit[nv - 1u] | ((unsigned long long)(seg_combine(uint32_t(v >> 32), s_agg) | 0x80000000u) << 32);
                else { // the halo was too short for this stretch of the stream: walk the own range again from the true state
                    sm.exit[h - 1u] = uint32_t(v);
                    sm.list[cur][0] = uint8_t(h);
                    atomicAdd(&b.sync->chain_redo, 1u);
                }
                s_pred = v;
                s_ok = ok ? 1u : 0u;
            }
            __syncthreads();
            pred = s_pred;
            if (s_ok)
                break;
            n_cur = 1;
        }
    }
    exc = seg_combine(uint32_t(pred >> 32), exc); // bit 31 of a published count is set: absolute
    B2J_STAMP(6);
    if (i == 0)
        atomicMax(&b.sync->rounds, round >= 1u ? round - 1u : 0u);
    if (redone)
        atomicAdd(&b.sync->redone, redone);
    if (!own)
        return;

    // ---- block pass: per block the DC difference and the bit position of its first AC symbol ----
    const uint32_t k = k0 + i;
    HuffState s;
    if (k == 0) {
        s.p = 0;
        s.z = 0;
        s.blk = 0;
        s.nblk = 0;
    } else {
        const uint32_t E = sm.exit[i - 1u];
        s.p = rel_pos(E, k - 1u, S);
        s.z = (E >> 16) & 63u;
        s.blk = E >> 22;
        s.nblk = (exc & 0x7FFFFFFFu) >> 6;
    }
    s.abs = true;
    BlkOut out;
    out.dc = b.dc + im.coef_off / 64;
    out.pos = b.blk_pos + im.coef_off / 64;
    const uint32_t nblk_in = s.nblk;
    point_env(i);
    const uint32_t err = decode_span<true, true>(env, s, (k + 1u) * S, &out);
    if (s.nblk == im.total_blocks && s.z == 0 && nblk_in < im.total_blocks)
        b.states[img].done_blocks = s.nblk; // this thread completed the last block
    if (err)
        atomicOr(&b.states[img].error, err);
#if defined(B2J_FUSED_TRACE)
    if (i == h) {
        unsigned long long t__;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t__));
        b.sub_exit[size_t(trace_cta) * 8 + 7] = t__;
    }
#endif
}

// ------------------------------------------------------------------------------------------------------------------
// Launchers
// ------------------------------------------------------------------------------------------------------------------
static size_t main_smem_bytes(const BatchDev& b) { return size_t(b.lut_slots) * kLutEntries * 2 + kMaxBlocksPerMcu * sizeof(TabSel) + size_t(kSubseqCta) * b.stream_words * 4; }
static size_t tail_smem_bytes(const BatchDev& b)
{
    return size_t(b.lut_slots) * kLutEntries * 2 + size_t(kTailCta) * kMaxBlocksPerMcu * sizeof(TabSel) + size_t(kTailCta) * b.stream_words * 4;
}

int huff_configure(const BatchDev& b, int device, int* tail_grid)
{
    const size_t ms = main_smem_bytes(b), ts = tail_smem_bytes(b);
    if (cudaFuncSetAttribute(k_huff_sync, cudaFuncAttributeMaxDynamicSharedMemorySize, int(ms)) != cudaSuccess ||
        cudaFuncSetAttribute(k_huff_write, cudaFuncAttributeMaxDynamicSharedMemorySize, int(ms)) != cudaSuccess ||
        cudaFuncSetAttribute(k_huff_sync_tail, cudaFuncAttributeMaxDynamicSharedMemorySize, int(ts)) != cudaSuccess ||
        cudaFuncSetAttribute(k_huff_tail_round, cudaFuncAttributeMaxDynamicSharedMemorySize, int(ts)) != cudaSuccess)
        return -1;
    int per_sm = 0, sms = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_huff_sync_tail, kTailCta, ts) != cudaSuccess || per_sm < 1)
        return -1;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || sms < 1)
        return -1;
    *tail_grid = per_sm * sms; // co-resident CTAs of the tail kernel on an otherwise empty device
    return 0;
}

int launch_huff_sync(const BatchDev& b, const LaunchDims& d, cudaStream_t st)
{
    const size_t ms = main_smem_bytes(b);
    k_huff_sync<<<d.subseq_ctas, kSubseqCta, ms, st>>>(b, 0);
    k_huff_sync<<<d.subseq_ctas, kSubseqCta, ms, st>>>(b, 1);
    return 2;
}

int launch_huff_tail(const BatchDev& b, const LaunchDims& d, cudaStream_t st, int tail_grid, int plain_rounds)
{
    const int plain_grid = std::max(tail_grid, int((size_t(d.subseq_ctas) * kSubseqCta + kTailCta - 1) / kTailCta));
    for (int r = 0; r < plain_rounds; ++r)
        k_huff_tail_round<<<plain_grid, kTailCta, tail_smem_bytes(b), st>>>(b, r);
    BatchDev bb = b;
    int first_round = plain_rounds;
    void* args[] = {&bb, &first_round};
    if (cudaLaunchCooperativeKernel(reinterpret_cast<void*>(k_huff_sync_tail), dim3(tail_grid), dim3(kTailCta), args, tail_smem_bytes(b), st) !=
        cudaSuccess)
        return -1;
    return 1 + plain_rounds;
}

int fused_configure(const BatchDev& b)
{
    return cudaFuncSetAttribute(k_huff_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, int(fused_smem_bytes(b))) == cudaSuccess ? 0 : -1;
}

int launch_huff_fused(const BatchDev& b, const LaunchDims& d, cudaStream_t st)
{
    k_huff_fused<<<d.fused_ctas, kSubseqCta, fused_smem_bytes(b), st>>>(b);
    return 1;
}

int launch_huff_write(const BatchDev& b, const LaunchDims& d, cudaStream_t st)
{
    k_subseq_scan<<<d.scan_ctas, kScanThreads, 0, st>>>(b);
    k_huff_write<<<d.subseq_ctas, kSubseqCta, main_smem_bytes(b), st>>>(b);
    return 2;
}

} // namespace b200jpeg
