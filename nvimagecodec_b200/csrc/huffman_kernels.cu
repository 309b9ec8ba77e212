// huffman_kernels.cu -- sm_100a kernels of the entropy-decode stage of b200jpeg.
//
//   H1 k_huff_sync x 2   self-synchronising subsequence decode (no output): round 0 from a guessed state, round 1 from the
//                        predecessor's exit state; subsequences whose exit state changed feed a work list
//   H2 k_huff_sync_tail  cooperative kernel: Jacobi rounds over the (compacted) work list until it is empty
//   H3 k_subseq_scan     per image: segmented scan of coefficient-slot counts -> absolute block position per subsequence
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

constexpr int kScanThreads = 256;
__global__ void __launch_bounds__(kScanThreads) k_subseq_scan(BatchDev b)
{
    const int img = blockIdx.x;
    const ImageDesc& im = b.images[img];
    const uint32_t count = im.subseq_cta_count * kSubseqCta;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ uint32_t s_w[kScanThreads / 32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0)
        s_carry = 0x80000000u;
    __syncthreads();
    bool bad = false;
    for (uint32_t c0 = 0; c0 < count; c0 += kScanThreads) {
        const uint32_t c = c0 + threadIdx.x;
        uint32_t v = c < count ? b.sub_rec[im.subseq_first + c] : 0u;
        if (c > 0 && c < count && b.sub_entry[im.subseq_first + c] != b.sub_exit[im.subseq_first + c - 1] &&
            uint64_t(c) * uint32_t(b.subseq_bits) < uint64_t(b.states[img].clean_len) * 8u)
            bad = true; // the chain of entry/exit states is not consistent: synchronisation did not finish
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t tprev = __shfl_up_sync(0xFFFFFFFFu, v, o);
            if (lane >= uint32_t(o))
                v = seg_combine(tprev, v);
        }
        if (lane == 31)
            s_w[warp] = v;
        __syncthreads();
        uint32_t pre = s_carry;
        for (int i = 0; i < int(warp); ++i)
            pre = seg_combine(pre, s_w[i]);
        v = seg_combine(pre, v);
        if (c < count)
            b.sub_abs[im.subseq_first + c] = v;
        __syncthreads();
        if (threadIdx.x == kScanThreads - 1)
            s_carry = v;
        __syncthreads();
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

int launch_huff_write(const BatchDev& b, const LaunchDims& d, cudaStream_t st)
{
    k_subseq_scan<<<d.num_images, kScanThreads, 0, st>>>(b);
    k_huff_write<<<d.subseq_ctas, kSubseqCta, main_smem_bytes(b), st>>>(b);
    return 2;
}

} // namespace b200jpeg
