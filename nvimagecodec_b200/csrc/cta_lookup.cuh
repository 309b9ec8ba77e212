// cta_lookup.cuh -- which image does a CTA (or any item of a per-image prefix array) belong to.
//
// Every descriptor-driven kernel starts with this lookup, and for the short-lived CTAs of the pixel kernels a binary search
// of dependent loads was a visible share of their life (ncu: 8-17 % of the stall samples of k_idct / k_color).  The whole
// warp searches instead: a binary search narrows the range to 64 entries (no step at all for the usual groups of a few
// dozen images), then every lane tests two entries with independent loads and two ballots count the entries <= v.
#pragma once
#include <cstdint>

namespace b200jpeg {

// map[0..n] ascending, map[0] <= v < map[n]: the i with map[i] <= v < map[i + 1].  Any thread, any argument.
__device__ __forceinline__ int find_image_scalar(const uint32_t* __restrict__ map, int n, uint32_t v)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(map + mid) <= v)
            lo = mid;
        else
            hi = mid - 1;
    }
    return lo;
}

// The same answer, found by the warp together.  Must be called by whole, converged
// warps (all lanes with the same arguments).
__device__ __forceinline__ int find_image_warp(const uint32_t* __restrict__ map, int n, uint32_t v)
{
    int lo = 0, hi = n - 1;
    while (hi - lo > 64) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(map + mid) <= v)
            lo = mid;
        else
            hi = mid - 1;
    }
    const int lane = int(threadIdx.x & 31u);
    const int i0 = lo + 1 + lane, i1 = i0 + 32;
    const bool a = i0 <= hi && __ldg(map + i0) <= v;
    const bool c = i1 <= hi && __ldg(map + i1) <= v;
    return lo + __popc(__ballot_sync(0xFFFFFFFFu, a)) + __popc(__ballot_sync(0xFFFFFFFFu, c));
}

} // namespace b200jpeg
