// noalloc_unstuff.cu -- b200jpeg_unstuff_noalloc: the unstuffing stage (U1-U3 of decode_kernels.cu) for ONE entropy-coded
// segment, as an entry point that allocates nothing.  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops).
//
// The decoder's unstuff kernels keep their per-chunk counts and the per-image state in scratch that Decoder::ensure
// allocates.  Here that scratch is a file-scope __device__ array of fixed size: every GPU that loads the library holds
// kScratchBytes of it until exit.  The array is shared by all calls and keeps the last call's contents: the header words
// are cleared by the first kernel of every call, and calls that overlap on the GPU (different streams, different host
// threads) race on it -- callers serialise.
//
// Host code sees `g_scratch` only as a stand-in symbol; the device address comes from cudaGetSymbolAddress, once per
// device, with that device current.
#include <cuda_runtime.h>
#include <mutex>
#include "../../include/b200jpeg.h"
#include "log.h"

namespace b200jpeg {
namespace {

constexpr uint32_t kNaThreads = 256;
constexpr uint32_t kNaBytesPerThread = 16;
constexpr uint32_t kNaChunk = kNaThreads * kNaBytesPerThread; // 4 KB of raw data per CTA, as the decoder's kernels
constexpr uint32_t kNaMaxRaw = 64u << 20;                     // B200JPEG_NOALLOC_MAX_BYTES
constexpr uint32_t kNaMaxChunks = kNaMaxRaw / kNaChunk;       // 16384
constexpr uint32_t kNaHeaderWords = 4;                        // uint64 words in front of the chunk counts
constexpr size_t kScratchWords = kNaHeaderWords + kNaMaxChunks;
constexpr size_t kScratchBytes = kScratchWords * sizeof(unsigned long long);
static_assert(kScratchBytes == 131104, "the limit documented in include/b200jpeg.h");

// [0] raw_len - position of the first marker that is not RSTn (0 = none), [1] effective raw length, [2..3] spare,
// [4 + c] chunk c: kept bytes | RSTn markers << 32, after the scan the exclusive prefix of the same
static __device__ __align__(256) unsigned long long g_scratch[kScratchWords];

struct Class16
{
    uint32_t keep, marker, bad; // bit i: byte i is data / the FF of an RSTn marker / the FF of any other marker
    uint8_t byte[16];
};

// Classification of the 16 bytes at [base, base + 16) of a segment of `len` bytes: removed are 00 after FF (stuffing), FF
// before FF (fill), FF Dn and the Dn behind it (restart marker); the same rules as classify16 of decode_kernels.cu, one
// byte at a time (this entry point is about where its scratch lives, not about instruction counts).
__device__ __forceinline__ Class16 classify(const uint8_t* __restrict__ raw, uint32_t len, uint32_t base)
{
    Class16 c;
    c.keep = c.marker = c.bad = 0;
    uint32_t prev = (base > 0 && base <= len) ? raw[base - 1] : 0;
    uint32_t cur = base < len ? raw[base] : 0;
#pragma unroll
    for (uint32_t i = 0; i < 16; ++i) {
        const uint32_t pos = base + i;
        const uint32_t next = pos + 1 < len ? raw[pos + 1] : 0;
        c.byte[i] = uint8_t(cur);
        if (pos < len) {
            const bool next_rst = (next & 0xF8u) == 0xD0u;
            const bool removed = (cur == 0 && prev == 0xFF) || (cur == 0xFF && (next == 0xFF || next_rst)) || ((cur & 0xF8u) == 0xD0u && prev == 0xFF);
            if (!removed)
                c.keep |= 1u << i;
            if (cur == 0xFF && next_rst)
                c.marker |= 1u << i;
            if (pos + 1 < len && cur == 0xFF && !(next == 0 || next == 0xFF || next_rst))
                c.bad |= 1u << i;
        }
        prev = cur;
        cur = next;
    }
    return c;
}

__global__ void k_na_clear(unsigned long long* scratch)
{
    if (threadIdx.x < kNaHeaderWords)
        scratch[threadIdx.x] = 0;
}

__global__ void __launch_bounds__(kNaThreads) k_na_count(const uint8_t* __restrict__ raw, uint32_t len, unsigned long long* scratch)
{
    const uint32_t base = blockIdx.x * kNaChunk + threadIdx.x * kNaBytesPerThread;
    uint32_t kept = 0, markers = 0;
    if (base < len) {
        const Class16 c = classify(raw, len, base);
        kept = __popc(c.keep);
        markers = __popc(c.marker);
        if (c.bad) // the data ends at the earliest marker that is not RSTn
            atomicMax(&scratch[0], (unsigned long long)(len - (base + uint32_t(__ffs(int(c.bad))) - 1u)));
    }
    __shared__ uint32_t s_k, s_m;
    if (threadIdx.x == 0)
        s_k = s_m = 0;
    __syncthreads();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        kept += __shfl_down_sync(0xFFFFFFFFu, kept, o);
        markers += __shfl_down_sync(0xFFFFFFFFu, markers, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s_k, kept);
        atomicAdd(&s_m, markers);
    }
    __syncthreads();
    if (threadIdx.x == 0)
        scratch[kNaHeaderWords + blockIdx.x] = (unsigned long long)s_k | ((unsigned long long)s_m << 32);
}

// one CTA: exclusive scan of the chunk counts; chunks behind the cut count nothing, the chunk of the cut is recounted
__global__ void __launch_bounds__(1024) k_na_scan(const uint8_t* __restrict__ raw, uint32_t len, uint32_t chunks, unsigned long long* scratch,
    uint32_t* __restrict__ result)
{
    const uint32_t trunc = uint32_t(scratch[0]);
    const uint32_t eff = len - trunc;
    const uint32_t cut = trunc ? eff / kNaChunk : 0xFFFFFFFFu;
    __shared__ unsigned long long s_cut, s_warp[32], s_carry;
    if (threadIdx.x == 0) {
        s_cut = 0;
        s_carry = 0;
    }
    __syncthreads();
    if (trunc && threadIdx.x < kNaThreads) {
        const uint32_t base = cut * kNaChunk + threadIdx.x * kNaBytesPerThread;
        if (base < eff) {
            const Class16 c = classify(raw, eff, base);
            atomicAdd(&s_cut, (unsigned long long)__popc(c.keep) | ((unsigned long long)__popc(c.marker) << 32));
        }
    }
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long* counts = scratch + kNaHeaderWords;
    for (uint32_t c0 = 0; c0 < chunks; c0 += blockDim.x) {
        const uint32_t c = c0 + threadIdx.x;
        unsigned long long v = c < chunks ? counts[c] : 0;
        if (c >= cut)
            v = c == cut ? s_cut : 0;
        unsigned long long inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= uint32_t(o))
                inc += t;
        }
        if (lane == 31)
            s_warp[warp] = inc;
        __syncthreads();
        unsigned long long pre = s_carry;
        for (uint32_t w = 0; w < warp; ++w)
            pre += s_warp[w];
        if (c < chunks)
            counts[c] = pre + inc - v;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1)
            s_carry = pre + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        scratch[1] = eff;
        result[0] = uint32_t(s_carry);       // unstuffed bytes
        result[1] = uint32_t(s_carry >> 32); // RSTn markers removed
        result[2] = eff;                     // raw bytes consumed
        result[3] = 0;                       // set to 1 by k_na_write if the output did not fit
    }
}

__global__ void __launch_bounds__(kNaThreads) k_na_write(const uint8_t* __restrict__ raw, const unsigned long long* __restrict__ scratch,
    uint8_t* __restrict__ clean, uint32_t capacity, uint32_t* __restrict__ result)
{
    const uint32_t eff = uint32_t(scratch[1]);
    const uint32_t base = blockIdx.x * kNaChunk + threadIdx.x * kNaBytesPerThread;
    Class16 c;
    c.keep = 0;
    if (base < eff)
        c = classify(raw, eff, base);
    const uint32_t mine = __popc(c.keep);
    uint32_t inc = mine;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= uint32_t(o))
            inc += t;
    }
    __shared__ uint32_t s_w[kNaThreads / 32];
    __shared__ uint8_t s_out[kNaChunk];
    if (lane == 31)
        s_w[warp] = inc;
    __syncthreads();
    uint32_t pos = inc - mine, total = 0;
#pragma unroll
    for (uint32_t i = 0; i < kNaThreads / 32; ++i) {
        if (i < warp)
            pos += s_w[i];
        total += s_w[i];
    }
#pragma unroll
    for (uint32_t i = 0; i < 16; ++i)
        if (c.keep & (1u << i))
            s_out[pos++] = c.byte[i];
    __syncthreads();
    const uint32_t gbase = uint32_t(scratch[kNaHeaderWords + blockIdx.x]);
    if (gbase + total > capacity) {
        if (threadIdx.x == 0)
            result[3] = 1;
        return;
    }
    for (uint32_t i = threadIdx.x; i < total; i += kNaThreads)
        clean[gbase + i] = s_out[i];
}

// device address of g_scratch per device ordinal, looked up once with that device current
constexpr int kMaxDevices = 64;
std::mutex g_addr_mutex;
unsigned long long* g_scratch_addr[kMaxDevices] = {};

int scratch_address(int device_id, unsigned long long** out)
{
    std::lock_guard<std::mutex> lock(g_addr_mutex);
    if (!g_scratch_addr[device_id]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
        if (e != cudaSuccess || !p) {
            log_message(2, "b200jpeg_unstuff_noalloc: cudaGetSymbolAddress(g_scratch) on device %d: %s", device_id, cudaGetErrorString(e));
            return B200JPEG_STATUS_CUDA_ERROR;
        }
        g_scratch_addr[device_id] = static_cast<unsigned long long*>(p);
    }
    *out = g_scratch_addr[device_id];
    return B200JPEG_STATUS_SUCCESS;
}

} // namespace
} // namespace b200jpeg

extern "C" B200JPEG_API size_t b200jpeg_unstuff_noalloc_max_bytes(void) { return b200jpeg::kNaMaxRaw; }

extern "C" B200JPEG_API int b200jpeg_unstuff_noalloc(int device_id, const uint8_t* d_raw, size_t raw_bytes, uint8_t* d_clean, size_t clean_capacity,
    uint32_t* d_result, void* cuda_stream)
{
    using namespace b200jpeg;
    if (device_id < 0 || device_id >= kMaxDevices || !d_raw || !d_clean || !d_result || raw_bytes == 0)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    if (raw_bytes > kNaMaxRaw) {
        log_message(2, "b200jpeg_unstuff_noalloc: %zu bytes of entropy-coded data exceed the fixed scratch (limit %u bytes)", raw_bytes, kNaMaxRaw);
        return B200JPEG_STATUS_INVALID_PARAMETER;
    }
    if (cudaSetDevice(device_id) != cudaSuccess)
        return B200JPEG_STATUS_CUDA_ERROR;
    unsigned long long* scratch = nullptr;
    const int rc = scratch_address(device_id, &scratch);
    if (rc != B200JPEG_STATUS_SUCCESS)
        return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const uint32_t len = uint32_t(raw_bytes);
    const uint32_t chunks = (len + kNaChunk - 1) / kNaChunk;
    const uint32_t cap = clean_capacity > 0xFFFFFFFFull ? 0xFFFFFFFFu : uint32_t(clean_capacity);
    k_na_clear<<<1, 32, 0, st>>>(scratch);
    k_na_count<<<chunks, kNaThreads, 0, st>>>(d_raw, len, scratch);
    k_na_scan<<<1, 1024, 0, st>>>(d_raw, len, chunks, scratch, d_result);
    k_na_write<<<chunks, kNaThreads, 0, st>>>(d_raw, scratch, d_clean, cap, d_result);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        log_message(2, "b200jpeg_unstuff_noalloc: launch failed: %s", cudaGetErrorString(e));
        return B200JPEG_STATUS_CUDA_ERROR;
    }
    return B200JPEG_STATUS_SUCCESS;
}
