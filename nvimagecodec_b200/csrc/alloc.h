// alloc.h -- where the scratch memory of the decoder / encoder comes from: cudaMalloc / cudaHostAlloc, or the caller's
// stream-ordered allocator callbacks (the nvimgcodecDeviceAllocator_t / nvimgcodecPinnedAllocator_t of
// nvimgcodecExecutionParams_t, reference include/nvimgcodec.h:232-302: (ctx, &ptr, size, stream) / (ctx, ptr, size, stream),
// 0 = success, allocations padded to a multiple of *_mem_padding).
#pragma once
#include <cuda_runtime_api.h>
#include <cstddef>

namespace b200jpeg {

typedef int (*AllocFn)(void* ctx, void** ptr, size_t size, cudaStream_t stream);
typedef int (*FreeFn)(void* ctx, void* ptr, size_t size, cudaStream_t stream);

struct Allocators
{
    AllocFn device_malloc = nullptr;
    FreeFn device_free = nullptr;
    void* device_ctx = nullptr;
    size_t device_padding = 0;
    AllocFn pinned_malloc = nullptr;
    FreeFn pinned_free = nullptr;
    void* pinned_ctx = nullptr;
    size_t pinned_padding = 0;

    static size_t pad(size_t bytes, size_t to) { return to > 1 ? (bytes + to - 1) / to * to : bytes; }

    // *size is rounded up to what was really requested (the free call must pass the same number)
    bool alloc(bool pinned, void** p, size_t* size, cudaStream_t st) const
    {
        *p = nullptr;
        if (pinned && pinned_malloc && pinned_free) {
            *size = pad(*size, pinned_padding);
            return pinned_malloc(pinned_ctx, p, *size, st) == 0 && *p;
        }
        if (!pinned && device_malloc && device_free) {
            *size = pad(*size, device_padding);
            return device_malloc(device_ctx, p, *size, st) == 0 && *p;
        }
        const cudaError_t e = pinned ? cudaHostAlloc(p, *size, cudaHostAllocDefault) : cudaMalloc(p, *size);
        if (e != cudaSuccess) {
            cudaGetLastError();
            *p = nullptr;
            return false;
        }
        return true;
    }
    void release(bool pinned, void* p, size_t size, cudaStream_t st) const
    {
        if (!p)
            return;
        if (pinned && pinned_malloc && pinned_free)
            pinned_free(pinned_ctx, p, size, st);
        else if (!pinned && device_malloc && device_free)
            device_free(device_ctx, p, size, st);
        else if (pinned)
            cudaFreeHost(p);
        else
            cudaFree(p);
    }
};

} // namespace b200jpeg
