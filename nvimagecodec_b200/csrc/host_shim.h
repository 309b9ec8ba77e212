// host_shim.h -- vector types for compiling jpeg_math.cuh with a plain C++ compiler (CPU emulation in tests only).
#pragma once
#include <cstdint>
namespace b200jpeg {
struct alignas(8) uint2 { uint32_t x, y; };
struct alignas(16) uint4 { uint32_t x, y, z, w; };
inline uint2 make_uint2(uint32_t x, uint32_t y) { return uint2{x, y}; }
inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
} // namespace b200jpeg
