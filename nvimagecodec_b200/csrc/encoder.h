// encoder.h -- batch baseline-JPEG encoder behind b200jpeg_encode_batch.
#pragma once
#include <cuda_runtime_api.h>
#include <cstddef>
#include <cstdint>
#include "../../include/b200jpeg.h"
#include "alloc.h"

namespace b200jpeg {

class Encoder
{
  public:
    Encoder(int device_id, const char* options);
    ~Encoder();
    int init();
    int encode(int n, const b200jpegEncodeInput_t* inputs, uint8_t* const* out, const size_t* out_capacity, size_t* out_sizes,
        cudaStream_t user_stream, int32_t* statuses);
    void get_stats(uint64_t* s8) const;
    int set_allocators(const Allocators& a); // before the first encode

  private:
    struct Impl;
    Impl* impl_;
};

} // namespace b200jpeg
