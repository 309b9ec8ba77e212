// encoder.cpp -- placeholder until the encode kernels land: every image reports B200JPEG_STATUS_UNSUPPORTED (there is no
// CPU fallback in this library).
#include "encoder.h"

namespace b200jpeg {

struct Encoder::Impl
{
    int device;
};

Encoder::Encoder(int device_id, const char*) : impl_(new Impl{device_id}) {}
Encoder::~Encoder() { delete impl_; }
int Encoder::init() { return B200JPEG_STATUS_SUCCESS; }
int Encoder::encode(int n, const b200jpegEncodeInput_t*, uint8_t* const*, const size_t*, size_t* out_sizes, cudaStream_t, int32_t* statuses)
{
    for (int i = 0; i < n; ++i) {
        statuses[i] = B200JPEG_STATUS_UNSUPPORTED;
        out_sizes[i] = 0;
    }
    return B200JPEG_STATUS_SUCCESS;
}
void Encoder::get_stats(uint64_t* s8) const
{
    for (int i = 0; i < 8; ++i)
        s8[i] = 0;
}

} // namespace b200jpeg
