// capi.cpp -- extern "C" boundary of libb200jpeg.so (include/b200jpeg.h).
#include <cuda_runtime_api.h>
#include <cstring>
#include <new>
#include <vector>
#include "../../include/b200jpeg.h"
#include "decoder.h"
#include "encoder.h"
#include "jpeg_parse.h"

using namespace b200jpeg;

struct b200jpegDecoder
{
    Decoder* impl;
};
struct b200jpegEncoder
{
    Encoder* impl;
};

extern "C" {

const char* b200jpeg_version(void) { return "b200jpeg 0.1.0 (sm_100a)"; }

const char* b200jpeg_status_string(int s)
{
    switch (s) {
    case B200JPEG_STATUS_SUCCESS: return "success";
    case B200JPEG_STATUS_INVALID_PARAMETER: return "invalid parameter";
    case B200JPEG_STATUS_BAD_JPEG: return "bad jpeg";
    case B200JPEG_STATUS_UNSUPPORTED: return "unsupported jpeg";
    case B200JPEG_STATUS_CORRUPT_STREAM: return "corrupt entropy-coded data";
    case B200JPEG_STATUS_CUDA_ERROR: return "cuda error";
    case B200JPEG_STATUS_ARCH_MISMATCH: return "device is not sm_100";
    case B200JPEG_STATUS_ALLOC_FAILED: return "allocation failed";
    case B200JPEG_STATUS_BUFFER_TOO_SMALL: return "output buffer too small";
    default: return "internal error";
    }
}

int b200jpeg_parse_header(const uint8_t* data, size_t size, b200jpegStreamInfo_t* info)
{
    if (!data || !info)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    std::memset(info, 0, sizeof(*info));
    ParsedJpeg p;
    const ParseResult r = parse_jpeg(data, size, p);
    if (r != PARSE_OK)
        return B200JPEG_STATUS_BAD_JPEG;
    info->width = p.width;
    info->height = p.height;
    info->num_components = p.ncomp;
    for (int c = 0; c < 3; ++c) {
        info->h_samp[c] = p.hs[c];
        info->v_samp[c] = p.vs[c];
    }
    info->subsampling = subsampling_of(p);
    info->sof_marker = p.sof;
    info->restart_interval = p.restart_interval;
    info->precision = p.precision;
    info->supported = is_supported(p) ? 1 : 0;
    info->scan_offset = p.scan_offset;
    info->scan_bytes = p.scan_bytes;
    return B200JPEG_STATUS_SUCCESS;
}

int b200jpeg_decoder_create(b200jpegDecoder_t* decoder, int device_id, const char* options)
{
    if (!decoder)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    *decoder = nullptr;
    try {
        Decoder* d = new Decoder(device_id, parse_decoder_options(options, "b200jpeg_decoder"));
        const int rc = d->init();
        if (rc != B200JPEG_STATUS_SUCCESS) {
            delete d;
            return rc;
        }
        *decoder = new b200jpegDecoder{d};
    } catch (const std::bad_alloc&) {
        return B200JPEG_STATUS_ALLOC_FAILED;
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
    return B200JPEG_STATUS_SUCCESS;
}

static Allocators to_allocators(const b200jpegAllocators_t* a)
{
    Allocators r;
    if (a) {
        r.device_malloc = reinterpret_cast<AllocFn>(a->device_malloc);
        r.device_free = reinterpret_cast<FreeFn>(a->device_free);
        r.device_ctx = a->device_ctx;
        r.device_padding = a->device_padding;
        r.pinned_malloc = reinterpret_cast<AllocFn>(a->pinned_malloc);
        r.pinned_free = reinterpret_cast<FreeFn>(a->pinned_free);
        r.pinned_ctx = a->pinned_ctx;
        r.pinned_padding = a->pinned_padding;
    }
    return r;
}

int b200jpeg_decoder_set_allocators(b200jpegDecoder_t decoder, const b200jpegAllocators_t* allocators)
{
    if (!decoder)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    return decoder->impl->set_allocators(to_allocators(allocators));
}

int b200jpeg_encoder_set_allocators(b200jpegEncoder_t encoder, const b200jpegAllocators_t* allocators)
{
    if (!encoder)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    return encoder->impl->set_allocators(to_allocators(allocators));
}

int b200jpeg_decoder_destroy(b200jpegDecoder_t decoder)
{
    if (!decoder)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    delete decoder->impl;
    delete decoder;
    return B200JPEG_STATUS_SUCCESS;
}

int b200jpeg_decode_prepare(b200jpegDecoder_t decoder, int n, const uint8_t* const* data, const size_t* sizes,
    const b200jpegOutput_t* outputs, int32_t* statuses)
{
    if (!decoder || n < 0)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    try {
        std::vector<Decoder::OutputDesc> outs(n);
        for (int i = 0; i < n; ++i)
            outs[i] = Decoder::OutputDesc{outputs[i].buffer, outputs[i].row_stride, outputs[i].format, outputs[i].on_host, outputs[i].roi_x,
                outputs[i].roi_y, outputs[i].roi_width, outputs[i].roi_height};
        return decoder->impl->prepare(n, data, sizes, outs.data(), statuses);
    } catch (const std::bad_alloc&) {
        return B200JPEG_STATUS_ALLOC_FAILED;
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
}

int b200jpeg_decode_run(b200jpegDecoder_t decoder, void* user_stream, int32_t* statuses)
{
    if (!decoder || !statuses)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    try {
        return decoder->impl->run(static_cast<cudaStream_t>(user_stream), statuses);
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
}

int b200jpeg_decode_batch(b200jpegDecoder_t decoder, int n, const uint8_t* const* data, const size_t* sizes,
    const b200jpegOutput_t* outputs, void* user_stream, int32_t* statuses)
{
    if (!decoder || n < 0)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    try {
        std::vector<Decoder::OutputDesc> outs(n);
        for (int i = 0; i < n; ++i)
            outs[i] = Decoder::OutputDesc{outputs[i].buffer, outputs[i].row_stride, outputs[i].format, outputs[i].on_host, outputs[i].roi_x,
                outputs[i].roi_y, outputs[i].roi_width, outputs[i].roi_height};
        return decoder->impl->decode(n, data, sizes, outs.data(), static_cast<cudaStream_t>(user_stream), statuses);
    } catch (const std::bad_alloc&) {
        return B200JPEG_STATUS_ALLOC_FAILED;
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
}

int b200jpeg_decoder_get_stats(b200jpegDecoder_t decoder, uint64_t* stats16)
{
    if (!decoder || !stats16)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    decoder->impl->get_stats(stats16);
    return B200JPEG_STATUS_SUCCESS;
}

int b200jpeg_decoder_get_stage_times(b200jpegDecoder_t decoder, float* ms, int n)
{
    if (!decoder || !ms)
        return -1;
    return decoder->impl->get_stage_times(ms, n);
}

int64_t b200jpeg_debug_fetch(b200jpegDecoder_t decoder, int what, int image, int component, void* host_dst, size_t capacity)
{
    if (!decoder)
        return -1;
    return decoder->impl->debug_fetch(what, image, component, host_dst, capacity);
}

int b200jpeg_encoder_create(b200jpegEncoder_t* encoder, int device_id, const char* options)
{
    if (!encoder)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    *encoder = nullptr;
    try {
        Encoder* e = new Encoder(device_id, options);
        const int rc = e->init();
        if (rc != B200JPEG_STATUS_SUCCESS) {
            delete e;
            return rc;
        }
        *encoder = new b200jpegEncoder{e};
    } catch (const std::bad_alloc&) {
        return B200JPEG_STATUS_ALLOC_FAILED;
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
    return B200JPEG_STATUS_SUCCESS;
}

int b200jpeg_encoder_destroy(b200jpegEncoder_t encoder)
{
    if (!encoder)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    delete encoder->impl;
    delete encoder;
    return B200JPEG_STATUS_SUCCESS;
}

int b200jpeg_encode_batch(b200jpegEncoder_t encoder, int n, const b200jpegEncodeInput_t* inputs, uint8_t* const* out,
    const size_t* out_capacity, size_t* out_sizes, void* user_stream, int32_t* statuses)
{
    if (!encoder || n < 0 || (n > 0 && (!inputs || !out || !out_capacity || !out_sizes || !statuses)))
        return B200JPEG_STATUS_INVALID_PARAMETER;
    try {
        return encoder->impl->encode(n, inputs, out, out_capacity, out_sizes, static_cast<cudaStream_t>(user_stream), statuses);
    } catch (const std::bad_alloc&) {
        return B200JPEG_STATUS_ALLOC_FAILED;
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
}

int b200jpeg_encoder_get_stats(b200jpegEncoder_t encoder, uint64_t* stats8)
{
    if (!encoder || !stats8)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    encoder->impl->get_stats(stats8);
    return B200JPEG_STATUS_SUCCESS;
}

} // extern "C"
