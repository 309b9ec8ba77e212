// capi.cpp -- extern "C" boundary of libb200jpeg.so (include/b200jpeg.h).
#include <cuda_runtime_api.h>
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <numeric>
#include <thread>
#include <cstring>
#include <memory>
#include <new>
#include <vector>
#include "../../include/b200jpeg.h"
#include "decoder.h"
#include "encoder.h"
#include "jpeg_parse.h"
#include "log.h"

using namespace b200jpeg;

// Two complete decoder contexts (streams, scratch, pinned staging) so that two batches can be in flight
// (b200jpeg_decode_submit / _wait); the second one is created when it is first needed.
struct b200jpegDecoder
{
    Decoder* slot[2] = {nullptr, nullptr};
    int device = 0;
    DecoderOptions opt;
    Allocators alloc;
    bool alloc_set = false;
    std::atomic<uint64_t> submitted{0}, completed{0}; // submit and wait may come from two threads (one each)
    int last = 0; // slot of the batch whose results get_stats / debug_fetch describe
    Decoder* impl() { return slot[last]; }
    ~b200jpegDecoder()
    {
        delete slot[0];
        delete slot[1];
    }
};

// One decoder and one persistent host thread per device; a call hands every thread its share of the batch and waits.
struct b200jpegMultiDecoder
{
    struct Worker
    {
        b200jpegDecoder* dec = nullptr;
        std::thread th;
        std::vector<const uint8_t*> data;
        std::vector<size_t> sizes;
        std::vector<b200jpegOutput_t> outs;
        std::vector<int32_t> statuses;
        std::vector<int> index; // position in the caller's batch
        int rc = 0;
    };
    std::vector<Worker> workers;
    std::mutex mu;
    std::condition_variable cv;
    uint64_t epoch = 0;
    int pending = 0;
    bool stop = false;
};

static int make_slot(b200jpegDecoder* d, int i)
{
    if (d->slot[i])
        return B200JPEG_STATUS_SUCCESS;
    Decoder* dec = new Decoder(d->device, d->opt);
    int rc = dec->init();
    if (rc == B200JPEG_STATUS_SUCCESS && d->alloc_set)
        rc = dec->set_allocators(d->alloc);
    if (rc != B200JPEG_STATUS_SUCCESS) {
        delete dec;
        return rc;
    }
    d->slot[i] = dec;
    return B200JPEG_STATUS_SUCCESS;
}

static std::vector<Decoder::OutputDesc> to_outputs(int n, const b200jpegOutput_t* outputs)
{
    std::vector<Decoder::OutputDesc> outs(n);
    for (int i = 0; i < n; ++i) {
        outs[i] = Decoder::OutputDesc{outputs[i].buffer, outputs[i].row_stride, outputs[i].format, outputs[i].on_host, outputs[i].roi_x,
            outputs[i].roi_y, outputs[i].roi_width, outputs[i].roi_height};
        outs[i].stream = static_cast<cudaStream_t>(outputs[i].cuda_stream);
        outs[i].has_stream = outputs[i].has_cuda_stream != 0;
        outs[i].orient = outputs[i].orientation & 15;
    }
    return outs;
}
struct b200jpegEncoder
{
    Encoder* impl;
};

extern "C" {

const char* b200jpeg_version(void) { return "b200jpeg 0.2.0 (sm_100a)"; }

void b200jpeg_set_log_callback(b200jpegLogFn_t fn, void* ctx) { set_log_callback(reinterpret_cast<LogFn>(fn), ctx); }

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
        const DecoderOptions opt = parse_decoder_options(options, "b200jpeg_decoder");
        if (opt.fast_idct)
            return B200JPEG_STATUS_UNSUPPORTED; // JDCT_IFAST pixels are not offered (see DecoderOptions::fast_idct)
        b200jpegDecoder* d = new b200jpegDecoder();
        d->device = device_id;
        d->opt = opt;
        const int rc = make_slot(d, 0);
        if (rc != B200JPEG_STATUS_SUCCESS) {
            delete d;
            return rc;
        }
        *decoder = d;
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
    if (!decoder || decoder->slot[1])
        return B200JPEG_STATUS_INVALID_PARAMETER;
    decoder->alloc = to_allocators(allocators);
    decoder->alloc_set = true;
    return decoder->slot[0]->set_allocators(decoder->alloc);
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
    delete decoder;
    return B200JPEG_STATUS_SUCCESS;
}

int b200jpeg_decode_prepare(b200jpegDecoder_t decoder, int n, const uint8_t* const* data, const size_t* sizes,
    const b200jpegOutput_t* outputs, int32_t* statuses)
{
    if (!decoder || n < 0 || decoder->submitted != decoder->completed)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    try {
        const std::vector<Decoder::OutputDesc> outs = to_outputs(n, outputs);
        decoder->last = 0;
        return decoder->slot[0]->prepare(n, data, sizes, outs.data(), statuses);
    } catch (const std::bad_alloc&) {
        return B200JPEG_STATUS_ALLOC_FAILED;
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
}

int b200jpeg_decode_run(b200jpegDecoder_t decoder, void* user_stream, int32_t* statuses)
{
    if (!decoder || !statuses || decoder->submitted != decoder->completed)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    try {
        decoder->last = 0;
        return decoder->slot[0]->run(static_cast<cudaStream_t>(user_stream), statuses);
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
}

int b200jpeg_decode_batch(b200jpegDecoder_t decoder, int n, const uint8_t* const* data, const size_t* sizes,
    const b200jpegOutput_t* outputs, void* user_stream, int32_t* statuses)
{
    if (!decoder || n < 0 || decoder->submitted != decoder->completed)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    try {
        const std::vector<Decoder::OutputDesc> outs = to_outputs(n, outputs);
        decoder->last = 0;
        return decoder->slot[0]->decode(n, data, sizes, outs.data(), static_cast<cudaStream_t>(user_stream), statuses);
    } catch (const std::bad_alloc&) {
        return B200JPEG_STATUS_ALLOC_FAILED;
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
}

int b200jpeg_decode_submit(b200jpegDecoder_t decoder, int n, const uint8_t* const* data, const size_t* sizes,
    const b200jpegOutput_t* outputs, void* user_stream, int32_t* statuses)
{
    if (!decoder || n < 0 || decoder->submitted - decoder->completed >= 2)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    try {
        const int s = int(decoder->submitted & 1);
        int rc = make_slot(decoder, s);
        if (rc != B200JPEG_STATUS_SUCCESS)
            return rc;
        const std::vector<Decoder::OutputDesc> outs = to_outputs(n, outputs);
        rc = decoder->slot[s]->submit(n, data, sizes, outs.data(), static_cast<cudaStream_t>(user_stream), statuses);
        if (rc != B200JPEG_STATUS_SUCCESS) {
            if (decoder->slot[s]->in_flight()) { // partly enqueued: drain it, the batch is not counted
                std::vector<int32_t> tmp(n > 0 ? n : 1);
                decoder->slot[s]->wait(tmp.data());
            }
            return rc;
        }
        ++decoder->submitted;
        return B200JPEG_STATUS_SUCCESS;
    } catch (const std::bad_alloc&) {
        return B200JPEG_STATUS_ALLOC_FAILED;
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
}

int b200jpeg_decode_wait(b200jpegDecoder_t decoder, int32_t* statuses)
{
    if (!decoder || !statuses || decoder->submitted == decoder->completed)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    try {
        const int s = int(decoder->completed & 1);
        ++decoder->completed;
        decoder->last = s;
        return decoder->slot[s]->wait(statuses);
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
}

// ---- multi-device ----
int b200jpeg_shard_batch(int n, const size_t* sizes, const int32_t* widths, const int32_t* heights, int num_devices, int32_t* device_of_image)
{
    if (n < 0 || num_devices < 1 || (n > 0 && (!sizes || !device_of_image)))
        return B200JPEG_STATUS_INVALID_PARAMETER;
    // Huffman work scales with the compressed bytes (about 8 symbols per byte-pixel), reconstruction with the pixels
    std::vector<uint64_t> cost(n);
    for (int i = 0; i < n; ++i)
        cost[i] = 8ull * sizes[i] + (widths && heights ? uint64_t(widths[i]) * uint64_t(heights[i]) : 0ull);
    std::vector<int> order(n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cost[a] > cost[b]; });
    std::vector<uint64_t> load(num_devices, 0);
    for (int i : order) {
        const int r = int(std::min_element(load.begin(), load.end()) - load.begin());
        device_of_image[i] = r;
        load[r] += cost[i];
    }
    return B200JPEG_STATUS_SUCCESS;
}

int b200jpeg_decoder_get_stats(b200jpegDecoder_t decoder, uint64_t* stats16)
{
    if (!decoder || !stats16)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    decoder->impl()->get_stats(stats16);
    return B200JPEG_STATUS_SUCCESS;
}

int b200jpeg_decoder_get_stage_times(b200jpegDecoder_t decoder, float* ms, int n)
{
    if (!decoder || !ms)
        return -1;
    return decoder->impl()->get_stage_times(ms, n);
}

int64_t b200jpeg_debug_fetch(b200jpegDecoder_t decoder, int what, int image, int component, void* host_dst, size_t capacity)
{
    if (!decoder)
        return -1;
    return decoder->impl()->debug_fetch(what, image, component, host_dst, capacity);
}

int b200jpeg_multi_decoder_create(b200jpegMultiDecoder_t* out, int num_devices, const int* device_ids, const char* options)
{
    if (!out || num_devices < 1 || !device_ids)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    *out = nullptr;
    try {
        std::unique_ptr<b200jpegMultiDecoder> md(new b200jpegMultiDecoder());
        md->workers.resize(num_devices);
        for (int w = 0; w < num_devices; ++w) {
            const int rc = b200jpeg_decoder_create(&md->workers[w].dec, device_ids[w], options);
            if (rc != B200JPEG_STATUS_SUCCESS) {
                for (auto& wk : md->workers)
                    if (wk.dec)
                        b200jpeg_decoder_destroy(wk.dec);
                return rc;
            }
        }
        b200jpegMultiDecoder* m = md.get();
        for (int w = 0; w < num_devices; ++w)
            m->workers[w].th = std::thread([m, w] {
                uint64_t seen = 0;
                b200jpegMultiDecoder::Worker& me = m->workers[w];
                for (;;) {
                    {
                        std::unique_lock<std::mutex> lk(m->mu);
                        m->cv.wait(lk, [&] { return m->stop || m->epoch != seen; });
                        if (m->stop)
                            return;
                        seen = m->epoch;
                    }
                    const int n = int(me.index.size());
                    me.rc = n ? b200jpeg_decode_batch(me.dec, n, me.data.data(), me.sizes.data(), me.outs.data(), nullptr, me.statuses.data())
                              : B200JPEG_STATUS_SUCCESS;
                    {
                        std::lock_guard<std::mutex> lk(m->mu);
                        --m->pending;
                    }
                    m->cv.notify_all();
                }
            });
        *out = md.release();
    } catch (const std::bad_alloc&) {
        return B200JPEG_STATUS_ALLOC_FAILED;
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
    return B200JPEG_STATUS_SUCCESS;
}

int b200jpeg_multi_decoder_destroy(b200jpegMultiDecoder_t md)
{
    if (!md)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    {
        std::lock_guard<std::mutex> lk(md->mu);
        md->stop = true;
    }
    md->cv.notify_all();
    for (auto& w : md->workers) {
        if (w.th.joinable())
            w.th.join();
        if (w.dec)
            b200jpeg_decoder_destroy(w.dec);
    }
    delete md;
    return B200JPEG_STATUS_SUCCESS;
}

int b200jpeg_multi_decode_batch(b200jpegMultiDecoder_t md, int n, const uint8_t* const* data, const size_t* sizes, const b200jpegOutput_t* outputs,
    const int32_t* device_of_image, int32_t* statuses)
{
    if (!md || n < 0 || (n > 0 && (!data || !sizes || !outputs || !device_of_image || !statuses)))
        return B200JPEG_STATUS_INVALID_PARAMETER;
    try {
        const int nd = int(md->workers.size());
        for (auto& w : md->workers) {
            w.data.clear();
            w.sizes.clear();
            w.outs.clear();
            w.index.clear();
        }
        for (int i = 0; i < n; ++i) {
            if (device_of_image[i] < 0 || device_of_image[i] >= nd)
                return B200JPEG_STATUS_INVALID_PARAMETER;
            auto& w = md->workers[device_of_image[i]];
            w.data.push_back(data[i]);
            w.sizes.push_back(sizes[i]);
            w.outs.push_back(outputs[i]);
            w.index.push_back(i);
        }
        for (auto& w : md->workers)
            w.statuses.assign(std::max<size_t>(1, w.index.size()), B200JPEG_STATUS_INTERNAL_ERROR);
        {
            std::lock_guard<std::mutex> lk(md->mu);
            md->pending = nd;
            ++md->epoch;
        }
        md->cv.notify_all();
        {
            std::unique_lock<std::mutex> lk(md->mu);
            md->cv.wait(lk, [&] { return md->pending == 0; });
        }
        int rc = B200JPEG_STATUS_SUCCESS;
        for (auto& w : md->workers) {
            if (w.rc != B200JPEG_STATUS_SUCCESS)
                rc = w.rc;
            for (size_t k = 0; k < w.index.size(); ++k)
                statuses[w.index[k]] = w.rc == B200JPEG_STATUS_SUCCESS ? w.statuses[k] : w.rc;
        }
        return rc;
    } catch (const std::bad_alloc&) {
        return B200JPEG_STATUS_ALLOC_FAILED;
    } catch (...) {
        return B200JPEG_STATUS_INTERNAL_ERROR;
    }
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
