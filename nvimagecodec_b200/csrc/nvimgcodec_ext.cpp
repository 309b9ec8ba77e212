// nvimgcodec_ext.cpp -- nvImageCodec extension module "b200jpeg_extension".
//
// Exports nvimgcodecExtensionModuleEntry (reference include/nvimgcodec.h:1356-1364) and registers
//   * decoder  id "b200jpeg_decoder", codec "jpeg", backend NVIMGCODEC_BACKEND_KIND_GPU_ONLY, priority 50
//   * encoder  id "b200jpeg_encoder", codec "jpeg", backend NVIMGCODEC_BACKEND_KIND_GPU_ONLY, priority 50
// through the unchanged plugin ABI, so nvimgcodecDecoderDecode / nvimgcodecEncoderEncode, the registry's priority and
// fallback logic (src/codec.cpp:119-124, src/decoder_worker.cpp:63-93,252-307) and the Python bindings pick them up as a
// drop-in for the JPEG path.  Structure follows the reference's own plugins (extensions/libjpeg_turbo/libjpeg_turbo_ext.cpp:
// 28-100 for registration, libjpeg_turbo_decoder.cpp:118-173 for the canDecode matrix, nvjpeg/hw_decoder.cpp:492-680 for
// batched GPU plumbing, nvjpeg/cuda_encoder.cpp:284-396 for encode) -- the arithmetic is ours (libb200jpeg, CUDA sm_100a).
// Anything outside the codec's scope is declined in canDecode/canEncode so the framework falls back to the next decoder
// (libjpeg_turbo_ext, priority 300); there is no CPU fallback in here.
#include <nvimgcodec.h>
#include <algorithm>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "../../include/b200jpeg.h"
#include <nvtx3/nvToolsExt.h>

#ifndef NVIMGCODEC_VER
#include "nvimgcodec_version.h"
#endif

namespace {

constexpr float kPriority = 50.0f; // ahead of nvjpeg HW (100) / nvjpeg CUDA (200) / libjpeg_turbo (300) / opencv (400)
const char* kDecoderId = "b200jpeg_decoder";
const char* kEncoderId = "b200jpeg_encoder";

void log_msg(const nvimgcodecFrameworkDesc_t* fw, nvimgcodecDebugMessageSeverity_t sev, const char* id, const std::string& msg)
{
    if (!fw || !fw->log)
        return;
    nvimgcodecDebugMessageData_t data{NVIMGCODEC_STRUCTURE_TYPE_DEBUG_MESSAGE_DATA, sizeof(nvimgcodecDebugMessageData_t), nullptr,
        msg.c_str(), 0, nullptr, id, NVIMGCODEC_VER};
    fw->log(fw->instance, sev, NVIMGCODEC_DEBUG_MESSAGE_CATEGORY_GENERAL, &data);
}

// Bytes of a code stream: zero-copy map when the io stream supports it, else read into `storage`
// (idiom of extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:373-400).
const uint8_t* stream_bytes(nvimgcodecIoStreamDesc_t* io, std::vector<uint8_t>& storage, size_t& size)
{
    size = 0;
    if (!io || io->size(io->instance, &size) != NVIMGCODEC_STATUS_SUCCESS || size == 0)
        return nullptr;
    void* ptr = nullptr;
    if (io->map && io->map(io->instance, &ptr, 0, size) == NVIMGCODEC_STATUS_SUCCESS && ptr)
        return static_cast<const uint8_t*>(ptr);
    storage.resize(size);
    size_t got = 0;
    if (io->seek(io->instance, 0, SEEK_SET) != NVIMGCODEC_STATUS_SUCCESS)
        return nullptr;
    if (io->read(io->instance, &got, storage.data(), size) != NVIMGCODEC_STATUS_SUCCESS || got != size)
        return nullptr;
    return storage.data();
}

// ---------------------------------------------------------------------------------------------------------------
// Decoder
// ---------------------------------------------------------------------------------------------------------------
nvimgcodecProcessingStatus_t to_processing_status(int rc, int32_t res)
{
    if (rc != B200JPEG_STATUS_SUCCESS)
        return NVIMGCODEC_PROCESSING_STATUS_FAIL;
    switch (res) {
    case B200JPEG_STATUS_SUCCESS: return NVIMGCODEC_PROCESSING_STATUS_SUCCESS;
    case B200JPEG_STATUS_UNSUPPORTED: return NVIMGCODEC_PROCESSING_STATUS_CODESTREAM_UNSUPPORTED;
    case B200JPEG_STATUS_BAD_JPEG:
    case B200JPEG_STATUS_CORRUPT_STREAM: return NVIMGCODEC_PROCESSING_STATUS_IMAGE_CORRUPTED;
    default: return NVIMGCODEC_PROCESSING_STATUS_FAIL;
    }
}

// decode() hands a batch to the library (b200jpeg_decode_submit: host work done, device work enqueued and ordered onto
// the images' streams) and returns; the statuses -- corrupt entropy-coded data is found on the device -- are collected by
// this thread, which then reports imageReady for every sample of the batch.  Reporting from another thread after decode()
// has returned is what the reference's own plugins do (their executor tasks call imageReady,
// extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:444-490).  Two batches can be in flight, so the host work of batch
// k + 1 (canDecode, header parse, staging) overlaps the kernels of batch k.
struct DecoderState
{
    const nvimgcodecFrameworkDesc_t* fw = nullptr;
    const nvimgcodecExecutionParams_t* exec = nullptr;
    b200jpegDecoder_t dec = nullptr;
    bool async = true;

    struct Pending
    {
        std::vector<nvimgcodecImageDesc_t*> imgs;
    };
    std::mutex mu;
    std::condition_variable cv;
    std::deque<Pending> queue; // batches submitted and not yet reported (at most two)
    bool stop = false;
    std::thread completer;

    void complete_one(Pending& p)
    {
        const int n = int(p.imgs.size());
        std::vector<int32_t> res(n > 0 ? n : 1, B200JPEG_STATUS_INTERNAL_ERROR);
        const int rc = b200jpeg_decode_wait(dec, res.data());
        if (rc != B200JPEG_STATUS_SUCCESS)
            log_msg(fw, NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_ERROR, kDecoderId, std::string("decode batch failed: ") + b200jpeg_status_string(rc));
        for (int i = 0; i < n; ++i)
            p.imgs[i]->imageReady(p.imgs[i]->instance, to_processing_status(rc, res[i]));
    }
    void run()
    {
        if (exec)
            cudaSetDevice(exec->device_id);
        for (;;) {
            Pending* p = nullptr;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return stop || !queue.empty(); });
                if (queue.empty())
                    return; // stop requested and nothing left
                p = &queue.front(); // stays in the queue until it is reported: the queue's length is what is in flight
            }
            complete_one(*p);
            {
                std::lock_guard<std::mutex> lk(mu);
                queue.pop_front();
            }
            cv.notify_all();
        }
    }
    void shutdown()
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            stop = true;
        }
        cv.notify_all();
        try {
            if (completer.joinable() && completer.get_id() != std::this_thread::get_id())
                completer.join();
        } catch (...) { // a process that is being torn down may have taken the thread away already
        }
    }
};

nvimgcodecProcessingStatus_t check_target(const nvimgcodecImageInfo_t& info, const b200jpegStreamInfo_t& si,
    const nvimgcodecDecodeParams_t* params)
{
    nvimgcodecProcessingStatus_t st = NVIMGCODEC_PROCESSING_STATUS_SUCCESS;
    const uint32_t planes = info.num_planes;
    const uint32_t ch0 = info.plane_info[0].num_channels;
    switch (info.sample_format) {
    case NVIMGCODEC_SAMPLEFORMAT_I_RGB:
    case NVIMGCODEC_SAMPLEFORMAT_I_BGR:
    case NVIMGCODEC_SAMPLEFORMAT_I_UNCHANGED:
        if (planes != 1)
            st |= NVIMGCODEC_PROCESSING_STATUS_NUM_PLANES_UNSUPPORTED;
        if (ch0 != 3)
            st |= NVIMGCODEC_PROCESSING_STATUS_NUM_CHANNELS_UNSUPPORTED;
        break;
    case NVIMGCODEC_SAMPLEFORMAT_P_RGB:
    case NVIMGCODEC_SAMPLEFORMAT_P_BGR:
    case NVIMGCODEC_SAMPLEFORMAT_P_UNCHANGED:
        if (planes != 3)
            st |= NVIMGCODEC_PROCESSING_STATUS_NUM_PLANES_UNSUPPORTED;
        if (ch0 != 1)
            st |= NVIMGCODEC_PROCESSING_STATUS_NUM_CHANNELS_UNSUPPORTED;
        break;
    case NVIMGCODEC_SAMPLEFORMAT_P_Y:
        if (planes != 1)
            st |= NVIMGCODEC_PROCESSING_STATUS_NUM_PLANES_UNSUPPORTED;
        if (ch0 != 1)
            st |= NVIMGCODEC_PROCESSING_STATUS_NUM_CHANNELS_UNSUPPORTED;
        break;
    default: // P_YUV, unknown
        st |= NVIMGCODEC_PROCESSING_STATUS_SAMPLE_FORMAT_UNSUPPORTED;
        break;
    }
    if (info.plane_info[0].sample_type != NVIMGCODEC_SAMPLE_DATA_TYPE_UINT8)
        st |= NVIMGCODEC_PROCESSING_STATUS_SAMPLE_TYPE_UNSUPPORTED;
    // region of interest: the reference's rule (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:355-369) -- two
    // dimensions, inside the image, and the target image has the region's size
    int want_w = si.width, want_h = si.height;
    if (info.region.ndim != 0 && info.region.ndim != 2) {
        st |= NVIMGCODEC_PROCESSING_STATUS_ROI_UNSUPPORTED;
    } else if (info.region.ndim == 2) {
        const int x = info.region.start[1], y = info.region.start[0];
        want_w = info.region.end[1] - x;
        want_h = info.region.end[0] - y;
        if (x < 0 || y < 0 || want_w <= 0 || want_h <= 0 || x + want_w > si.width || y + want_h > si.height)
            st |= NVIMGCODEC_PROCESSING_STATUS_ROI_UNSUPPORTED;
    }
    // EXIF orientation is applied by the colour kernel (quarter turns and mirrors; the CPU plugin declines it,
    // libjpeg_turbo_decoder.cpp:164-165; nvjpeg's plugins apply it, cuda_decoder.cpp:443-460) -- not together with a region
    if (params->apply_exif_orientation && (info.orientation.flip_x || info.orientation.flip_y || info.orientation.rotated != 0)) {
        if (info.region.ndim != 0 || info.orientation.rotated % 90 != 0 || info.orientation.rotated < 0 || info.orientation.rotated >= 360)
            st |= NVIMGCODEC_PROCESSING_STATUS_ORIENTATION_UNSUPPORTED;
        else if ((info.orientation.rotated / 90) & 1)
            std::swap(want_w, want_h);
    }
    if (int(info.plane_info[0].width) != want_w || int(info.plane_info[0].height) != want_h)
        st |= NVIMGCODEC_PROCESSING_STATUS_ROI_UNSUPPORTED; // a resized target is not ours
    // the buffer kind is not looked at: a host buffer handed to a GPU backend is staged by the framework after canDecode
    // (src/decoder_worker.cpp:298-302, src/work.h:144-190), and decode() takes either kind anyway
    return st;
}

nvimgcodecStatus_t dec_can_decode(nvimgcodecDecoder_t decoder, nvimgcodecProcessingStatus_t* status, nvimgcodecCodeStreamDesc_t** css,
    nvimgcodecImageDesc_t** imgs, int n, const nvimgcodecDecodeParams_t* params)
{
    if (!decoder || !status || !css || !imgs || !params)
        return NVIMGCODEC_STATUS_EXTENSION_INVALID_PARAMETER;
    nvtxRangePushA("b200jpeg_can_decode");
    struct Pop { ~Pop() { nvtxRangePop(); } } pop;
    try {
        std::vector<uint8_t> storage;
        for (int i = 0; i < n; ++i) {
            storage.clear();
            status[i] = NVIMGCODEC_PROCESSING_STATUS_SUCCESS;
            nvimgcodecJpegImageInfo_t jinfo{NVIMGCODEC_STRUCTURE_TYPE_JPEG_IMAGE_INFO, sizeof(nvimgcodecJpegImageInfo_t), nullptr};
            nvimgcodecImageInfo_t cs_info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), &jinfo};
            css[i]->getImageInfo(css[i]->instance, &cs_info);
            if (std::strcmp(cs_info.codec_name, "jpeg") != 0) {
                status[i] = NVIMGCODEC_PROCESSING_STATUS_CODEC_UNSUPPORTED;
                continue;
            }
            size_t size = 0;
            const uint8_t* data = stream_bytes(css[i]->io_stream, storage, size);
            b200jpegStreamInfo_t si;
            const bool parsed = data && b200jpeg_parse_header(data, size, &si) == B200JPEG_STATUS_SUCCESS;
            if (data && storage.empty() && css[i]->io_stream->unmap) // a mapping is handed back as soon as the header is read
                css[i]->io_stream->unmap(css[i]->io_stream->instance, const_cast<uint8_t*>(data), size);
            if (!parsed) {
                status[i] = NVIMGCODEC_PROCESSING_STATUS_FAIL;
                continue;
            }
            if (!si.supported) {
                // progressive / arithmetic / lossless => ENCODING_UNSUPPORTED, everything else (CMYK, 4:1:0, 12-bit,
                // multi-scan ...) => CODESTREAM_UNSUPPORTED; both send the sample to the next decoder
                status[i] = (si.sof_marker != 0xC0 && si.sof_marker != 0xC1) ? NVIMGCODEC_PROCESSING_STATUS_ENCODING_UNSUPPORTED
                                                                             : NVIMGCODEC_PROCESSING_STATUS_CODESTREAM_UNSUPPORTED;
                continue;
            }
            nvimgcodecImageInfo_t info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), nullptr};
            imgs[i]->getImageInfo(imgs[i]->instance, &info);
            status[i] = check_target(info, si, params);
        }
    } catch (...) {
        return NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
    }
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t dec_decode(nvimgcodecDecoder_t decoder, nvimgcodecCodeStreamDesc_t** css, nvimgcodecImageDesc_t** imgs, int n,
    const nvimgcodecDecodeParams_t* params)
{
    auto* st = reinterpret_cast<DecoderState*>(decoder);
    if (!st || !css || !imgs || n < 0)
        return NVIMGCODEC_STATUS_EXTENSION_INVALID_PARAMETER;
    nvtxRangePushA("b200jpeg_decode_batch");
    struct Pop { ~Pop() { nvtxRangePop(); } } pop;
    std::vector<char> reported(n, 0);
    auto report = [&](int i, nvimgcodecProcessingStatus_t s) {
        if (!reported[i]) { // imageReady exactly once per sample (src/processing_results.cpp:104-115)
            reported[i] = 1;
            imgs[i]->imageReady(imgs[i]->instance, s);
        }
    };
    try {
        std::vector<std::vector<uint8_t>> storage(n);
        std::vector<const uint8_t*> data(n, nullptr);
        std::vector<size_t> sizes(n, 0);
        std::vector<b200jpegOutput_t> outs(n);
        for (int i = 0; i < n; ++i) {
            data[i] = stream_bytes(css[i]->io_stream, storage[i], sizes[i]);
            nvimgcodecImageInfo_t info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), nullptr};
            imgs[i]->getImageInfo(imgs[i]->instance, &info);
            outs[i] = b200jpegOutput_t{};
            outs[i].buffer = info.buffer;
            outs[i].row_stride = int64_t(info.plane_info[0].row_stride);
            outs[i].format = int32_t(info.sample_format);
            outs[i].on_host = info.buffer_kind == NVIMGCODEC_IMAGE_BUFFER_KIND_STRIDED_HOST ? 1 : 0;
            if (info.region.ndim == 2) {
                outs[i].roi_x = info.region.start[1];
                outs[i].roi_y = info.region.start[0];
                outs[i].roi_width = info.region.end[1] - info.region.start[1];
                outs[i].roi_height = info.region.end[0] - info.region.start[0];
            }
            if (params && params->apply_exif_orientation)
                outs[i].orientation = ((info.orientation.rotated / 90) & 3) | (info.orientation.flip_x ? 4 : 0) | (info.orientation.flip_y ? 8 : 0);
            outs[i].cuda_stream = info.cuda_stream; // every distinct stream is waited for, and made to wait for the decode
            outs[i].has_cuda_stream = 1;
            if (!data[i]) {
                static const uint8_t empty[2] = {0, 0};
                data[i] = empty;
                sizes[i] = 0;
            }
        }
        auto unmap_all = [&] { // after submit the bytes sit in the library's pinned staging
            for (int i = 0; i < n; ++i)
                if (css[i]->io_stream && css[i]->io_stream->unmap && storage[i].empty() && sizes[i])
                    css[i]->io_stream->unmap(css[i]->io_stream->instance, const_cast<uint8_t*>(data[i]), sizes[i]);
        };
        std::vector<int32_t> res(n > 0 ? n : 1, B200JPEG_STATUS_INTERNAL_ERROR);
        int rc;
        if (st->async) {
            {
                std::unique_lock<std::mutex> lk(st->mu);
                st->cv.wait(lk, [&] { return st->queue.size() < 2; }); // the library takes two batches in flight
            }
            rc = b200jpeg_decode_submit(st->dec, n, data.data(), sizes.data(), outs.data(), nullptr, res.data());
            unmap_all();
            if (rc == B200JPEG_STATUS_SUCCESS) {
                DecoderState::Pending p;
                p.imgs.assign(imgs, imgs + n);
                for (int i = 0; i < n; ++i)
                    reported[i] = 1; // from here on the completion thread reports
                {
                    std::lock_guard<std::mutex> lk(st->mu);
                    st->queue.push_back(std::move(p));
                }
                st->cv.notify_all();
                return NVIMGCODEC_STATUS_SUCCESS;
            }
        } else {
            rc = b200jpeg_decode_batch(st->dec, n, data.data(), sizes.data(), outs.data(), nullptr, res.data());
            unmap_all();
            // b200jpeg_decode_batch returns with the device work complete
        }
        for (int i = 0; i < n; ++i)
            report(i, to_processing_status(rc, res[i]));
        if (rc != B200JPEG_STATUS_SUCCESS) {
            log_msg(st->fw, NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_ERROR, kDecoderId, std::string("decode batch failed: ") + b200jpeg_status_string(rc));
            return rc == B200JPEG_STATUS_CUDA_ERROR ? NVIMGCODEC_STATUS_EXTENSION_CUDA_CALL_ERROR : NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
        }
    } catch (...) {
        for (int i = 0; i < n; ++i)
            report(i, NVIMGCODEC_PROCESSING_STATUS_FAIL);
        return NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
    }
    (void)params;
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t dec_destroy(nvimgcodecDecoder_t decoder)
{
    auto* st = reinterpret_cast<DecoderState*>(decoder);
    if (!st)
        return NVIMGCODEC_STATUS_EXTENSION_INVALID_PARAMETER;
    st->shutdown(); // reports whatever is still in flight
    if (st->dec)
        b200jpeg_decoder_destroy(st->dec);
    delete st;
    return NVIMGCODEC_STATUS_SUCCESS;
}


// ---------------------------------------------------------------------------------------------------------------
// Encoder (interface precedent: extensions/nvjpeg/cuda_encoder.cpp:54-147 canEncode, :284-396 encode)
// ---------------------------------------------------------------------------------------------------------------
struct EncoderState
{
    const nvimgcodecFrameworkDesc_t* fw = nullptr;
    const nvimgcodecExecutionParams_t* exec = nullptr;
    b200jpegEncoder_t enc = nullptr;
    // where the library puts the files before they are handed to the code streams: one page-locked allocation kept across
    // calls (the device -> host copies go straight into it), never cleared
    uint8_t* bounce = nullptr;
    size_t bounce_cap = 0;
    bool bounce_pinned = false;
    uint8_t* ensure_bounce(size_t bytes)
    {
        if (bytes <= bounce_cap)
            return bounce;
        release_bounce();
        const size_t want = bytes + bytes / 4;
        void* p = nullptr;
        if (cudaHostAlloc(&p, want, cudaHostAllocDefault) == cudaSuccess) {
            bounce_pinned = true;
        } else {
            cudaGetLastError();
            p = std::malloc(want);
            bounce_pinned = false;
        }
        bounce = static_cast<uint8_t*>(p);
        bounce_cap = p ? want : 0;
        return bounce;
    }
    void release_bounce()
    {
        if (bounce) {
            if (bounce_pinned)
                cudaFreeHost(bounce);
            else
                std::free(bounce);
        }
        bounce = nullptr;
        bounce_cap = 0;
    }
};

// what the caller wants written: baseline only, standard Huffman tables only
bool wants_out_of_scope_jpeg(const nvimgcodecImageInfo_t& cs_info, const nvimgcodecEncodeParams_t* params)
{
    for (auto* p = static_cast<const nvimgcodecJpegImageInfo_t*>(cs_info.struct_next); p; p = static_cast<const nvimgcodecJpegImageInfo_t*>(p->struct_next))
        if (p->struct_type == NVIMGCODEC_STRUCTURE_TYPE_JPEG_IMAGE_INFO && p->encoding != NVIMGCODEC_JPEG_ENCODING_BASELINE_DCT &&
            p->encoding != NVIMGCODEC_JPEG_ENCODING_UNKNOWN)
            return true;
    for (auto* p = static_cast<const nvimgcodecJpegEncodeParams_t*>(params->struct_next); p; p = static_cast<const nvimgcodecJpegEncodeParams_t*>(p->struct_next))
        if (p->struct_type == NVIMGCODEC_STRUCTURE_TYPE_JPEG_ENCODE_PARAMS && p->optimized_huffman)
            return true;
    return false;
}

nvimgcodecProcessingStatus_t check_source(const nvimgcodecImageInfo_t& info, const nvimgcodecImageInfo_t& out_info)
{
    nvimgcodecProcessingStatus_t st = NVIMGCODEC_PROCESSING_STATUS_SUCCESS;
    const bool gray = info.sample_format == NVIMGCODEC_SAMPLEFORMAT_P_Y;
    switch (info.sample_format) {
    case NVIMGCODEC_SAMPLEFORMAT_I_RGB:
        if (info.num_planes != 1 || info.plane_info[0].num_channels != 3)
            st |= NVIMGCODEC_PROCESSING_STATUS_NUM_CHANNELS_UNSUPPORTED;
        break;
    case NVIMGCODEC_SAMPLEFORMAT_P_RGB:
        if (info.num_planes != 3 || info.plane_info[0].num_channels != 1)
            st |= NVIMGCODEC_PROCESSING_STATUS_NUM_PLANES_UNSUPPORTED;
        break;
    case NVIMGCODEC_SAMPLEFORMAT_P_Y:
        if (info.num_planes != 1 || info.plane_info[0].num_channels != 1)
            st |= NVIMGCODEC_PROCESSING_STATUS_NUM_CHANNELS_UNSUPPORTED;
        break;
    default:
        st |= NVIMGCODEC_PROCESSING_STATUS_SAMPLE_FORMAT_UNSUPPORTED;
        break;
    }
    for (uint32_t p = 0; p < info.num_planes; ++p)
        if (info.plane_info[p].sample_type != NVIMGCODEC_SAMPLE_DATA_TYPE_UINT8)
            st |= NVIMGCODEC_PROCESSING_STATUS_SAMPLE_TYPE_UNSUPPORTED;
    switch (out_info.chroma_subsampling) {
    case NVIMGCODEC_SAMPLING_444:
    case NVIMGCODEC_SAMPLING_422:
    case NVIMGCODEC_SAMPLING_420:
    case NVIMGCODEC_SAMPLING_440:
        if (gray)
            st |= NVIMGCODEC_PROCESSING_STATUS_SAMPLING_UNSUPPORTED;
        break;
    case NVIMGCODEC_SAMPLING_GRAY:
        if (!gray)
            st |= NVIMGCODEC_PROCESSING_STATUS_SAMPLING_UNSUPPORTED; // RGB -> gray conversion is not offered
        break;
    default:
        st |= NVIMGCODEC_PROCESSING_STATUS_SAMPLING_UNSUPPORTED;
        break;
    }
    if (info.color_spec != NVIMGCODEC_COLORSPEC_UNCHANGED && info.color_spec != NVIMGCODEC_COLORSPEC_SRGB &&
        !(gray && (info.color_spec == NVIMGCODEC_COLORSPEC_GRAY || info.color_spec == NVIMGCODEC_COLORSPEC_SYCC)))
        st |= NVIMGCODEC_PROCESSING_STATUS_COLOR_SPEC_UNSUPPORTED;
    if (info.buffer_kind != NVIMGCODEC_IMAGE_BUFFER_KIND_STRIDED_DEVICE && info.buffer_kind != NVIMGCODEC_IMAGE_BUFFER_KIND_STRIDED_HOST)
        st |= NVIMGCODEC_PROCESSING_STATUS_BACKEND_UNSUPPORTED;
    return st;
}

nvimgcodecStatus_t enc_can_encode(nvimgcodecEncoder_t encoder, nvimgcodecProcessingStatus_t* status, nvimgcodecImageDesc_t** imgs,
    nvimgcodecCodeStreamDesc_t** css, int n, const nvimgcodecEncodeParams_t* params)
{
    if (!encoder || !status || !css || !imgs || !params)
        return NVIMGCODEC_STATUS_EXTENSION_INVALID_PARAMETER;
    try {
        for (int i = 0; i < n; ++i) {
            // an output code stream has no parser: ask without a chained structure and look at the chain that comes back
            // (src/code_stream.cpp:75-98; same idiom as extensions/nvjpeg/cuda_encoder.cpp:64-83)
            nvimgcodecImageInfo_t cs_info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), nullptr};
            css[i]->getImageInfo(css[i]->instance, &cs_info);
            if (std::strcmp(cs_info.codec_name, "jpeg") != 0) {
                status[i] = NVIMGCODEC_PROCESSING_STATUS_CODEC_UNSUPPORTED;
                continue;
            }
            if (wants_out_of_scope_jpeg(cs_info, params)) {
                status[i] = NVIMGCODEC_PROCESSING_STATUS_ENCODING_UNSUPPORTED;
                continue;
            }
            nvimgcodecImageInfo_t info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), nullptr};
            imgs[i]->getImageInfo(imgs[i]->instance, &info);
            status[i] = check_source(info, cs_info);
        }
    } catch (...) {
        return NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
    }
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t enc_encode(nvimgcodecEncoder_t encoder, nvimgcodecImageDesc_t** imgs, nvimgcodecCodeStreamDesc_t** css, int n,
    const nvimgcodecEncodeParams_t* params)
{
    auto* st = reinterpret_cast<EncoderState*>(encoder);
    if (!st || !css || !imgs || !params || n < 0)
        return NVIMGCODEC_STATUS_EXTENSION_INVALID_PARAMETER;
    std::vector<char> reported(n, 0);
    auto report = [&](int i, nvimgcodecProcessingStatus_t s) {
        if (!reported[i]) {
            reported[i] = 1;
            imgs[i]->imageReady(imgs[i]->instance, s);
        }
    };
    try {
        std::vector<b200jpegEncodeInput_t> in(n);
        std::vector<uint8_t*> outs(n);
        std::vector<size_t> offs(n + 1, 0);
        std::vector<size_t> caps(n), sizes(n, 0);
        std::vector<int32_t> res(n, B200JPEG_STATUS_INTERNAL_ERROR);
        cudaStream_t user_stream = nullptr;
        const int quality = std::max(1, std::min(100, int(params->quality + 0.5f)));
        for (int i = 0; i < n; ++i) {
            nvimgcodecImageInfo_t info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), nullptr};
            imgs[i]->getImageInfo(imgs[i]->instance, &info);
            nvimgcodecImageInfo_t cs_info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), nullptr};
            css[i]->getImageInfo(css[i]->instance, &cs_info);
            in[i].buffer = info.buffer;
            in[i].row_stride = int64_t(info.plane_info[0].row_stride);
            in[i].width = int32_t(info.plane_info[0].width);
            in[i].height = int32_t(info.plane_info[0].height);
            in[i].format = int32_t(info.sample_format);
            in[i].on_host = info.buffer_kind == NVIMGCODEC_IMAGE_BUFFER_KIND_STRIDED_HOST ? 1 : 0;
            in[i].subsampling = int32_t(cs_info.chroma_subsampling);
            in[i].quality = quality;
            in[i].restart_interval = 0;
            if (i == 0)
                user_stream = info.cuda_stream;
            // a baseline file never needs more than about the raw size (16 bits per coefficient would be 2x; quality 100 noise
            // stays under 1.3x): twice the samples plus the headers
            caps[i] = size_t(in[i].width) * in[i].height * (in[i].format == B200JPEG_FMT_P_Y ? 1 : 3) * 2 + 4096;
            offs[i + 1] = offs[i] + ((caps[i] + 255) & ~size_t(255));
        }
        uint8_t* bounce = st->ensure_bounce(offs[n]);
        if (!bounce && n > 0) {
            for (int i = 0; i < n; ++i)
                report(i, NVIMGCODEC_PROCESSING_STATUS_FAIL);
            return NVIMGCODEC_STATUS_EXTENSION_ALLOCATOR_FAILURE;
        }
        for (int i = 0; i < n; ++i)
            outs[i] = bounce + offs[i];
        const int rc = b200jpeg_encode_batch(st->enc, n, in.data(), outs.data(), caps.data(), sizes.data(), user_stream, res.data());
        for (int i = 0; i < n; ++i) {
            if (rc != B200JPEG_STATUS_SUCCESS || res[i] != B200JPEG_STATUS_SUCCESS) {
                report(i, NVIMGCODEC_PROCESSING_STATUS_FAIL);
                continue;
            }
            // hand the file to the code stream (idiom of extensions/nvjpeg/cuda_encoder.cpp:383-388)
            nvimgcodecIoStreamDesc_t* io = css[i]->io_stream;
            size_t written = 0;
            io->reserve(io->instance, sizes[i]);
            io->seek(io->instance, 0, SEEK_SET);
            io->write(io->instance, &written, outs[i], sizes[i]);
            io->flush(io->instance);
            report(i, written == sizes[i] ? NVIMGCODEC_PROCESSING_STATUS_SUCCESS : NVIMGCODEC_PROCESSING_STATUS_FAIL);
        }
        if (rc != B200JPEG_STATUS_SUCCESS) {
            log_msg(st->fw, NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_ERROR, kEncoderId, std::string("encode batch failed: ") + b200jpeg_status_string(rc));
            return rc == B200JPEG_STATUS_CUDA_ERROR ? NVIMGCODEC_STATUS_EXTENSION_CUDA_CALL_ERROR : NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
        }
    } catch (...) {
        for (int i = 0; i < n; ++i)
            report(i, NVIMGCODEC_PROCESSING_STATUS_FAIL);
        return NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
    }
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t enc_destroy(nvimgcodecEncoder_t encoder)
{
    auto* st = reinterpret_cast<EncoderState*>(encoder);
    if (!st)
        return NVIMGCODEC_STATUS_EXTENSION_INVALID_PARAMETER;
    if (st->enc)
        b200jpeg_encoder_destroy(st->enc);
    st->release_bounce();
    delete st;
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t enc_create(void* instance, nvimgcodecEncoder_t* encoder, const nvimgcodecExecutionParams_t* exec, const char* options);

struct Extension;
nvimgcodecStatus_t dec_create(void* instance, nvimgcodecDecoder_t* decoder, const nvimgcodecExecutionParams_t* exec, const char* options);

// ---------------------------------------------------------------------------------------------------------------
// Extension object: owns the descriptors (the framework keeps pointers to them, src/image_decoder_factory.cpp:26-29)
// ---------------------------------------------------------------------------------------------------------------
struct Extension
{
    const nvimgcodecFrameworkDesc_t* fw;
    nvimgcodecDecoderDesc_t dec_desc;
    nvimgcodecEncoderDesc_t enc_desc;
    bool dec_registered = false, enc_registered = false;

    static void lib_log(void* ctx, int severity, const char* message)
    {
        auto* fw = static_cast<const nvimgcodecFrameworkDesc_t*>(ctx);
        log_msg(fw, severity >= 2 ? NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_ERROR : severity == 1 ? NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_WARNING
                                                                                             : NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_INFO,
            "b200jpeg", message);
    }

    explicit Extension(const nvimgcodecFrameworkDesc_t* f) : fw(f)
    {
        b200jpeg_set_log_callback(&Extension::lib_log, const_cast<nvimgcodecFrameworkDesc_t*>(fw));
        dec_desc = nvimgcodecDecoderDesc_t{NVIMGCODEC_STRUCTURE_TYPE_DECODER_DESC, sizeof(nvimgcodecDecoderDesc_t), nullptr, this, kDecoderId,
            "jpeg", NVIMGCODEC_BACKEND_KIND_GPU_ONLY, dec_create, dec_destroy, dec_can_decode, dec_decode};
        if (fw->registerDecoder(fw->instance, &dec_desc, kPriority) == NVIMGCODEC_STATUS_SUCCESS)
            dec_registered = true;
        enc_desc = nvimgcodecEncoderDesc_t{NVIMGCODEC_STRUCTURE_TYPE_ENCODER_DESC, sizeof(nvimgcodecEncoderDesc_t), nullptr, this, kEncoderId,
            "jpeg", NVIMGCODEC_BACKEND_KIND_GPU_ONLY, enc_create, enc_destroy, enc_can_encode, enc_encode};
        if (fw->registerEncoder(fw->instance, &enc_desc, kPriority) == NVIMGCODEC_STATUS_SUCCESS)
            enc_registered = true;
    }
    ~Extension()
    {
        b200jpeg_set_log_callback(nullptr, nullptr);
        if (enc_registered)
            fw->unregisterEncoder(fw->instance, &enc_desc);
        if (dec_registered)
            fw->unregisterDecoder(fw->instance, &dec_desc);
    }
};

// the framework's allocators (include/nvimgcodec.h:232-302) in the C ABI's form; cudaStream_t and void* are the same at the ABI level
b200jpegAllocators_t exec_allocators(const nvimgcodecExecutionParams_t* exec)
{
    b200jpegAllocators_t a{};
    if (exec->device_allocator && exec->device_allocator->device_malloc && exec->device_allocator->device_free) {
        a.device_malloc = reinterpret_cast<b200jpegMalloc_t>(exec->device_allocator->device_malloc);
        a.device_free = reinterpret_cast<b200jpegFree_t>(exec->device_allocator->device_free);
        a.device_ctx = exec->device_allocator->device_ctx;
        a.device_padding = exec->device_allocator->device_mem_padding;
    }
    if (exec->pinned_allocator && exec->pinned_allocator->pinned_malloc && exec->pinned_allocator->pinned_free) {
        a.pinned_malloc = reinterpret_cast<b200jpegMalloc_t>(exec->pinned_allocator->pinned_malloc);
        a.pinned_free = reinterpret_cast<b200jpegFree_t>(exec->pinned_allocator->pinned_free);
        a.pinned_ctx = exec->pinned_allocator->pinned_ctx;
        a.pinned_padding = exec->pinned_allocator->pinned_mem_padding;
    }
    return a;
}

nvimgcodecStatus_t dec_create(void* instance, nvimgcodecDecoder_t* decoder, const nvimgcodecExecutionParams_t* exec, const char* options)
{
    auto* ext = reinterpret_cast<Extension*>(instance);
    if (!ext || !decoder || !exec)
        return NVIMGCODEC_STATUS_EXTENSION_INVALID_PARAMETER;
    if (exec->device_id == NVIMGCODEC_DEVICE_CPU_ONLY)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER; // like extensions/nvjpeg/cuda_decoder.cpp:272-273
    auto st = std::make_unique<DecoderState>();
    st->fw = ext->fw;
    st->exec = exec;
    const int rc = b200jpeg_decoder_create(&st->dec, exec->device_id, options ? options : "");
    if (rc != B200JPEG_STATUS_SUCCESS) { // the framework then treats the decoder as absent and falls back (src/image_decoder.cpp:29-36,62-68)
        log_msg(ext->fw, NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_WARNING, kDecoderId, std::string("decoder unavailable: ") + b200jpeg_status_string(rc));
        if (rc == B200JPEG_STATUS_UNSUPPORTED)
            return NVIMGCODEC_STATUS_EXTENSION_IMPLEMENTATION_NOT_SUPPORTED; // an option asks for what is not offered (fast_idct=1)
        return rc == B200JPEG_STATUS_ARCH_MISMATCH ? NVIMGCODEC_STATUS_EXTENSION_ARCH_MISMATCH : NVIMGCODEC_STATUS_EXTENSION_CUDA_CALL_ERROR;
    }
    const b200jpegAllocators_t al = exec_allocators(exec);
    b200jpeg_decoder_set_allocators(st->dec, &al);
    // "async=0" (optionally prefixed like every option): decode() returns only when the device is done
    st->async = !(options && std::strstr(options, "async=0"));
    if (st->async)
        st->completer = std::thread([p = st.get()] { p->run(); });
    *decoder = reinterpret_cast<nvimgcodecDecoder_t>(st.release());
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t enc_create(void* instance, nvimgcodecEncoder_t* encoder, const nvimgcodecExecutionParams_t* exec, const char* options)
{
    auto* ext = reinterpret_cast<Extension*>(instance);
    if (!ext || !encoder || !exec)
        return NVIMGCODEC_STATUS_EXTENSION_INVALID_PARAMETER;
    if (exec->device_id == NVIMGCODEC_DEVICE_CPU_ONLY)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER; // like extensions/nvjpeg/cuda_encoder.cpp:203-204
    auto st = std::make_unique<EncoderState>();
    st->fw = ext->fw;
    st->exec = exec;
    const int rc = b200jpeg_encoder_create(&st->enc, exec->device_id, options ? options : "");
    if (rc != B200JPEG_STATUS_SUCCESS) {
        log_msg(ext->fw, NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_WARNING, kEncoderId, std::string("encoder unavailable: ") + b200jpeg_status_string(rc));
        return rc == B200JPEG_STATUS_ARCH_MISMATCH ? NVIMGCODEC_STATUS_EXTENSION_ARCH_MISMATCH : NVIMGCODEC_STATUS_EXTENSION_CUDA_CALL_ERROR;
    }
    const b200jpegAllocators_t al = exec_allocators(exec);
    b200jpeg_encoder_set_allocators(st->enc, &al);
    *encoder = reinterpret_cast<nvimgcodecEncoder_t>(st.release());
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t ext_create(void* instance, nvimgcodecExtension_t* extension, const nvimgcodecFrameworkDesc_t* framework)
{
    (void)instance;
    if (!extension || !framework)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER;
    try {
        *extension = reinterpret_cast<nvimgcodecExtension_t>(new Extension(framework));
    } catch (...) {
        return NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
    }
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t ext_destroy(nvimgcodecExtension_t extension)
{
    if (!extension)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER;
    delete reinterpret_cast<Extension*>(extension);
    return NVIMGCODEC_STATUS_SUCCESS;
}

// clang-format off
const nvimgcodecExtensionDesc_t kExtensionDesc = {
    NVIMGCODEC_STRUCTURE_TYPE_EXTENSION_DESC,
    sizeof(nvimgcodecExtensionDesc_t),
    nullptr,
    nullptr,
    "b200jpeg_extension",
    NVIMGCODEC_VER,
    NVIMGCODEC_EXT_API_VER,
    ext_create,
    ext_destroy
};
// clang-format on

} // namespace

extern "C" {

// static-registration seam used by tests (nvimgcodecExtensionCreate, src/nvimgcodec_capi.cpp:224-243)
__attribute__((visibility("default"))) nvimgcodecStatus_t get_b200jpeg_extension_desc(nvimgcodecExtensionDesc_t* ext_desc)
{
    if (!ext_desc || ext_desc->struct_type != NVIMGCODEC_STRUCTURE_TYPE_EXTENSION_DESC)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER;
    *ext_desc = kExtensionDesc;
    return NVIMGCODEC_STATUS_SUCCESS;
}

__attribute__((visibility("default"))) nvimgcodecStatus_t nvimgcodecExtensionModuleEntry(nvimgcodecExtensionDesc_t* ext_desc)
{
    return get_b200jpeg_extension_desc(ext_desc);
}

} // extern "C"
