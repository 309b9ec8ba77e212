// oracle/harness.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT.
//
// A small C-ABI driver (for ctypes) around the *unmodified* reference, compiled into
// oracle/_ref/libnvimgcodec_ref.so together with the reference core, the reference's
// libjpeg_turbo_ext and libjpeg-turbo 3.0.3.  It does three things:
//   1. refh_*        : drive the public nvimgcodec C API (include/nvimgcodec.h:1364-1706) exactly the way
//                      a user / the Python bindings do (python/decoder.cpp:148-252): instance -> decoder ->
//                      code streams from host memory -> images -> nvimgcodecDecoderDecode -> future.
//                      The CPU oracle is reached with backends=[CPU_ONLY]; the b200jpeg plugin is reached
//                      by pointing extension_modules_path at the directory holding libb200jpeg_ext.so.
//   2. refh_ljt_*    : call libjpeg-turbo directly to *make* the synthetic corpus (jpeg_compress with exact
//                      control of sampling / quality / DRI / tables) and to dump quantised DCT coefficients
//                      (jpeg_read_coefficients) for the "down to the DCT coefficients" parity bar.
//   3. wall-clock timing of the decode call (CLOCK_MONOTONIC, like example/nvimtrans/main.cpp:84-85).
#include <nvimgcodec.h>
#include <cuda_runtime_api.h>
#include <stdio.h>
#include <string.h>
#include <setjmp.h>
#include <time.h>
#include <vector>
#include <string>

extern "C" {
#include <jpeglib.h>
}

#include "libjpeg_turbo_ext.h"

#define REFH_API extern "C" __attribute__((visibility("default")))

namespace {

double now_s()
{
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

struct Session
{
    nvimgcodecInstance_t instance = nullptr;
    nvimgcodecExtension_t ljt_ext = nullptr;
    std::string ext_path;
};

struct Decoder
{
    Session* s = nullptr;
    nvimgcodecDecoder_t dec = nullptr;
    std::vector<nvimgcodecBackend_t> backends;
    nvimgcodecExecutionParams_t exec{};
};

struct Encoder
{
    Session* s = nullptr;
    nvimgcodecEncoder_t enc = nullptr;
    std::vector<nvimgcodecBackend_t> backends;
    nvimgcodecExecutionParams_t exec{};
};

int channels_of(int sample_format, int fallback)
{
    switch (sample_format) {
    case NVIMGCODEC_SAMPLEFORMAT_P_Y:
        return 1;
    case NVIMGCODEC_SAMPLEFORMAT_I_RGB:
    case NVIMGCODEC_SAMPLEFORMAT_I_BGR:
    case NVIMGCODEC_SAMPLEFORMAT_P_RGB:
    case NVIMGCODEC_SAMPLEFORMAT_P_BGR:
        return 3;
    default:
        return fallback;
    }
}

bool is_planar(int sample_format)
{
    return sample_format == NVIMGCODEC_SAMPLEFORMAT_P_RGB || sample_format == NVIMGCODEC_SAMPLEFORMAT_P_BGR ||
           sample_format == NVIMGCODEC_SAMPLEFORMAT_P_Y || sample_format == NVIMGCODEC_SAMPLEFORMAT_P_UNCHANGED ||
           sample_format == NVIMGCODEC_SAMPLEFORMAT_P_YUV;
}

void fill_image_info(nvimgcodecImageInfo_t& info, int width, int height, int sample_format, int64_t row_stride, int buffer_kind,
    void* buffer, void* stream, int unchanged_channels)
{
    memset(&info, 0, sizeof(info));
    info.struct_type = NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO;
    info.struct_size = sizeof(info);
    info.sample_format = (nvimgcodecSampleFormat_t)sample_format;
    int ch = channels_of(sample_format, unchanged_channels);
    info.color_spec = ch == 1 ? NVIMGCODEC_COLORSPEC_GRAY : NVIMGCODEC_COLORSPEC_SRGB;
    info.chroma_subsampling = ch == 1 ? NVIMGCODEC_SAMPLING_GRAY : NVIMGCODEC_SAMPLING_NONE;
    bool planar = is_planar(sample_format);
    info.num_planes = planar ? ch : 1;
    for (uint32_t p = 0; p < info.num_planes; ++p) {
        info.plane_info[p].struct_type = NVIMGCODEC_STRUCTURE_TYPE_IMAGE_PLANE_INFO;
        info.plane_info[p].struct_size = sizeof(nvimgcodecImagePlaneInfo_t);
        info.plane_info[p].width = width;
        info.plane_info[p].height = height;
        info.plane_info[p].num_channels = planar ? 1 : ch;
        info.plane_info[p].row_stride = row_stride;
        info.plane_info[p].sample_type = NVIMGCODEC_SAMPLE_DATA_TYPE_UINT8;
        info.plane_info[p].precision = 0;
    }
    info.buffer = buffer;
    info.buffer_size = (size_t)row_stride * height * info.num_planes;
    info.buffer_kind = (nvimgcodecImageBufferKind_t)buffer_kind;
    info.cuda_stream = (cudaStream_t)stream;
}

} // namespace

// ------------------------------------------------------------------------------------------------
// 1. nvimgcodec public API driver
// ------------------------------------------------------------------------------------------------

// ext_dir: directory scanned for extension modules (NULL/"" => only builtin parsers + static libjpeg_turbo_ext).
// with_ljt: register the reference libjpeg_turbo_ext statically (nvimgcodecExtensionCreate, the seam used by
//           test/extensions/libjpeg_turbo_ext_decoder_test.cpp:42-50).
REFH_API void* refh_create(const char* ext_dir, int with_ljt, int verbose)
{
    auto* s = new Session();
    nvimgcodecInstanceCreateInfo_t ci{NVIMGCODEC_STRUCTURE_TYPE_INSTANCE_CREATE_INFO, sizeof(nvimgcodecInstanceCreateInfo_t), 0};
    ci.load_builtin_modules = 1;
    if (ext_dir && ext_dir[0]) {
        s->ext_path = ext_dir;
        ci.load_extension_modules = 1;
        ci.extension_modules_path = s->ext_path.c_str();
    }
    ci.create_debug_messenger = verbose ? 1 : 0;
    ci.message_severity = verbose > 1 ? NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_ALL : NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_DEFAULT;
    ci.message_category = NVIMGCODEC_DEBUG_MESSAGE_CATEGORY_ALL;
    if (nvimgcodecInstanceCreate(&s->instance, &ci) != NVIMGCODEC_STATUS_SUCCESS) {
        delete s;
        return nullptr;
    }
    if (with_ljt) {
        nvimgcodecExtensionDesc_t d{NVIMGCODEC_STRUCTURE_TYPE_EXTENSION_DESC, sizeof(nvimgcodecExtensionDesc_t), 0};
        get_libjpeg_turbo_extension_desc(&d);
        if (nvimgcodecExtensionCreate(s->instance, &s->ljt_ext, &d) != NVIMGCODEC_STATUS_SUCCESS) {
            nvimgcodecInstanceDestroy(s->instance);
            delete s;
            return nullptr;
        }
    }
    return s;
}

REFH_API void refh_destroy(void* h)
{
    auto* s = (Session*)h;
    if (!s)
        return;
    if (s->ljt_ext)
        nvimgcodecExtensionDestroy(s->ljt_ext);
    nvimgcodecInstanceDestroy(s->instance);
    delete s;
}

// backend_mask: bit (kind-1) set => allow nvimgcodecBackendKind_t kind; 0 => all backends.
REFH_API void* refh_decoder_create(void* h, int device_id, int backend_mask, int num_threads, const char* options)
{
    auto* d = new Decoder();
    d->s = (Session*)h;
    for (int k = 1; k <= 4; ++k)
        if (backend_mask & (1 << (k - 1))) {
            nvimgcodecBackend_t b{NVIMGCODEC_STRUCTURE_TYPE_BACKEND, sizeof(nvimgcodecBackend_t), 0};
            b.kind = (nvimgcodecBackendKind_t)k;
            b.params = {NVIMGCODEC_STRUCTURE_TYPE_BACKEND_PARAMS, sizeof(nvimgcodecBackendParams_t), 0, 1.0f};
            d->backends.push_back(b);
        }
    d->exec = {NVIMGCODEC_STRUCTURE_TYPE_EXECUTION_PARAMS, sizeof(nvimgcodecExecutionParams_t), 0};
    d->exec.device_id = device_id;
    d->exec.max_num_cpu_threads = num_threads;
    d->exec.num_backends = (int)d->backends.size();
    d->exec.backends = d->backends.empty() ? nullptr : d->backends.data();
    if (nvimgcodecDecoderCreate(d->s->instance, &d->dec, &d->exec, options) != NVIMGCODEC_STATUS_SUCCESS) {
        delete d;
        return nullptr;
    }
    return d;
}

REFH_API void refh_decoder_destroy(void* hd)
{
    auto* d = (Decoder*)hd;
    if (!d)
        return;
    nvimgcodecDecoderDestroy(d->dec);
    delete d;
}

// Probe a code stream with the reference's builtin parser (src/parsers/jpeg.cpp:202-361).
// out[0..5] = width, height, num_planes, channels(plane0), chroma_subsampling, jpeg encoding (SOF marker)
REFH_API int refh_get_info(void* h, const unsigned char* data, size_t size, int* out)
{
    auto* s = (Session*)h;
    nvimgcodecCodeStream_t cs = nullptr;
    if (nvimgcodecCodeStreamCreateFromHostMem(s->instance, &cs, data, size) != NVIMGCODEC_STATUS_SUCCESS)
        return -1;
    nvimgcodecJpegImageInfo_t ji{NVIMGCODEC_STRUCTURE_TYPE_JPEG_IMAGE_INFO, sizeof(nvimgcodecJpegImageInfo_t), 0};
    nvimgcodecImageInfo_t info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), &ji};
    int rc = nvimgcodecCodeStreamGetImageInfo(cs, &info);
    if (rc == NVIMGCODEC_STATUS_SUCCESS) {
        out[0] = info.plane_info[0].width;
        out[1] = info.plane_info[0].height;
        out[2] = info.num_planes;
        out[3] = info.plane_info[0].num_channels;
        out[4] = info.chroma_subsampling;
        out[5] = ji.encoding;
    }
    nvimgcodecCodeStreamDestroy(cs);
    return rc;
}

// Decode a batch through nvimgcodecDecoderDecode (src/nvimgcodec_capi.cpp:430-455).
//   buffer_kind : 1 = STRIDED_DEVICE, 2 = STRIDED_HOST (include/nvimgcodec.h:464-471)
//   reps        : the decode call is repeated `reps` times on the same streams/images; seconds[r] gets the
//                 wall time of Decode + FutureWaitForAll (+ cudaStreamSynchronize for device buffers)
//   status_out  : per-sample nvimgcodecProcessingStatus_t of the last repetition
static int decode_impl(void* hd, int n, const unsigned char** data, const size_t* sizes, void** buffers, const int* widths,
    const int* heights, const int* sample_formats, const int64_t* row_strides, int buffer_kind, void* cuda_stream, int reps,
    unsigned int* status_out, double* seconds, const int* rois, const int* orient = nullptr, int apply_exif = 0)
{
    auto* d = (Decoder*)hd;
    auto inst = d->s->instance;
    std::vector<nvimgcodecCodeStream_t> cs(n, nullptr);
    std::vector<nvimgcodecImage_t> img(n, nullptr);
    int rc = 0;
    for (int i = 0; i < n && rc == 0; ++i) {
        if (nvimgcodecCodeStreamCreateFromHostMem(inst, &cs[i], data[i], sizes[i]) != NVIMGCODEC_STATUS_SUCCESS) {
            rc = -1;
            break;
        }
        nvimgcodecImageInfo_t csinfo{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), 0};
        // *_UNCHANGED means 3-channel RGB in the reference plugin even for gray streams
        // (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:336-345,400-412), so the buffer always has 3 channels
        const int unchanged_channels = 3;
        nvimgcodecCodeStreamGetImageInfo(cs[i], &csinfo);
        nvimgcodecImageInfo_t info;
        fill_image_info(info, widths[i], heights[i], sample_formats[i], row_strides[i], buffer_kind, buffers[i], cuda_stream,
            unchanged_channels);
        if (rois && rois[4 * i + 2] > 0) { // region of interest: x, y, width, height; widths / heights are the region's
            info.region.struct_type = NVIMGCODEC_STRUCTURE_TYPE_REGION;
            info.region.struct_size = sizeof(nvimgcodecRegion_t);
            info.region.ndim = 2;
            info.region.start[0] = rois[4 * i + 1];
            info.region.start[1] = rois[4 * i];
            info.region.end[0] = rois[4 * i + 1] + rois[4 * i + 3];
            info.region.end[1] = rois[4 * i] + rois[4 * i + 2];
        }
        if (orient) { // what the Python layer does: the code stream's EXIF orientation goes into the target's info (python/decoder.cpp:190-205)
            info.orientation.struct_type = NVIMGCODEC_STRUCTURE_TYPE_ORIENTATION;
            info.orientation.struct_size = sizeof(nvimgcodecOrientation_t);
            info.orientation.rotated = orient[3 * i];
            info.orientation.flip_x = orient[3 * i + 1];
            info.orientation.flip_y = orient[3 * i + 2];
        }
        if (nvimgcodecImageCreate(inst, &img[i], &info) != NVIMGCODEC_STATUS_SUCCESS)
            rc = -2;
    }
    nvimgcodecDecodeParams_t params{NVIMGCODEC_STRUCTURE_TYPE_DECODE_PARAMS, sizeof(nvimgcodecDecodeParams_t), 0};
    params.apply_exif_orientation = apply_exif;
    params.enable_roi = rois ? 1 : 0;
    for (int r = 0; r < reps && rc == 0; ++r) {
        double t0 = now_s();
        nvimgcodecFuture_t fut = nullptr;
        if (nvimgcodecDecoderDecode(d->dec, cs.data(), img.data(), n, &params, &fut) != NVIMGCODEC_STATUS_SUCCESS) {
            rc = -3;
            break;
        }
        nvimgcodecFutureWaitForAll(fut);
        if (buffer_kind == NVIMGCODEC_IMAGE_BUFFER_KIND_STRIDED_DEVICE)
            cudaStreamSynchronize((cudaStream_t)cuda_stream);
        double t1 = now_s();
        if (seconds)
            seconds[r] = t1 - t0;
        size_t sz = 0;
        std::vector<nvimgcodecProcessingStatus_t> st(n, 0);
        nvimgcodecFutureGetProcessingStatus(fut, st.data(), &sz);
        if (status_out)
            for (int i = 0; i < n; ++i)
                status_out[i] = st[i];
        nvimgcodecFutureDestroy(fut);
    }
    for (int i = 0; i < n; ++i) {
        if (img[i])
            nvimgcodecImageDestroy(img[i]);
        if (cs[i])
            nvimgcodecCodeStreamDestroy(cs[i]);
    }
    return rc;
}

REFH_API int refh_decode(void* hd, int n, const unsigned char** data, const size_t* sizes, void** buffers, const int* widths,
    const int* heights, const int* sample_formats, const int64_t* row_strides, int buffer_kind, void* cuda_stream, int reps,
    unsigned int* status_out, double* seconds)
{
    return decode_impl(hd, n, data, sizes, buffers, widths, heights, sample_formats, row_strides, buffer_kind, cuda_stream, reps, status_out,
        seconds, nullptr);
}

// The same with a region of interest per image (rois: x, y, width, height; width 0 = whole image); widths / heights
// describe the target buffers, i.e. the regions (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:355-369).
REFH_API int refh_decode_roi(void* hd, int n, const unsigned char** data, const size_t* sizes, void** buffers, const int* widths,
    const int* heights, const int* sample_formats, const int64_t* row_strides, int buffer_kind, void* cuda_stream, int reps,
    unsigned int* status_out, double* seconds, const int* rois)
{
    return decode_impl(hd, n, data, sizes, buffers, widths, heights, sample_formats, row_strides, buffer_kind, cuda_stream, reps, status_out,
        seconds, rois);
}

// The same with an EXIF orientation per image (orient: rotated degrees counter-clockwise, flip_x, flip_y) and
// nvimgcodecDecodeParams_t::apply_exif_orientation; widths / heights describe the target buffers (swapped for 90 / 270).
REFH_API int refh_decode_ex(void* hd, int n, const unsigned char** data, const size_t* sizes, void** buffers, const int* widths,
    const int* heights, const int* sample_formats, const int64_t* row_strides, int buffer_kind, void* cuda_stream, int reps,
    unsigned int* status_out, double* seconds, const int* rois, const int* orient, int apply_exif)
{
    return decode_impl(hd, n, data, sizes, buffers, widths, heights, sample_formats, row_strides, buffer_kind, cuda_stream, reps, status_out,
        seconds, rois, orient, apply_exif);
}

// Steady-state decode loop for the benchmark and the pipelining tests: `rounds` decode calls of the same n code streams,
// alternating between two sets of target buffers, with `depth` calls in flight (1: Decode -> FutureWaitForAll -> next;
// 2: batch r + 1 is handed in before the future of batch r is waited for -- a loader that keeps one batch ahead).
// With d2h_bytes != 0 every finished batch is followed by one cudaMemcpyAsync(device -> host) of that many bytes on
// cuda_stream (what a user of the Python layer does with Image.cpu(); the targets of a set then sit back to back in one
// allocation starting at d2h_src_*).  seconds[0] = wall time from the first Decode to everything done (stream synchronised).
// hd_b: the decoder that takes the odd rounds.  The reference's own plugins keep per-decoder batch state
// (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp decodeBatch refills decode_state_batch_->samples_), so two batches in
// flight on ONE nvimgcodecDecoder_t are outside what the framework promises; two decoders, one batch each, are not.
REFH_API int refh_decode_pipelined(void* hd, void* hd_b, int n, const unsigned char** data, const size_t* sizes, void** buffers_a, void** buffers_b,
    const int* widths, const int* heights, const int* sample_formats, const int64_t* row_strides, int buffer_kind, void* cuda_stream,
    int rounds, int depth, void* d2h_dst_a, void* d2h_dst_b, const void* d2h_src_a, const void* d2h_src_b, size_t d2h_bytes,
    unsigned int* status_out, double* seconds)
{
    auto* d = (Decoder*)hd;
    auto* d2 = hd_b ? (Decoder*)hd_b : d;
    auto inst = d->s->instance;
    std::vector<nvimgcodecCodeStream_t> cs(n, nullptr);
    std::vector<nvimgcodecImage_t> img[2];
    img[0].assign(n, nullptr);
    img[1].assign(n, nullptr);
    int rc = 0;
    for (int i = 0; i < n && rc == 0; ++i) {
        if (nvimgcodecCodeStreamCreateFromHostMem(inst, &cs[i], data[i], sizes[i]) != NVIMGCODEC_STATUS_SUCCESS) {
            rc = -1;
            break;
        }
        for (int k = 0; k < 2; ++k) {
            nvimgcodecImageInfo_t info;
            fill_image_info(info, widths[i], heights[i], sample_formats[i], row_strides[i], buffer_kind, (k ? buffers_b : buffers_a)[i], cuda_stream, 3);
            if (nvimgcodecImageCreate(inst, &img[k][i], &info) != NVIMGCODEC_STATUS_SUCCESS)
                rc = -2;
        }
    }
    nvimgcodecDecodeParams_t params{NVIMGCODEC_STRUCTURE_TYPE_DECODE_PARAMS, sizeof(nvimgcodecDecodeParams_t), 0};
    if (status_out)
        for (int i = 0; i < n; ++i)
            status_out[i] = NVIMGCODEC_PROCESSING_STATUS_SUCCESS;
    // The copies run on a stream of their own: on the images' stream the copy of batch r would queue up behind the wait for
    // batch r + 1 (a GPU plugin makes the images' stream wait for the decode when the call is made).
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t decoded[2] = {nullptr, nullptr};
    if (d2h_bytes) {
        cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking);
        cudaEventCreateWithFlags(&decoded[0], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&decoded[1], cudaEventDisableTiming);
    }
    auto finish = [&](nvimgcodecFuture_t fut, int k) {
        nvimgcodecFutureWaitForAll(fut);
        size_t sz = 0;
        std::vector<nvimgcodecProcessingStatus_t> st(n, 0);
        nvimgcodecFutureGetProcessingStatus(fut, st.data(), &sz);
        for (int i = 0; i < n; ++i)
            if (status_out && st[i] != NVIMGCODEC_PROCESSING_STATUS_SUCCESS)
                status_out[i] = st[i];
        nvimgcodecFutureDestroy(fut);
        if (d2h_bytes) { // behind the decode of THIS batch only: the event was recorded on the images' stream right after Decode returned
            cudaStreamWaitEvent(copy_stream, decoded[k], 0);
            cudaMemcpyAsync(k ? d2h_dst_b : d2h_dst_a, k ? d2h_src_b : d2h_src_a, d2h_bytes, cudaMemcpyDeviceToHost, copy_stream);
        }
    };
    if (rc == 0) {
        double t0 = now_s();
        nvimgcodecFuture_t fut[2] = {nullptr, nullptr};
        for (int r = 0; r < rounds && rc == 0; ++r) {
            const int k = r & 1;
            if (nvimgcodecDecoderDecode((k ? d2 : d)->dec, cs.data(), img[k].data(), n, &params, &fut[k]) != NVIMGCODEC_STATUS_SUCCESS) {
                rc = -3;
                break;
            }
            if (d2h_bytes)
                cudaEventRecord(decoded[k], (cudaStream_t)cuda_stream);
            if (depth <= 1) {
                finish(fut[k], k);
                fut[k] = nullptr;
            } else if (r > 0) {
                finish(fut[k ^ 1], k ^ 1);
                fut[k ^ 1] = nullptr;
            }
        }
        for (int k = 0; k < 2; ++k) {
            const int kk = (rounds & 1) ? (k ^ 1) : k; // the older one first
            if (fut[kk])
                finish(fut[kk], kk);
            fut[kk] = nullptr;
        }
        cudaStreamSynchronize((cudaStream_t)cuda_stream);
        if (copy_stream)
            cudaStreamSynchronize(copy_stream);
        if (seconds)
            seconds[0] = now_s() - t0;
    }
    if (copy_stream) {
        cudaStreamDestroy(copy_stream);
        cudaEventDestroy(decoded[0]);
        cudaEventDestroy(decoded[1]);
    }
    for (int i = 0; i < n; ++i) {
        for (int k = 0; k < 2; ++k)
            if (img[k][i])
                nvimgcodecImageDestroy(img[k][i]);
        if (cs[i])
            nvimgcodecCodeStreamDestroy(cs[i]);
    }
    return rc;
}

// nvimgcodecDecoderCanDecode with every field of the target's image info spelled out: the rejection matrix of a plugin's
// canDecode (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:118-173).  orient: 3 ints per image or NULL.
REFH_API int refh_can_decode_ex(void* hd, int n, const unsigned char** data, const size_t* sizes, const int* widths, const int* heights,
    const int* sample_formats, const int* num_planes, const int* num_channels, const int* sample_types, const int* orient, int apply_exif,
    int buffer_kind, int force_format, unsigned int* status_out)
{
    auto* d = (Decoder*)hd;
    auto inst = d->s->instance;
    std::vector<nvimgcodecCodeStream_t> cs(n, nullptr);
    std::vector<nvimgcodecImage_t> img(n, nullptr);
    int rc = 0;
    static unsigned char dummy[16];
    for (int i = 0; i < n && rc == 0; ++i) {
        if (nvimgcodecCodeStreamCreateFromHostMem(inst, &cs[i], data[i], sizes[i]) != NVIMGCODEC_STATUS_SUCCESS) {
            rc = -1;
            break;
        }
        nvimgcodecImageInfo_t info;
        fill_image_info(info, widths[i], heights[i], sample_formats[i], (int64_t)widths[i] * num_channels[i], buffer_kind, dummy, nullptr, 3);
        info.num_planes = num_planes[i];
        for (int p = 0; p < num_planes[i] && p < NVIMGCODEC_MAX_NUM_PLANES; ++p) {
            info.plane_info[p].struct_type = NVIMGCODEC_STRUCTURE_TYPE_IMAGE_PLANE_INFO;
            info.plane_info[p].struct_size = sizeof(nvimgcodecImagePlaneInfo_t);
            info.plane_info[p].width = widths[i];
            info.plane_info[p].height = heights[i];
            info.plane_info[p].num_channels = num_channels[i];
            info.plane_info[p].row_stride = (size_t)widths[i] * num_channels[i];
            info.plane_info[p].sample_type = (nvimgcodecSampleDataType_t)sample_types[i];
        }
        if (orient) {
            info.orientation.struct_type = NVIMGCODEC_STRUCTURE_TYPE_ORIENTATION;
            info.orientation.struct_size = sizeof(nvimgcodecOrientation_t);
            info.orientation.rotated = orient[3 * i];
            info.orientation.flip_x = orient[3 * i + 1];
            info.orientation.flip_y = orient[3 * i + 2];
        }
        if (nvimgcodecImageCreate(inst, &img[i], &info) != NVIMGCODEC_STATUS_SUCCESS)
            rc = -2;
    }
    nvimgcodecDecodeParams_t params{NVIMGCODEC_STRUCTURE_TYPE_DECODE_PARAMS, sizeof(nvimgcodecDecodeParams_t), 0};
    params.apply_exif_orientation = apply_exif;
    if (rc == 0) {
        std::vector<nvimgcodecProcessingStatus_t> st(n, 0);
        if (nvimgcodecDecoderCanDecode(d->dec, cs.data(), img.data(), n, &params, st.data(), force_format) != NVIMGCODEC_STATUS_SUCCESS)
            rc = -3;
        for (int i = 0; i < n; ++i)
            status_out[i] = st[i];
    }
    for (int i = 0; i < n; ++i) {
        if (img[i])
            nvimgcodecImageDestroy(img[i]);
        if (cs[i])
            nvimgcodecCodeStreamDestroy(cs[i]);
    }
    return rc;
}

// nvimgcodecDecoderCanDecode for a batch (src/image_generic_decoder.cpp:84-131)
REFH_API int refh_can_decode(void* hd, int n, const unsigned char** data, const size_t* sizes, const int* widths, const int* heights,
    const int* sample_formats, int buffer_kind, int force_format, unsigned int* status_out)
{
    auto* d = (Decoder*)hd;
    auto inst = d->s->instance;
    std::vector<nvimgcodecCodeStream_t> cs(n, nullptr);
    std::vector<nvimgcodecImage_t> img(n, nullptr);
    int rc = 0;
    static unsigned char dummy[16];
    for (int i = 0; i < n && rc == 0; ++i) {
        if (nvimgcodecCodeStreamCreateFromHostMem(inst, &cs[i], data[i], sizes[i]) != NVIMGCODEC_STATUS_SUCCESS) {
            rc = -1;
            break;
        }
        nvimgcodecImageInfo_t info;
        int ch = channels_of(sample_formats[i], 3);
        fill_image_info(info, widths[i], heights[i], sample_formats[i], (int64_t)widths[i] * (is_planar(sample_formats[i]) ? 1 : ch),
            buffer_kind, dummy, nullptr, 3);
        if (nvimgcodecImageCreate(inst, &img[i], &info) != NVIMGCODEC_STATUS_SUCCESS)
            rc = -2;
    }
    nvimgcodecDecodeParams_t params{NVIMGCODEC_STRUCTURE_TYPE_DECODE_PARAMS, sizeof(nvimgcodecDecodeParams_t), 0};
    if (rc == 0) {
        std::vector<nvimgcodecProcessingStatus_t> st(n, 0);
        if (nvimgcodecDecoderCanDecode(d->dec, cs.data(), img.data(), n, &params, st.data(), force_format) != NVIMGCODEC_STATUS_SUCCESS)
            rc = -3;
        for (int i = 0; i < n; ++i)
            status_out[i] = st[i];
    }
    for (int i = 0; i < n; ++i) {
        if (img[i])
            nvimgcodecImageDestroy(img[i]);
        if (cs[i])
            nvimgcodecCodeStreamDestroy(cs[i]);
    }
    return rc;
}

// ---- encoder through the public API (src/nvimgcodec_capi.cpp:563-590); only useful with an encoder plugin loaded
REFH_API void* refh_encoder_create(void* h, int device_id, int backend_mask, int num_threads, const char* options)
{
    auto* e = new Encoder();
    e->s = (Session*)h;
    for (int k = 1; k <= 4; ++k)
        if (backend_mask & (1 << (k - 1))) {
            nvimgcodecBackend_t b{NVIMGCODEC_STRUCTURE_TYPE_BACKEND, sizeof(nvimgcodecBackend_t), 0};
            b.kind = (nvimgcodecBackendKind_t)k;
            b.params = {NVIMGCODEC_STRUCTURE_TYPE_BACKEND_PARAMS, sizeof(nvimgcodecBackendParams_t), 0, 1.0f};
            e->backends.push_back(b);
        }
    e->exec = {NVIMGCODEC_STRUCTURE_TYPE_EXECUTION_PARAMS, sizeof(nvimgcodecExecutionParams_t), 0};
    e->exec.device_id = device_id;
    e->exec.max_num_cpu_threads = num_threads;
    e->exec.num_backends = (int)e->backends.size();
    e->exec.backends = e->backends.empty() ? nullptr : e->backends.data();
    if (nvimgcodecEncoderCreate(e->s->instance, &e->enc, &e->exec, options) != NVIMGCODEC_STATUS_SUCCESS) {
        delete e;
        return nullptr;
    }
    return e;
}

REFH_API void refh_encoder_destroy(void* he)
{
    auto* e = (Encoder*)he;
    if (!e)
        return;
    nvimgcodecEncoderDestroy(e->enc);
    delete e;
}

namespace {
struct OutBuf
{
    unsigned char* ptr;
    size_t cap;
    size_t used;
};
unsigned char* resize_cb(void* ctx, size_t req)
{
    auto* o = (OutBuf*)ctx;
    if (req > o->cap)
        return nullptr;
    o->used = req;
    return o->ptr;
}
} // namespace

// Encode a batch of interleaved-RGB (or P_Y gray) images through nvimgcodecEncoderEncode.
// chroma_subsampling: nvimgcodecChromaSubsampling_t for the *code stream* (nvjpeg/cuda_encoder.cpp:358-361).
REFH_API int refh_encode(void* he, int n, void** buffers, const int* widths, const int* heights, const int* sample_formats,
    const int64_t* row_strides, int buffer_kind, void* cuda_stream, float quality, const int* chroma_subsampling,
    unsigned char** outs, const size_t* out_caps, size_t* out_sizes, int reps, unsigned int* status_out, double* seconds)
{
    auto* e = (Encoder*)he;
    auto inst = e->s->instance;
    std::vector<nvimgcodecCodeStream_t> cs(n, nullptr);
    std::vector<nvimgcodecImage_t> img(n, nullptr);
    std::vector<OutBuf> ob(n);
    int rc = 0;
    for (int i = 0; i < n && rc == 0; ++i) {
        nvimgcodecImageInfo_t info;
        fill_image_info(info, widths[i], heights[i], sample_formats[i], row_strides[i], buffer_kind, buffers[i], cuda_stream, 3);
        if (nvimgcodecImageCreate(inst, &img[i], &info) != NVIMGCODEC_STATUS_SUCCESS) {
            rc = -2;
            break;
        }
        nvimgcodecImageInfo_t out_info = info;
        strcpy(out_info.codec_name, "jpeg");
        out_info.chroma_subsampling = (nvimgcodecChromaSubsampling_t)chroma_subsampling[i];
        out_info.buffer = nullptr;
        ob[i] = {outs[i], out_caps[i], 0};
        if (nvimgcodecCodeStreamCreateToHostMem(inst, &cs[i], &ob[i], resize_cb, &out_info) != NVIMGCODEC_STATUS_SUCCESS)
            rc = -1;
    }
    nvimgcodecJpegEncodeParams_t jp{NVIMGCODEC_STRUCTURE_TYPE_JPEG_ENCODE_PARAMS, sizeof(nvimgcodecJpegEncodeParams_t), 0};
    jp.optimized_huffman = 0;
    nvimgcodecEncodeParams_t params{NVIMGCODEC_STRUCTURE_TYPE_ENCODE_PARAMS, sizeof(nvimgcodecEncodeParams_t), &jp};
    params.quality = quality;
    params.target_psnr = 0;
    for (int r = 0; r < reps && rc == 0; ++r) {
        double t0 = now_s();
        nvimgcodecFuture_t fut = nullptr;
        if (nvimgcodecEncoderEncode(e->enc, img.data(), cs.data(), n, &params, &fut) != NVIMGCODEC_STATUS_SUCCESS) {
            rc = -3;
            break;
        }
        nvimgcodecFutureWaitForAll(fut);
        double t1 = now_s();
        if (seconds)
            seconds[r] = t1 - t0;
        size_t sz = 0;
        std::vector<nvimgcodecProcessingStatus_t> st(n, 0);
        nvimgcodecFutureGetProcessingStatus(fut, st.data(), &sz);
        for (int i = 0; i < n; ++i) {
            if (status_out)
                status_out[i] = st[i];
            out_sizes[i] = ob[i].used;
        }
        nvimgcodecFutureDestroy(fut);
    }
    for (int i = 0; i < n; ++i) {
        if (img[i])
            nvimgcodecImageDestroy(img[i]);
        if (cs[i])
            nvimgcodecCodeStreamDestroy(cs[i]);
    }
    return rc;
}

// ------------------------------------------------------------------------------------------------
// 2. libjpeg-turbo direct: corpus generator + coefficient dump + plain decode
// ------------------------------------------------------------------------------------------------
namespace {
struct ErrMgr
{
    jpeg_error_mgr pub;
    jmp_buf jb;
    char msg[JMSG_LENGTH_MAX];
};
void err_exit(j_common_ptr cinfo)
{
    auto* e = (ErrMgr*)cinfo->err;
    (*cinfo->err->format_message)(cinfo, e->msg);
    longjmp(e->jb, 1);
}
void err_silent(j_common_ptr, int) {}
void out_silent(j_common_ptr) {}
} // namespace

// Compress interleaved 8-bit pixels (channels = 3 RGB or 1 gray) with libjpeg-turbo.
//   h_samp/v_samp : luma sampling factors (chroma 1x1) ; ignored for gray ; h_samp == 0 with RGB input => cjpeg -gray
//   restart_interval (MCUs) / restart_in_rows : as in jpeglib.h ; 0 = none
//   optimize : optimize_coding ; progressive : jpeg_simple_progression ; arith: arithmetic coding
// returns number of bytes written to out (<= cap) or -1
REFH_API long refh_ljt_encode(const unsigned char* pixels, int width, int height, int channels, int quality, int h_samp, int v_samp,
    int restart_interval, int restart_in_rows, int optimize, int progressive, int arith, unsigned char* out, size_t cap)
{
    jpeg_compress_struct c;
    ErrMgr em;
    c.err = jpeg_std_error(&em.pub);
    em.pub.error_exit = err_exit;
    unsigned char* mem = nullptr;
    unsigned long memsz = 0;
    if (setjmp(em.jb)) {
        jpeg_destroy_compress(&c);
        free(mem);
        return -1;
    }
    jpeg_create_compress(&c);
    jpeg_mem_dest(&c, &mem, &memsz);
    c.image_width = width;
    c.image_height = height;
    c.input_components = channels;
    c.in_color_space = channels == 3 ? JCS_RGB : JCS_GRAYSCALE;
    jpeg_set_defaults(&c);
    jpeg_set_quality(&c, quality, TRUE);
    c.dct_method = JDCT_ISLOW;
    c.optimize_coding = optimize ? TRUE : FALSE;
    c.arith_code = arith ? TRUE : FALSE;
    if (channels == 3 && h_samp == 0) {
        jpeg_set_colorspace(&c, JCS_GRAYSCALE); // cjpeg -gray on RGB input
    } else if (channels == 3) {
        c.comp_info[0].h_samp_factor = h_samp;
        c.comp_info[0].v_samp_factor = v_samp;
        c.comp_info[1].h_samp_factor = c.comp_info[1].v_samp_factor = 1;
        c.comp_info[2].h_samp_factor = c.comp_info[2].v_samp_factor = 1;
    }
    c.restart_interval = restart_interval;
    c.restart_in_rows = restart_in_rows;
    if (progressive)
        jpeg_simple_progression(&c);
    jpeg_start_compress(&c, TRUE);
    while (c.next_scanline < c.image_height) {
        JSAMPROW row = (JSAMPROW)(pixels + (size_t)c.next_scanline * width * channels);
        jpeg_write_scanlines(&c, &row, 1);
    }
    jpeg_finish_compress(&c);
    jpeg_destroy_compress(&c);
    long n = -1;
    if (memsz <= cap) {
        memcpy(out, mem, memsz);
        n = (long)memsz;
    }
    free(mem);
    return n;
}

// Dump quantised coefficients of component `comp` (jpeg_read_coefficients) in natural (de-zigzagged) order,
// block-row-major over the *padded* component block grid (width_in_blocks rounded up to the MCU):
// out[(by*wib + bx)*64 + k].  dims[0..1] = blocks per row / block rows written.  Returns 0 or -1.
REFH_API int refh_ljt_read_coefficients(const unsigned char* data, size_t size, int comp, short* out, size_t cap_blocks, int* dims)
{
    jpeg_decompress_struct d;
    ErrMgr em;
    d.err = jpeg_std_error(&em.pub);
    em.pub.error_exit = err_exit;
    em.pub.emit_message = err_silent;
    em.pub.output_message = out_silent;
    if (setjmp(em.jb)) {
        jpeg_destroy_decompress(&d);
        return -1;
    }
    jpeg_create_decompress(&d);
    jpeg_mem_src(&d, data, size);
    jpeg_read_header(&d, TRUE);
    jvirt_barray_ptr* arrays = jpeg_read_coefficients(&d);
    if (comp >= d.num_components) {
        jpeg_destroy_decompress(&d);
        return -1;
    }
    jpeg_component_info* ci = &d.comp_info[comp];
    // padded grid as allocated by jdcoefct.c (full iMCU rows/cols)
    JDIMENSION wib = (JDIMENSION)((ci->width_in_blocks + ci->h_samp_factor - 1) / ci->h_samp_factor * ci->h_samp_factor);
    JDIMENSION hib = (JDIMENSION)((ci->height_in_blocks + ci->v_samp_factor - 1) / ci->v_samp_factor * ci->v_samp_factor);
    dims[0] = (int)wib;
    dims[1] = (int)hib;
    if ((size_t)wib * hib > cap_blocks) {
        jpeg_destroy_decompress(&d);
        return -2;
    }
    for (JDIMENSION by = 0; by < hib; ++by) {
        JBLOCKARRAY rows = (*d.mem->access_virt_barray)((j_common_ptr)&d, arrays[comp], by, 1, FALSE);
        memcpy(out + (size_t)by * wib * 64, rows[0], (size_t)wib * 64 * sizeof(short));
    }
    jpeg_finish_decompress(&d);
    jpeg_destroy_decompress(&d);
    return 0;
}

// Plain libjpeg decode (the same call sequence as extensions/libjpeg_turbo/jpeg_mem.cpp:143-199,290-376,445 without
// the nvimgcodec plumbing) -- used to cross-check the plugin path and to dump raw (pre-upsampling) component planes.
//   out_channels 3 => JCS_RGB (or JCS_EXT_BGR if bgr), 1 => JCS_GRAYSCALE ; fancy = do_fancy_upsampling ; dct 0 = ISLOW
REFH_API int refh_ljt_decode(const unsigned char* data, size_t size, int out_channels, int bgr, int fancy, unsigned char* out,
    int64_t row_stride)
{
    jpeg_decompress_struct d;
    ErrMgr em;
    d.err = jpeg_std_error(&em.pub);
    em.pub.error_exit = err_exit;
    em.pub.emit_message = err_silent;
    em.pub.output_message = out_silent;
    if (setjmp(em.jb)) {
        jpeg_destroy_decompress(&d);
        return -1;
    }
    jpeg_create_decompress(&d);
    jpeg_mem_src(&d, data, size);
    jpeg_read_header(&d, TRUE);
    d.out_color_space = out_channels == 1 ? JCS_GRAYSCALE : (bgr ? JCS_EXT_BGR : JCS_RGB);
    d.do_fancy_upsampling = fancy ? TRUE : FALSE;
    d.dct_method = JDCT_ISLOW;
    jpeg_start_decompress(&d);
    while (d.output_scanline < d.output_height) {
        JSAMPROW row = out + (size_t)d.output_scanline * row_stride;
        jpeg_read_scanlines(&d, &row, 1);
    }
    jpeg_finish_decompress(&d);
    jpeg_destroy_decompress(&d);
    return 0;
}

// Raw component planes after dequant+IDCT (raw_data_out; jdapistd.c jpeg_read_raw_data) for stage-wise bisecting.
// plane c is written at out + offsets[c], row pitch pitches[c] (must hold padded dims: width_in_blocks*8 x
// iMCU-rounded rows).
REFH_API int refh_ljt_decode_raw(const unsigned char* data, size_t size, unsigned char* out, const size_t* offsets, const int* pitches)
{
    jpeg_decompress_struct d;
    ErrMgr em;
    d.err = jpeg_std_error(&em.pub);
    em.pub.error_exit = err_exit;
    em.pub.emit_message = err_silent;
    em.pub.output_message = out_silent;
    if (setjmp(em.jb)) {
        jpeg_destroy_decompress(&d);
        return -1;
    }
    jpeg_create_decompress(&d);
    jpeg_mem_src(&d, data, size);
    jpeg_read_header(&d, TRUE);
    d.raw_data_out = TRUE;
    d.dct_method = JDCT_ISLOW;
    jpeg_start_decompress(&d);
    int ncomp = d.num_components;
    int mcu_rows = d.max_v_samp_factor * DCTSIZE;
    std::vector<std::vector<JSAMPROW>> rows(ncomp);
    std::vector<JSAMPARRAY> planes(ncomp);
    while (d.output_scanline < d.output_height) {
        int imcu = d.output_scanline / mcu_rows;
        for (int c = 0; c < ncomp; ++c) {
            int vr = d.comp_info[c].v_samp_factor * DCTSIZE;
            rows[c].resize(vr);
            for (int r = 0; r < vr; ++r)
                rows[c][r] = out + offsets[c] + (size_t)(imcu * vr + r) * pitches[c];
            planes[c] = rows[c].data();
        }
        jpeg_read_raw_data(&d, planes.data(), mcu_rows);
    }
    jpeg_finish_decompress(&d);
    jpeg_destroy_decompress(&d);
    return 0;
}
