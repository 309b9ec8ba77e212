/* b200jpeg.h -- C ABI of the B200-native baseline-JPEG codec (libb200jpeg.so).
 *
 * This is the drop-in boundary below nvImageCodec's plugin interface.  Every entry point takes plain pointers and sizes
 * (no C++/torch types).  The nvImageCodec extension module (libb200jpeg_ext.so, exporting nvimgcodecExtensionModuleEntry,
 * reference include/nvimgcodec.h:1356-1364) is a thin adapter that forwards the plugin callbacks
 *     nvimgcodecDecoderDesc_t::create/destroy/canDecode/decode   (reference include/nvimgcodec.h:1150-1209)
 *     nvimgcodecEncoderDesc_t::create/destroy/canEncode/encode   (reference include/nvimgcodec.h:1087-1145)
 * to the functions below; the reference implementation they replace is extensions/libjpeg_turbo/
 * libjpeg_turbo_decoder.cpp:118-173 (canDecode), :324-490 (decode/decodeBatch) and, for encode, the libjpeg-turbo
 * jpeg_compress path (the reference ships no CPU JPEG encoder plugin; extensions/nvjpeg/cuda_encoder.cpp:284-396 is the
 * interface precedent).
 *
 * All compute is CUDA on sm_100a.  There is no CPU fallback: every call that needs the device fails with
 * B200JPEG_STATUS_CUDA_ERROR / ARCH_MISMATCH when it is missing.
 */
#ifndef B200JPEG_H
#define B200JPEG_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200JPEG_API __attribute__((visibility("default")))
#else
#define B200JPEG_API
#endif

typedef struct b200jpegDecoder* b200jpegDecoder_t;
typedef struct b200jpegEncoder* b200jpegEncoder_t;

typedef enum
{
    B200JPEG_STATUS_SUCCESS = 0,
    B200JPEG_STATUS_INVALID_PARAMETER = 1,
    B200JPEG_STATUS_BAD_JPEG = 2,         /* marker syntax broken / truncated header */
    B200JPEG_STATUS_UNSUPPORTED = 3,      /* valid JPEG outside this codec's scope (progressive, arithmetic, 12-bit, CMYK, 4:1:0 ...) */
    B200JPEG_STATUS_CORRUPT_STREAM = 4,   /* entropy-coded data inconsistent with the header (detected on the GPU) */
    B200JPEG_STATUS_CUDA_ERROR = 5,
    B200JPEG_STATUS_ARCH_MISMATCH = 6,    /* device is not sm_100 */
    B200JPEG_STATUS_ALLOC_FAILED = 7,
    B200JPEG_STATUS_BUFFER_TOO_SMALL = 8, /* encoder output capacity */
    B200JPEG_STATUS_INTERNAL_ERROR = 9
} b200jpegStatus_t;

/* Output sample formats; the numeric values equal nvimgcodecSampleFormat_t (reference include/nvimgcodec.h:384-392). */
typedef enum
{
    B200JPEG_FMT_P_UNCHANGED = 1, /* planar RGB, also for gray streams (what libjpeg_turbo_ext produces) */
    B200JPEG_FMT_I_UNCHANGED = 2, /* interleaved RGB */
    B200JPEG_FMT_P_RGB = 3,
    B200JPEG_FMT_I_RGB = 4,
    B200JPEG_FMT_P_BGR = 5,
    B200JPEG_FMT_I_BGR = 6,
    B200JPEG_FMT_P_Y = 7
} b200jpegFormat_t;

/* Chroma subsampling; numeric values equal nvimgcodecChromaSubsampling_t (reference include/nvimgcodec.h:360-372). */
typedef enum
{
    B200JPEG_CSS_444 = 0,
    B200JPEG_CSS_422 = 2,
    B200JPEG_CSS_420 = 3,
    B200JPEG_CSS_440 = 4,
    B200JPEG_CSS_GRAY = 7
} b200jpegSubsampling_t;

/* What the header parse found (the plugin's canDecode decides on this; mirrors what the reference learns from
 * jpeg_read_header, external/libjpeg-turbo/jdmarker.c get_sof:242 / get_sos:308 / get_dri:569). */
typedef struct
{
    int32_t width, height;
    int32_t num_components;       /* 1 or 3 */
    int32_t h_samp[3], v_samp[3];
    int32_t subsampling;          /* b200jpegSubsampling_t, -1 if not one of them */
    int32_t sof_marker;           /* 0xC0 .. 0xCF */
    int32_t restart_interval;     /* MCUs, 0 = none */
    int32_t precision;
    int32_t supported;            /* 1 if b200jpeg_decode_batch can decode it */
    uint64_t scan_offset, scan_bytes; /* entropy-coded segment inside the file */
} b200jpegStreamInfo_t;

/* One output image of a decode batch: device (or host) pixel buffer description, same meaning as the
 * nvimgcodecImageInfo_t fields buffer / plane_info[].row_stride / sample_format (reference include/nvimgcodec.h:474-497).
 * Planar formats: plane p starts at buffer + p * row_stride * height (reference extensions/nvjpeg/cuda_decoder.cpp:532-538). */
typedef struct
{
    void* buffer;
    int64_t row_stride;  /* bytes */
    int32_t format;      /* b200jpegFormat_t */
    int32_t on_host;     /* 0: buffer is device memory (plugin case); 1: host memory, the library stages + copies back */
    /* Region of interest = nvimgcodecImageInfo_t::region (reference include/nvimgcodec.h:446-457, consumed by
     * extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:355-369 and jpeg_mem.cpp:206-240): the buffer receives pixels
     * [roi_x, roi_x + roi_width) x [roi_y, roi_y + roi_height) of the image, bit-identical to the same pixels of a full
     * decode (what the reference produces).  roi_width == roi_height == 0: the whole image.  Planar formats: plane p
     * starts at buffer + p * row_stride * roi_height. */
    int32_t roi_x, roi_y, roi_width, roi_height;
    /* nvimgcodecImageInfo_t::cuda_stream of this image (reference include/nvimgcodec.h:474-497): with has_cuda_stream != 0 the
     * decoder waits for what is in flight on that stream before it writes the buffer, and makes the stream wait for the
     * decode (reference idiom: extensions/nvjpeg/hw_decoder.cpp:649-655, cuda_decoder.cpp:551-556).  Otherwise the
     * user_stream argument of the call plays that role. */
    void* cuda_stream;
    int32_t has_cuda_stream;
    /* EXIF orientation applied to the output = nvimgcodecImageInfo_t::orientation with nvimgcodecDecodeParams_t::
     * apply_exif_orientation (reference include/nvimgcodec.h; the CPU plugin declines it, libjpeg_turbo_decoder.cpp:164-165;
     * the GPU precedent is extensions/nvjpeg/cuda_decoder.cpp:443-460): rotated / 90 (counter-clockwise quarter turns, 0..3)
     * | flip_x << 2 | flip_y << 3.  The buffer has the oriented size (width and height swapped for odd quarter turns).
     * Not combined with a region of interest. */
    int32_t orientation;
} b200jpegOutput_t;

/* ---- library ---- */
B200JPEG_API const char* b200jpeg_version(void);
B200JPEG_API const char* b200jpeg_status_string(int status);
/* Diagnostics (CUDA failures, configuration problems) go to stderr unless a callback is installed; severity 0 info,
 * 1 warning, 2 error.  The nvImageCodec extension installs one that forwards to nvimgcodecFrameworkDesc_t::log. */
typedef void (*b200jpegLogFn_t)(void* ctx, int severity, const char* message);
B200JPEG_API void b200jpeg_set_log_callback(b200jpegLogFn_t fn, void* ctx);

/* ---- stream probe (host only, no CUDA) ---- */
B200JPEG_API int b200jpeg_parse_header(const uint8_t* data, size_t size, b200jpegStreamInfo_t* info);

/* ---- decoder ----
 * options: the reference's option string, "name=value" tokens separated by spaces, optional "b200jpeg_decoder:" or ":"
 * prefix (reference extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:250-276): fancy_upsampling=0|1 (default 1),
 * fast_idct=0|1 (the reference's switch to JDCT_IFAST, libjpeg_turbo_decoder.cpp:272-273: create() answers
 * B200JPEG_STATUS_UNSUPPORTED for 1, so that the framework keeps its CPU decoder for such a session),
 * subseq_bytes=N (bytes of unstuffed stream per Huffman decoding thread, default 128), profile=0|1; tuning knobs, none of
 * which changes a result: groups=N / groups_host=N (image groups on separate streams), copy_threads=N, tail_launches=N
 * (work-list rounds run as ordinary launches before the cooperative kernel, default 1), tail_spread=0|1 (work-list
 * items dealt lane-major, default 1), stream_priority=0|1, stage_stream=0|1, dump_coef=0|1 (debug tap), fused=0|1 (the
 * entropy-decode stage as one kernel, k_huff_fused, instead of six launches: same results, slower on B200, default 0).
 * B200JPEG_TRACE=1 in the environment prints the host timeline, the work-list rounds and the error flags of failed images. */
B200JPEG_API int b200jpeg_decoder_create(b200jpegDecoder_t* decoder, int device_id, const char* options);
B200JPEG_API int b200jpeg_decoder_destroy(b200jpegDecoder_t decoder);

/* Memory callbacks: every device / page-locked host allocation of the decoder (encoder) made after this call goes through
 * them; call it right after create.  Same signatures and meaning as nvimgcodecDeviceAllocator_t / nvimgcodecPinnedAllocator_t
 * of nvimgcodecExecutionParams_t (reference include/nvimgcodec.h:232-302): stream-ordered (ctx, &ptr, size, stream) /
 * (ctx, ptr, size, stream), 0 = success, sizes padded to a multiple of *_padding.  A NULL pair keeps cudaMalloc /
 * cudaHostAlloc for that kind of memory. */
typedef int (*b200jpegMalloc_t)(void* ctx, void** ptr, size_t size, void* cuda_stream);
typedef int (*b200jpegFree_t)(void* ctx, void* ptr, size_t size, void* cuda_stream);
typedef struct
{
    b200jpegMalloc_t device_malloc;
    b200jpegFree_t device_free;
    void* device_ctx;
    size_t device_padding;
    b200jpegMalloc_t pinned_malloc;
    b200jpegFree_t pinned_free;
    void* pinned_ctx;
    size_t pinned_padding;
} b200jpegAllocators_t;
B200JPEG_API int b200jpeg_decoder_set_allocators(b200jpegDecoder_t decoder, const b200jpegAllocators_t* allocators);

/* Decode a batch.  data[i]/sizes[i]: whole JPEG files in host memory.  outputs[i]: where the pixels go.
 * user_stream: CUDA stream (cudaStream_t as void*) the outputs are ordered on -- on return the work has been enqueued and
 * synchronised enough that statuses[i] is final (B200JPEG_STATUS_*) and device outputs are complete.
 * Replaces DecoderImpl::decodeBatch (reference extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:444-490). */
B200JPEG_API int b200jpeg_decode_batch(b200jpegDecoder_t decoder, int batch_size, const uint8_t* const* data, const size_t* sizes,
    const b200jpegOutput_t* outputs, void* user_stream, int32_t* statuses);

/* The same call in two halves, so that the host work of batch k + 1 overlaps the device work of batch k (the plugin's
 * decode() returns after submit and reports imageReady from a completion thread, the way the reference's plugins hand
 * their work to the executor, extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:444-490).
 *   submit: parse, plan, copy the entropy-coded bytes to pinned staging, enqueue every copy and kernel, make the outputs'
 *           streams wait for the result.  On return data[]/sizes[] are not needed any more; statuses[] holds what the
 *           host already knows (bad / unsupported streams).  Up to two batches may be in flight (each has its own scratch).
 *   wait:   blocks until the oldest submitted batch is done on the device; statuses[] final (same batch_size entries). */
B200JPEG_API int b200jpeg_decode_submit(b200jpegDecoder_t decoder, int batch_size, const uint8_t* const* data, const size_t* sizes,
    const b200jpegOutput_t* outputs, void* user_stream, int32_t* statuses);
B200JPEG_API int b200jpeg_decode_wait(b200jpegDecoder_t decoder, int32_t* statuses);

/* ---- one batch over several GPUs of a box (reference hook: nvimgcodecExecutionParams_t::device_id, one decoder per
 * device, include/nvimgcodec.h:845).  Images are independent, so the batch is dealt to the devices -- greedy
 * longest-processing-time on an estimated cost of 8 x compressed bytes + pixels, the same deal on every caller without
 * communication -- and every device decodes its share with its own decoder, host thread, streams and pinned staging.
 * There is no collective: outputs stay on the GPU that decoded them (or go to host memory with on_host). ---- */
typedef struct b200jpegMultiDecoder* b200jpegMultiDecoder_t;
/* device_of_image[i] = index (0 .. num_devices-1) of the device that gets image i.  widths/heights may be NULL (cost =
 * compressed bytes only).  Pure host function. */
B200JPEG_API int b200jpeg_shard_batch(int batch_size, const size_t* sizes, const int32_t* widths, const int32_t* heights, int num_devices,
    int32_t* device_of_image);
B200JPEG_API int b200jpeg_multi_decoder_create(b200jpegMultiDecoder_t* decoder, int num_devices, const int* device_ids, const char* options);
B200JPEG_API int b200jpeg_multi_decoder_destroy(b200jpegMultiDecoder_t decoder);
/* outputs[i].buffer must live on device device_ids[device_of_image[i]] (or in host memory with on_host).  Every device's
 * share runs concurrently (one host thread each); returns when all are done; statuses[i] as for b200jpeg_decode_batch. */
B200JPEG_API int b200jpeg_multi_decode_batch(b200jpegMultiDecoder_t decoder, int batch_size, const uint8_t* const* data, const size_t* sizes,
    const b200jpegOutput_t* outputs, const int32_t* device_of_image, int32_t* statuses);

/* Two-phase variant used to time the device work with inputs already resident in HBM:
 * prepare = parse + table building + H2D of bitstreams/descriptors; run = every kernel of the decode, nothing else. */
B200JPEG_API int b200jpeg_decode_prepare(b200jpegDecoder_t decoder, int batch_size, const uint8_t* const* data, const size_t* sizes,
    const b200jpegOutput_t* outputs, int32_t* statuses);
B200JPEG_API int b200jpeg_decode_run(b200jpegDecoder_t decoder, void* user_stream, int32_t* statuses);

/* Counters of the last run: [0] kernels launched, [1] Huffman sync rounds (2 full rounds + work-list rounds), [2] entropy-coded
 * bytes, [3] decoded blocks, [4] subsequences, [5] device scratch bytes, [6] subsequences re-decoded in the work-list rounds,
 * [7] output bytes, [8] bytes of coefficient records the Huffman write pass produced, [9..15] reserved (0) */
B200JPEG_API int b200jpeg_decoder_get_stats(b200jpegDecoder_t decoder, uint64_t* stats16);

/* Per-stage device times (ms) of the last run, measured with CUDA events on the decoder's own stream; needs the option
 * "profile=1" (single group).  Stages: 0 unstuff (3 kernels + memsets), 1 Huffman sync rounds 0 and 1 (2 launches of
 * k_huff_sync), 2 work-list sync rounds (k_huff_sync_tail), 3 slot/record scan + Huffman write (k_subseq_scan, k_huff_write),
 * 4 DC prediction, 5 records -> dequant + IDCT, 6 upsample + colour.  Returns the number of stages or -1. */
B200JPEG_API int b200jpeg_decoder_get_stage_times(b200jpegDecoder_t decoder, float* ms, int n);

/* Debug/verification taps on the last prepared+run batch (device -> host copies of intermediates):
 *   what = 0: quantised coefficients of image i in MCU scan order, int16[blocks*64] (natural order inside a block);
 *             needs the decoder option "dump_coef=1" (the product path keeps coefficients as compact records only)
 *   what = 1: component plane c of image i after dequant+IDCT (MCU-padded), uint8
 *   what = 2: unstuffed entropy-coded bytes of image i
 * returns bytes written or <0 */
B200JPEG_API int64_t b200jpeg_debug_fetch(b200jpegDecoder_t decoder, int what, int image, int component, void* host_dst, size_t capacity);

/* ---- allocation-free entry point ----
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The decoder's unstuffing stage (k_unstuff_count / scan / write, the first stage of b200jpeg_decode_batch) for ONE
 * entropy-coded segment, with its scratch (per-4-KB-chunk counts) in a fixed __device__ array of the library instead of
 * allocated memory: the call makes no cudaMalloc / cudaFree / allocating library call, it only enqueues four kernels on
 * cuda_stream and returns.
 *   d_raw[raw_bytes]   device: the entropy-coded bytes of a scan (file bytes [scan_offset, scan_offset + scan_bytes) of
 *                      b200jpeg_parse_header), raw_bytes <= b200jpeg_unstuff_noalloc_max_bytes() = 64 MiB, else
 *                      B200JPEG_STATUS_INVALID_PARAMETER (and a log line naming the limit)
 *   d_clean[capacity]  device: receives the bytes without FF00 stuffing, FF fill bytes and RSTn markers, cut at the first
 *                      other marker -- what b200jpeg_debug_fetch(what = 2) shows of a decode of the same file
 *   d_result[4]        device uint32: [0] bytes written to d_clean, [1] RSTn markers removed, [2] raw bytes consumed,
 *                      [3] 1 if capacity was too small (d_clean then incomplete)
 * Limits: the scratch is 131104 bytes per GPU, held from library load to exit, shared by every call and not cleared
 * between calls beyond what the call itself resets -- calls that overlap on the GPU (two streams, two threads) race on it
 * and must be serialised by the caller. */
B200JPEG_API size_t b200jpeg_unstuff_noalloc_max_bytes(void);
B200JPEG_API int b200jpeg_unstuff_noalloc(int device_id, const uint8_t* d_raw, size_t raw_bytes, uint8_t* d_clean, size_t clean_capacity,
    uint32_t* d_result, void* cuda_stream);

/* ---- encoder (baseline, standard Huffman tables, JFIF header; mirrors libjpeg-turbo jpeg_compress with
 * jpeg_set_quality(q, TRUE), JDCT_ISLOW, optimize_coding=FALSE) ---- */
typedef struct
{
    const void* buffer;  /* interleaved RGB (format I_RGB) or gray (P_Y) uint8 */
    int64_t row_stride;
    int32_t width, height;
    int32_t format;      /* B200JPEG_FMT_I_RGB | B200JPEG_FMT_P_RGB | B200JPEG_FMT_P_Y */
    int32_t on_host;
    int32_t subsampling; /* b200jpegSubsampling_t of the produced stream */
    int32_t quality;     /* 1..100 */
    int32_t restart_interval;
} b200jpegEncodeInput_t;

B200JPEG_API int b200jpeg_encoder_create(b200jpegEncoder_t* encoder, int device_id, const char* options);
B200JPEG_API int b200jpeg_encoder_destroy(b200jpegEncoder_t encoder);
B200JPEG_API int b200jpeg_encoder_set_allocators(b200jpegEncoder_t encoder, const b200jpegAllocators_t* allocators);
/* out[i] (host memory, capacity out_capacity[i]) receives a complete JFIF file; out_sizes[i] its length. */
B200JPEG_API int b200jpeg_encode_batch(b200jpegEncoder_t encoder, int batch_size, const b200jpegEncodeInput_t* inputs, uint8_t* const* out,
    const size_t* out_capacity, size_t* out_sizes, void* user_stream, int32_t* statuses);
B200JPEG_API int b200jpeg_encoder_get_stats(b200jpegEncoder_t encoder, uint64_t* stats8);

#ifdef __cplusplus
}
#endif
#endif /* B200JPEG_H */
