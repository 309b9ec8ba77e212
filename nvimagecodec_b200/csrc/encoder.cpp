// encoder.cpp -- host side of the b200jpeg batch encoder: per-image planning (geometry, quality -> reciprocal tables,
// JFIF header bytes), pinned staging of host pixels, the kernel sequence of encode_kernels.cu on the encoder's own stream,
// and the read-back of the entropy-coded bytes behind the host-written headers.
// Interface precedent in the reference: extensions/nvjpeg/cuda_encoder.cpp:284-396 (batched GPU encoder plugin); the
// arithmetic mirrors libjpeg-turbo's jpeg_compress (see enc_math.cuh / enc_plan.cpp for the per-function citations).
#include "encoder.h"
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include "enc_plan.h"
#include "encode_kernels.h"

namespace b200jpeg {

namespace {
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
#define B2J_CUDA(call)                                                                                              \
    do {                                                                                                            \
        cudaError_t e__ = (call);                                                                                   \
        if (e__ != cudaSuccess) {                                                                                   \
            std::fprintf(stderr, "b200jpeg: CUDA error %s at %s:%d\n", cudaGetErrorString(e__), __FILE__, __LINE__); \
            return B200JPEG_STATUS_CUDA_ERROR;                                                                      \
        }                                                                                                           \
    } while (0)

// is this host pointer page-locked memory CUDA knows about (then async copies can use it directly)?
inline bool is_pinned(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

// run job(i) for i in [0, n) on up to `threads` host threads
template <class F>
inline void parallel_for(int n, int threads, const F& job)
{
    threads = std::max(1, std::min(threads, n));
    if (threads == 1) {
        for (int i = 0; i < n; ++i)
            job(i);
        return;
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t] {
            for (int i = t; i < n; i += threads)
                job(i);
        });
    for (auto& th : pool)
        th.join();
}

struct Buf
{
    void* p = nullptr;
    size_t cap = 0;
};
} // namespace

struct Encoder::Impl
{
    int device = 0;
    bool profile = false;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_in = nullptr;
    cudaEvent_t ev[8] = {};
    Buf d_pixels, d_meta, d_dyn, d_coef, d_dc, d_acbits, d_blk_off, d_seg_bits, d_seg_off, d_img_bytes, d_stream, d_chunks, d_out;
    Buf h_pixels, h_meta, h_dyn, h_sizes, h_out;
    uint64_t stats[8] = {};
    float stage_ms[4] = {};

    int ensure(Buf& b, size_t bytes, bool pinned)
    {
        if (bytes <= b.cap)
            return 0;
        if (b.p) {
            B2J_CUDA(cudaStreamSynchronize(stream));
            B2J_CUDA(pinned ? cudaFreeHost(b.p) : cudaFree(b.p));
            b.p = nullptr;
            b.cap = 0;
        }
        const size_t want = align_up(bytes + bytes / 8 + 256, 1 << 16);
        const cudaError_t e = pinned ? cudaHostAlloc(&b.p, want, cudaHostAllocDefault) : cudaMalloc(&b.p, want);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return B200JPEG_STATUS_ALLOC_FAILED;
        }
        b.cap = want;
        return 0;
    }
};

Encoder::Encoder(int device_id, const char* options) : impl_(new Impl)
{
    impl_->device = device_id;
    if (options && std::string(options).find("profile=1") != std::string::npos)
        impl_->profile = true;
}

Encoder::~Encoder()
{
    Impl& s = *impl_;
    cudaSetDevice(s.device);
    Buf* dev[] = {&s.d_pixels, &s.d_meta, &s.d_dyn, &s.d_coef, &s.d_dc, &s.d_acbits, &s.d_blk_off, &s.d_seg_bits, &s.d_seg_off, &s.d_img_bytes,
        &s.d_stream, &s.d_chunks, &s.d_out};
    for (Buf* b : dev)
        if (b->p)
            cudaFree(b->p);
    Buf* host[] = {&s.h_pixels, &s.h_meta, &s.h_dyn, &s.h_sizes, &s.h_out};
    for (Buf* b : host)
        if (b->p)
            cudaFreeHost(b->p);
    for (cudaEvent_t e : s.ev)
        if (e)
            cudaEventDestroy(e);
    if (s.ev_in)
        cudaEventDestroy(s.ev_in);
    if (s.stream)
        cudaStreamDestroy(s.stream);
    delete impl_;
}

int Encoder::init()
{
    Impl& s = *impl_;
    int count = 0;
    B2J_CUDA(cudaGetDeviceCount(&count));
    if (s.device < 0 || s.device >= count)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    B2J_CUDA(cudaSetDevice(s.device));
    cudaDeviceProp prop;
    B2J_CUDA(cudaGetDeviceProperties(&prop, s.device));
    if (prop.major != 10) {
        std::fprintf(stderr, "b200jpeg: device %d is sm_%d%d; this library only carries sm_100a code\n", s.device, prop.major, prop.minor);
        return B200JPEG_STATUS_ARCH_MISMATCH;
    }
    B2J_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    B2J_CUDA(cudaEventCreateWithFlags(&s.ev_in, cudaEventDisableTiming));
    for (cudaEvent_t& e : s.ev)
        B2J_CUDA(cudaEventCreate(&e));
    return B200JPEG_STATUS_SUCCESS;
}

int Encoder::encode(int n, const b200jpegEncodeInput_t* in, uint8_t* const* out, const size_t* out_capacity, size_t* out_sizes,
    cudaStream_t user_stream, int32_t* statuses)
{
    Impl& s = *impl_;
    static const bool trace = std::getenv("B200JPEG_TRACE") != nullptr; // developer aid: host-side timeline on stderr
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_begin = now();
    B2J_CUDA(cudaSetDevice(s.device));
    std::memset(s.stats, 0, sizeof(s.stats));
    std::vector<EncImageDesc> images;
    std::vector<int> batch_index;
    std::vector<EncQuantTable> quant;
    std::vector<std::vector<uint8_t>> headers;
    std::vector<uint32_t> map_dct(1, 0), map_emit(1, 0), seg_map(1, 0);
    std::vector<size_t> host_pix_off; // staging offset of host inputs (SIZE_MAX for device inputs)
    size_t host_pix_bytes = 0;
    uint64_t total_blocks = 0, in_bytes = 0;
    for (int i = 0; i < n; ++i) {
        out_sizes[i] = 0;
        statuses[i] = B200JPEG_STATUS_SUCCESS;
        EncImageDesc im;
        if (!in[i].buffer || !out[i] || in[i].quality < 1 || in[i].quality > 100) {
            statuses[i] = B200JPEG_STATUS_INVALID_PARAMETER;
            continue;
        }
        if (!enc_plan_image(in[i].width, in[i].height, in[i].format, in[i].subsampling, in[i].restart_interval, im)) {
            statuses[i] = B200JPEG_STATUS_UNSUPPORTED;
            continue;
        }
        const int64_t min_stride = int64_t(im.width) * (im.in_fmt == B200JPEG_FMT_I_RGB ? 3 : 1);
        if (in[i].row_stride < min_stride) {
            statuses[i] = B200JPEG_STATUS_INVALID_PARAMETER;
            continue;
        }
        im.pixels = static_cast<const uint8_t*>(in[i].buffer);
        im.stride = in[i].row_stride;
        const size_t planes = im.in_fmt == B200JPEG_FMT_P_RGB ? 3 : 1;
        const size_t bytes = size_t(im.stride) * im.height * planes;
        in_bytes += size_t(min_stride) * im.height * planes;
        if (in[i].on_host) {
            host_pix_off.push_back(host_pix_bytes);
            host_pix_bytes += align_up(bytes, 256);
        } else {
            host_pix_off.push_back(SIZE_MAX);
        }
        im.qt = int32_t(quant.size());
        quant.emplace_back();
        enc_build_quant(in[i].quality, 0, quant.back());
        quant.emplace_back();
        enc_build_quant(in[i].quality, 1, quant.back());
        headers.emplace_back();
        enc_write_headers(im, quant[im.qt], quant[im.qt + 1], headers.back());
        im.dct_cta_first = map_dct.back();
        im.emit_cta_first = map_emit.back();
        im.blk_first = total_blocks;
        im.seg_first = seg_map.back();
        map_dct.push_back(map_dct.back() + uint32_t(im.tiles_per_row) * uint32_t(im.mcus_y));
        map_emit.push_back(map_emit.back() + (im.total_blocks + kEncCta - 1) / kEncCta);
        seg_map.push_back(seg_map.back() + im.num_segments);
        total_blocks += im.total_blocks;
        images.push_back(im);
        batch_index.push_back(i);
    }
    const int m = int(images.size());
    if (m == 0)
        return B200JPEG_STATUS_SUCCESS;
    const uint32_t total_segs = seg_map.back();
    const double t_plan = now();

    // ---- scratch ----
    int rc = 0;
    const size_t off_img = 0;
    const size_t off_quant = align_up(off_img + sizeof(EncImageDesc) * m, 256);
    const size_t off_huff = align_up(off_quant + sizeof(EncQuantTable) * quant.size(), 256);
    const size_t map_bytes = align_up(sizeof(uint32_t) * (m + 1), 256);
    const size_t off_maps = align_up(off_huff + sizeof(EncHuffTables), 256);
    const size_t meta_bytes = off_maps + 3 * map_bytes;
    const size_t dyn_bytes = align_up(sizeof(EncImageDyn) * m, 256) + map_bytes;
    if ((rc = s.ensure(s.d_meta, meta_bytes, false)) || (rc = s.ensure(s.h_meta, meta_bytes, true)) || (rc = s.ensure(s.d_dyn, dyn_bytes, false)) ||
        (rc = s.ensure(s.h_dyn, dyn_bytes, true)) || (rc = s.ensure(s.d_coef, total_blocks * 128, false)) ||
        (rc = s.ensure(s.d_dc, total_blocks * 2, false)) || (rc = s.ensure(s.d_acbits, total_blocks * 2, false)) ||
        (rc = s.ensure(s.d_blk_off, total_blocks * 4, false)) || (rc = s.ensure(s.d_seg_bits, size_t(map_emit.back()) * 4 + size_t(m) * 4, false)) ||
        (rc = s.ensure(s.d_seg_off, size_t(total_segs) * 4, false)) || (rc = s.ensure(s.d_img_bytes, size_t(m) * 8, false)) ||
        (rc = s.ensure(s.h_sizes, size_t(m) * 8, true)) || (host_pix_bytes && (rc = s.ensure(s.d_pixels, host_pix_bytes, false))))
        return rc;

    // ---- host pixels -> pinned staging -> device ----
    B2J_CUDA(cudaEventRecord(s.ev_in, user_stream));
    B2J_CUDA(cudaStreamWaitEvent(s.stream, s.ev_in, 0));
    if (s.profile)
        cudaEventRecord(s.ev[0], s.stream);
    if (host_pix_bytes) {
        // page-locked inputs go up directly; pageable ones pass through the pinned staging buffer (copied by a few threads)
        std::vector<int> staged;
        for (int k = 0; k < m; ++k) {
            if (host_pix_off[k] == SIZE_MAX)
                continue;
            EncImageDesc& im = images[k];
            const size_t planes = im.in_fmt == B200JPEG_FMT_P_RGB ? 3 : 1;
            const size_t bytes = size_t(im.stride) * im.height * planes;
            uint8_t* dst = static_cast<uint8_t*>(s.d_pixels.p) + host_pix_off[k];
            if (is_pinned(im.pixels))
                B2J_CUDA(cudaMemcpyAsync(dst, im.pixels, bytes, cudaMemcpyHostToDevice, s.stream));
            else
                staged.push_back(k);
        }
        if (!staged.empty()) {
            if ((rc = s.ensure(s.h_pixels, host_pix_bytes, true)))
                return rc;
            parallel_for(int(staged.size()), 8, [&](int j) {
                const int k = staged[j];
                const EncImageDesc& im = images[k];
                const size_t planes = im.in_fmt == B200JPEG_FMT_P_RGB ? 3 : 1;
                std::memcpy(static_cast<uint8_t*>(s.h_pixels.p) + host_pix_off[k], im.pixels, size_t(im.stride) * im.height * planes);
            });
            for (int k : staged) {
                const EncImageDesc& im = images[k];
                const size_t planes = im.in_fmt == B200JPEG_FMT_P_RGB ? 3 : 1;
                B2J_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(s.d_pixels.p) + host_pix_off[k], static_cast<uint8_t*>(s.h_pixels.p) + host_pix_off[k],
                    size_t(im.stride) * im.height * planes, cudaMemcpyHostToDevice, s.stream));
            }
        }
        for (int k = 0; k < m; ++k)
            if (host_pix_off[k] != SIZE_MAX)
                images[k].pixels = static_cast<const uint8_t*>(s.d_pixels.p) + host_pix_off[k];
    }
    uint8_t* hm = static_cast<uint8_t*>(s.h_meta.p);
    std::memcpy(hm + off_img, images.data(), sizeof(EncImageDesc) * m);
    std::memcpy(hm + off_quant, quant.data(), sizeof(EncQuantTable) * quant.size());
    enc_build_huff(*reinterpret_cast<EncHuffTables*>(hm + off_huff));
    std::memcpy(hm + off_maps, map_dct.data(), sizeof(uint32_t) * (m + 1));
    std::memcpy(hm + off_maps + map_bytes, map_emit.data(), sizeof(uint32_t) * (m + 1));
    std::memcpy(hm + off_maps + 2 * map_bytes, seg_map.data(), sizeof(uint32_t) * (m + 1));
    B2J_CUDA(cudaMemcpyAsync(s.d_meta.p, s.h_meta.p, meta_bytes, cudaMemcpyHostToDevice, s.stream));

    uint8_t* dm = static_cast<uint8_t*>(s.d_meta.p);
    EncBatchDev b{};
    b.images = reinterpret_cast<const EncImageDesc*>(dm + off_img);
    b.dyn = static_cast<const EncImageDyn*>(s.d_dyn.p);
    b.num_images = m;
    b.quant = reinterpret_cast<const EncQuantTable*>(dm + off_quant);
    b.huff = reinterpret_cast<const EncHuffTables*>(dm + off_huff);
    b.coef = static_cast<int16_t*>(s.d_coef.p);
    b.dc = static_cast<int16_t*>(s.d_dc.p);
    b.acbits = static_cast<uint16_t*>(s.d_acbits.p);
    b.blk_off = static_cast<uint32_t*>(s.d_blk_off.p);
    b.tile_off = static_cast<uint32_t*>(s.d_seg_bits.p);
    b.img_bits = b.tile_off + map_emit.back();
    b.seg_byte_off = static_cast<uint32_t*>(s.d_seg_off.p);
    b.img_stream_bytes = static_cast<uint32_t*>(s.d_img_bytes.p);
    b.img_out_bytes = static_cast<uint32_t*>(s.d_img_bytes.p) + m;
    b.cta_map_dct = reinterpret_cast<const uint32_t*>(dm + off_maps);
    b.cta_map_emit = reinterpret_cast<const uint32_t*>(dm + off_maps + map_bytes);
    b.seg_map = reinterpret_cast<const uint32_t*>(dm + off_maps + 2 * map_bytes);
    b.cta_map_stuff = reinterpret_cast<const uint32_t*>(static_cast<uint8_t*>(s.d_dyn.p) + align_up(sizeof(EncImageDyn) * m, 256));

    EncLaunchDims d;
    d.num_images = m;
    d.dct_ctas = map_dct.back();
    d.emit_ctas = map_emit.back();
    d.total_segments = total_segs;
    uint64_t launches = 0;
    launches += launch_enc_dct(b, d, s.stream);
    if (s.profile)
        cudaEventRecord(s.ev[1], s.stream);
    launches += launch_enc_scan(b, d, s.stream);
    B2J_CUDA(cudaGetLastError());
    uint32_t* hsizes = static_cast<uint32_t*>(s.h_sizes.p);
    B2J_CUDA(cudaMemcpyAsync(hsizes, b.img_stream_bytes, sizeof(uint32_t) * m, cudaMemcpyDeviceToHost, s.stream));
    B2J_CUDA(cudaStreamSynchronize(s.stream)); // the stream lengths decide the layout of everything that follows
    const double t_half = now();

    // ---- second half: layout of the streams, Huffman emit, stuffing ----
    EncImageDyn* dyn = static_cast<EncImageDyn*>(s.h_dyn.p);
    uint32_t* map_stuff = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(s.h_dyn.p) + align_up(sizeof(EncImageDyn) * m, 256));
    uint64_t stream_bytes = 0, out_bytes = 0;
    uint32_t chunks = 0;
    for (int k = 0; k < m; ++k) {
        dyn[k].stream_off = stream_bytes;
        dyn[k].stream_bytes = hsizes[k];
        dyn[k].chunk_first = chunks;
        dyn[k].out_off = out_bytes;
        map_stuff[k] = chunks;
        stream_bytes += align_up(size_t(hsizes[k]) + 16, 256);
        out_bytes += align_up(2 * size_t(hsizes[k]) + 2 * size_t(images[k].num_segments) + 16, 256);
        chunks += (hsizes[k] + kEncStuffChunk - 1) / kEncStuffChunk;
    }
    map_stuff[m] = chunks;
    if ((rc = s.ensure(s.d_stream, stream_bytes, false)) || (rc = s.ensure(s.d_out, out_bytes, false)) ||
        (rc = s.ensure(s.d_chunks, size_t(chunks) * 4 + 4, false)))
        return rc;
    b.stream = static_cast<uint8_t*>(s.d_stream.p);
    b.out = static_cast<uint8_t*>(s.d_out.p);
    b.chunk_out = static_cast<uint32_t*>(s.d_chunks.p);
    d.stuff_ctas = chunks;
    B2J_CUDA(cudaMemcpyAsync(s.d_dyn.p, s.h_dyn.p, dyn_bytes, cudaMemcpyHostToDevice, s.stream));
    if (s.profile)
        cudaEventRecord(s.ev[2], s.stream);
    launches += launch_enc_emit(b, d, s.stream);
    if (s.profile)
        cudaEventRecord(s.ev[3], s.stream);
    launches += launch_enc_stuff(b, d, s.stream);
    if (s.profile)
        cudaEventRecord(s.ev[4], s.stream);
    B2J_CUDA(cudaGetLastError());
    B2J_CUDA(cudaMemcpyAsync(hsizes + m, b.img_out_bytes, sizeof(uint32_t) * m, cudaMemcpyDeviceToHost, s.stream));
    B2J_CUDA(cudaStreamSynchronize(s.stream));

    const double t_dev = now();
    // ---- headers (host) + entropy-coded bytes (device -> caller's buffer) + EOI ----
    // page-locked destinations receive the bytes directly; pageable ones go through a pinned bounce buffer and a threaded
    // host copy (an asynchronous copy into pageable memory would be staged by the driver one image at a time)
    uint64_t total_out = 0, bounce_bytes = 0;
    std::vector<uint64_t> bounce_off(m, UINT64_MAX);
    std::vector<char> ok(m, 0);
    for (int k = 0; k < m; ++k) {
        const int bi = batch_index[k];
        const size_t hdr = headers[k].size(), body = hsizes[m + k], need = hdr + body + 2;
        out_sizes[bi] = need;
        if (need > out_capacity[bi]) {
            statuses[bi] = B200JPEG_STATUS_BUFFER_TOO_SMALL;
            continue;
        }
        ok[k] = 1;
        total_out += need;
        if (!is_pinned(out[bi])) {
            bounce_off[k] = bounce_bytes;
            bounce_bytes += align_up(body, 256);
        }
    }
    if (bounce_bytes && (rc = s.ensure(s.h_out, bounce_bytes, true)))
        return rc;
    for (int k = 0; k < m; ++k) {
        if (!ok[k])
            continue;
        const int bi = batch_index[k];
        const size_t hdr = headers[k].size(), body = hsizes[m + k];
        uint8_t* dst = bounce_off[k] == UINT64_MAX ? out[bi] + hdr : static_cast<uint8_t*>(s.h_out.p) + bounce_off[k];
        B2J_CUDA(cudaMemcpyAsync(dst, static_cast<uint8_t*>(s.d_out.p) + dyn[k].out_off, body, cudaMemcpyDeviceToHost, s.stream));
        std::memcpy(out[bi], headers[k].data(), hdr);
        out[bi][hdr + body] = 0xFF;
        out[bi][hdr + body + 1] = 0xD9;
    }
    if (bounce_bytes) {
        B2J_CUDA(cudaStreamSynchronize(s.stream));
        parallel_for(m, 8, [&](int k) {
            if (ok[k] && bounce_off[k] != UINT64_MAX)
                std::memcpy(out[batch_index[k]] + headers[k].size(), static_cast<uint8_t*>(s.h_out.p) + bounce_off[k], hsizes[m + k]);
        });
    }
    B2J_CUDA(cudaStreamSynchronize(s.stream));
    if (s.profile)
        for (int i = 0; i < 4; ++i)
            cudaEventElapsedTime(&s.stage_ms[i], s.ev[i], s.ev[i + 1]);
    if (trace)
        std::fprintf(stderr, "b200jpeg encode trace: plan %.3f, first half (pixels up, DCT, scan) %.3f, second half (emit, stuff) %.3f, copy out %.3f ms\n",
            t_plan - t_begin, t_half - t_plan, t_dev - t_half, now() - t_dev);
    s.stats[0] = launches;
    s.stats[1] = total_blocks;
    s.stats[2] = in_bytes;
    s.stats[3] = total_out;
    s.stats[4] = stream_bytes;
    s.stats[5] = uint64_t(s.stage_ms[0] * 1e3f);
    s.stats[6] = uint64_t((s.stage_ms[1] + s.stage_ms[2]) * 1e3f);
    s.stats[7] = uint64_t(s.stage_ms[3] * 1e3f);
    return B200JPEG_STATUS_SUCCESS;
}

void Encoder::get_stats(uint64_t* s8) const
{
    for (int i = 0; i < 8; ++i)
        s8[i] = impl_->stats[i];
}

} // namespace b200jpeg
