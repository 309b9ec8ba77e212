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
constexpr int kEncGroups = 4; // a batch is cut into up to this many groups of images, each on its own stream
} // namespace

struct Encoder::Impl
{
    int device = 0;
    bool profile = false;
    int groups = kEncGroups;
    cudaStream_t stream = nullptr;           // group 0's stream; the others follow
    cudaStream_t gstream[kEncGroups] = {};
    cudaEvent_t ev_in = nullptr, ev_meta = nullptr;
    cudaEvent_t ev_half[kEncGroups] = {}, ev_sizes[kEncGroups] = {}, ev_done[kEncGroups] = {};
    cudaEvent_t ev[8] = {};
    Buf d_nzmask, d_pixels, d_planes, d_meta, d_dyn, d_coef, d_dc, d_acbits, d_blk_off, d_tile_off, d_seg_off, d_img_bytes;
    Buf d_stream[kEncGroups], d_chunks[kEncGroups], d_out[kEncGroups]; // per group: they grow while other groups are in flight
    Buf h_pixels, h_meta, h_dyn, h_sizes, h_out;
    uint64_t stats[8] = {};
    float stage_ms[4] = {};

    Allocators alloc;
    bool allocated = false;

    int ensure(Buf& b, size_t bytes, bool pinned)
    {
        if (bytes <= b.cap)
            return 0;
        if (b.p) { // nothing in flight may still use the buffer: whole-batch buffers are sized before the first launch of a call,
            // per-group ones are only used by their own group, whose stream has drained when it gets here again
            B2J_CUDA(cudaDeviceSynchronize());
            alloc.release(pinned, b.p, b.cap, stream);
            b.p = nullptr;
            b.cap = 0;
        }
        size_t want = align_up(bytes + bytes / 8 + 256, 1 << 16);
        if (!alloc.alloc(pinned, &b.p, &want, stream))
            return B200JPEG_STATUS_ALLOC_FAILED;
        B2J_CUDA(cudaStreamSynchronize(stream)); // a stream-ordered allocation is used on every group's stream
        b.cap = want;
        allocated = true;
        return 0;
    }
};

Encoder::Encoder(int device_id, const char* options) : impl_(new Impl)
{
    impl_->device = device_id;
    const std::string o = options ? options : "";
    if (o.find("profile=1") != std::string::npos)
        impl_->profile = true;
    const size_t g = o.find("groups=");
    if (g != std::string::npos)
        impl_->groups = std::max(1, std::min(kEncGroups, std::atoi(o.c_str() + g + 7)));
}

Encoder::~Encoder()
{
    Impl& s = *impl_;
    cudaSetDevice(s.device);
    Buf* dev[] = {&s.d_nzmask, &s.d_pixels, &s.d_planes, &s.d_meta, &s.d_dyn, &s.d_coef, &s.d_dc, &s.d_acbits, &s.d_blk_off, &s.d_tile_off, &s.d_seg_off, &s.d_img_bytes};
    cudaDeviceSynchronize();
    for (Buf* b : dev)
        s.alloc.release(false, b->p, b->cap, s.stream);
    for (int g = 0; g < kEncGroups; ++g)
        for (Buf* b : {&s.d_stream[g], &s.d_chunks[g], &s.d_out[g]})
            s.alloc.release(false, b->p, b->cap, s.stream);
    Buf* host[] = {&s.h_pixels, &s.h_meta, &s.h_dyn, &s.h_sizes, &s.h_out};
    for (Buf* b : host)
        s.alloc.release(true, b->p, b->cap, s.stream);
    if (s.stream)
        cudaStreamSynchronize(s.stream); // stream-ordered frees
    for (cudaEvent_t e : s.ev)
        if (e)
            cudaEventDestroy(e);
    for (cudaEvent_t e : {s.ev_in, s.ev_meta})
        if (e)
            cudaEventDestroy(e);
    for (int g = 0; g < kEncGroups; ++g) {
        for (cudaEvent_t e : {s.ev_half[g], s.ev_sizes[g], s.ev_done[g]})
            if (e)
                cudaEventDestroy(e);
        if (s.gstream[g])
            cudaStreamDestroy(s.gstream[g]);
    }
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
    // earlier groups run at higher priority: the block scheduler then finishes group g's kernels before it spends SMs on
    // group g + 1 (equal priorities would interleave the groups' CTAs and every group would finish at the end)
    int prio_lo = 0, prio_hi = 0;
    B2J_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    for (int g = 0; g < kEncGroups; ++g) {
        B2J_CUDA(cudaStreamCreateWithPriority(&s.gstream[g], cudaStreamNonBlocking, std::max(prio_hi, std::min(prio_lo, prio_lo - (kEncGroups - 1 - g)))));
        B2J_CUDA(cudaEventCreate(&s.ev_half[g])); // with timing: B200JPEG_TRACE prints the groups' timeline
        B2J_CUDA(cudaEventCreate(&s.ev_sizes[g]));
        B2J_CUDA(cudaEventCreate(&s.ev_done[g]));
    }
    s.stream = s.gstream[0];
    B2J_CUDA(cudaEventCreateWithFlags(&s.ev_in, cudaEventDisableTiming));
    B2J_CUDA(cudaEventCreateWithFlags(&s.ev_meta, cudaEventDisableTiming));
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
    std::vector<uint32_t> map_dct(1, 0), map_emit(1, 0);
    uint32_t total_segs = 0;
    std::vector<size_t> host_pix_off; // staging offset of host inputs (SIZE_MAX for device inputs)
    size_t host_pix_bytes = 0;
    uint64_t total_blocks = 0, in_bytes = 0;
    size_t plane_bytes = 0;
    int prev_quality = -1;
    images.reserve(n);
    headers.reserve(n);
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
        // tables and header bytes depend on (quality, geometry) only: a run of equal images shares them
        const bool same = !images.empty() && prev_quality == in[i].quality && images.back().width == im.width && images.back().height == im.height &&
            images.back().ncomp == im.ncomp && images.back().hs == im.hs && images.back().vs == im.vs &&
            images.back().restart_interval == im.restart_interval;
        if (same) {
            im.qt = images.back().qt;
            headers.push_back(headers.back());
        } else {
            im.qt = int32_t(quant.size());
            quant.emplace_back();
            enc_build_quant(in[i].quality, 0, quant.back());
            quant.emplace_back();
            enc_build_quant(in[i].quality, 1, quant.back());
            headers.emplace_back();
            enc_write_headers(im, quant[im.qt], quant[im.qt + 1], headers.back());
        }
        prev_quality = in[i].quality;
        im.color_cta_first = map_dct.back();
        {
            const size_t ybytes = size_t(im.plane_pitch[0]) * size_t(im.mcus_y) * 8 * (im.ncomp == 3 ? im.vs : 1);
            const size_t cbytes = size_t(im.plane_pitch[1]) * size_t(im.mcus_y) * 8;
            im.plane_off[0] = plane_bytes;
            im.plane_off[1] = im.plane_off[0] + align_up(ybytes, 256);
            im.plane_off[2] = im.plane_off[1] + align_up(cbytes, 256);
            plane_bytes = im.plane_off[2] + align_up(cbytes, 256);
        }
        im.emit_cta_first = map_emit.back();
        im.blk_first = total_blocks;
        im.seg_first = total_segs;
        map_dct.push_back(map_dct.back() + uint32_t((uint64_t(im.plane_pitch[0] / 4) * uint64_t(im.mcus_y) * 8 + kEncColorCta - 1) / kEncColorCta));
        map_emit.push_back(map_emit.back() + (im.total_blocks + kEncCta - 1) / kEncCta);
        total_segs += im.num_segments;
        total_blocks += im.total_blocks;
        images.push_back(im);
        batch_index.push_back(i);
    }
    const int m = int(images.size());
    if (m == 0)
        return B200JPEG_STATUS_SUCCESS;
    const double t_plan = now();

    // ---- groups of consecutive images, balanced by blocks; each runs on its own stream so that the uploads, the two
    // host round trips (stream sizes, stuffed sizes) and the copy-out of one group overlap the kernels of the others ----
    const int G = s.profile ? 1 : std::max(1, std::min(s.groups, m));
    // The groups shrink (weights G+1, G, .., 2): what is left when the GPU is done is the last group's copy-out.
    std::vector<int> gfirst(G + 1, m);
    gfirst[0] = 0;
    {
        int k = 0;
        uint64_t acc = 0, wsum = 0, wtot = 0;
        for (int g = 0; g < G; ++g)
            wtot += uint64_t(G + 1 - g);
        for (int g = 1; g < G; ++g) {
            wsum += uint64_t(G + 2 - g);
            const uint64_t target = total_blocks * wsum / wtot;
            while (k < m - (G - g) && (k < gfirst[g - 1] + 1 || acc + images[k].total_blocks / 2 <= target))
                acc += images[k++].total_blocks;
            gfirst[g] = k;
        }
    }

    // ---- scratch (whole batch; sized before anything is launched) ----
    int rc = 0;
    const size_t off_img = 0;
    const size_t off_quant = align_up(off_img + sizeof(EncImageDesc) * m, 256);
    const size_t off_huff = align_up(off_quant + sizeof(EncQuantTable) * quant.size(), 256);
    const size_t map_bytes = align_up(sizeof(uint32_t) * (m + 1), 256);
    const size_t off_maps = align_up(off_huff + sizeof(EncHuffTables), 256);
    const size_t meta_bytes = off_maps + 2 * map_bytes;
    const size_t dyn_arr_bytes = align_up(sizeof(EncImageDyn) * m, 256);
    const size_t dyn_bytes = dyn_arr_bytes + sizeof(uint32_t) * (m + G); // per group a chunk map of (images + 1) entries
    if ((rc = s.ensure(s.d_meta, meta_bytes, false)) || (rc = s.ensure(s.h_meta, meta_bytes, true)) || (rc = s.ensure(s.d_dyn, dyn_bytes, false)) ||
        (rc = s.ensure(s.h_dyn, dyn_bytes, true)) || (rc = s.ensure(s.d_coef, total_blocks * 128, false)) || (rc = s.ensure(s.d_planes, plane_bytes, false)) ||
        (rc = s.ensure(s.d_dc, total_blocks * 2, false)) || (rc = s.ensure(s.d_acbits, total_blocks * 2, false)) || (rc = s.ensure(s.d_nzmask, total_blocks * 8, false)) ||
        (rc = s.ensure(s.d_blk_off, total_blocks * 4, false)) ||
        (rc = s.ensure(s.d_tile_off, size_t(map_emit.back()) * 4 + size_t(m) * 4, false)) ||
        (rc = s.ensure(s.d_seg_off, size_t(total_segs) * 4, false)) || (rc = s.ensure(s.d_img_bytes, size_t(m) * 8, false)) ||
        (rc = s.ensure(s.h_sizes, size_t(m) * 8, true)) || (host_pix_bytes && (rc = s.ensure(s.d_pixels, host_pix_bytes, false))))
        return rc;
    bool any_staged = false;
    std::vector<char> staged(m, 0);
    for (int k = 0; k < m; ++k)
        if (host_pix_off[k] != SIZE_MAX && !is_pinned(images[k].pixels))
            staged[k] = 1, any_staged = true;
    if (any_staged && (rc = s.ensure(s.h_pixels, host_pix_bytes, true)))
        return rc;
    auto pixel_bytes = [&](const EncImageDesc& im) { return size_t(im.stride) * im.height * (im.in_fmt == B200JPEG_FMT_P_RGB ? 3 : 1); };
    std::vector<const uint8_t*> host_src(m, nullptr);
    for (int k = 0; k < m; ++k)
        if (host_pix_off[k] != SIZE_MAX) {
            host_src[k] = images[k].pixels;
            images[k].pixels = static_cast<const uint8_t*>(s.d_pixels.p) + host_pix_off[k];
        }

    // ---- descriptors, tables, maps ----
    B2J_CUDA(cudaEventRecord(s.ev_in, user_stream));
    if (trace)
        cudaEventRecord(s.ev[5], s.gstream[0]);
    uint8_t* hm = static_cast<uint8_t*>(s.h_meta.p);
    std::memcpy(hm + off_img, images.data(), sizeof(EncImageDesc) * m);
    std::memcpy(hm + off_quant, quant.data(), sizeof(EncQuantTable) * quant.size());
    enc_build_huff(*reinterpret_cast<EncHuffTables*>(hm + off_huff));
    std::memcpy(hm + off_maps, map_dct.data(), sizeof(uint32_t) * (m + 1));
    std::memcpy(hm + off_maps + map_bytes, map_emit.data(), sizeof(uint32_t) * (m + 1));
    B2J_CUDA(cudaMemcpyAsync(s.d_meta.p, s.h_meta.p, meta_bytes, cudaMemcpyHostToDevice, s.gstream[0]));
    B2J_CUDA(cudaEventRecord(s.ev_meta, s.gstream[0]));

    uint8_t* dm = static_cast<uint8_t*>(s.d_meta.p);
    uint32_t* hsizes = static_cast<uint32_t*>(s.h_sizes.p);
    auto group_view = [&](int g, EncBatchDev& b, EncLaunchDims& d) {
        const int k0 = gfirst[g], cnt = gfirst[g + 1] - k0;
        b = EncBatchDev{};
        b.images = reinterpret_cast<const EncImageDesc*>(dm + off_img) + k0;
        b.dyn = static_cast<const EncImageDyn*>(s.d_dyn.p) + k0;
        b.num_images = cnt;
        b.quant = reinterpret_cast<const EncQuantTable*>(dm + off_quant);
        b.huff = reinterpret_cast<const EncHuffTables*>(dm + off_huff);
        b.planes = static_cast<uint8_t*>(s.d_planes.p);
        b.coef = static_cast<int16_t*>(s.d_coef.p);
        b.dc = static_cast<int16_t*>(s.d_dc.p);
        b.acbits = static_cast<uint16_t*>(s.d_acbits.p);
        b.nzmask = static_cast<uint64_t*>(s.d_nzmask.p);
        b.blk_off = static_cast<uint32_t*>(s.d_blk_off.p);
        b.tile_off = static_cast<uint32_t*>(s.d_tile_off.p);
        b.img_bits = b.tile_off + map_emit.back() + k0;
        b.seg_byte_off = static_cast<uint32_t*>(s.d_seg_off.p);
        b.img_stream_bytes = static_cast<uint32_t*>(s.d_img_bytes.p) + k0;
        b.img_out_bytes = static_cast<uint32_t*>(s.d_img_bytes.p) + m + k0;
        b.cta_map_color = reinterpret_cast<const uint32_t*>(dm + off_maps) + k0;
        b.cta_map_emit = reinterpret_cast<const uint32_t*>(dm + off_maps + map_bytes) + k0;
        b.cta_map_stuff = reinterpret_cast<const uint32_t*>(static_cast<uint8_t*>(s.d_dyn.p) + dyn_arr_bytes) + k0 + g;
        b.cta_base_color = map_dct[k0];
        b.cta_base_emit = map_emit[k0];
        b.stream = static_cast<uint8_t*>(s.d_stream[g].p);
        b.out = static_cast<uint8_t*>(s.d_out[g].p);
        b.chunk_out = static_cast<uint32_t*>(s.d_chunks[g].p);
        d = EncLaunchDims{};
        d.num_images = cnt;
        d.color_ctas = map_dct[k0 + cnt] - map_dct[k0];
        d.emit_ctas = map_emit[k0 + cnt] - map_emit[k0];
    };

    // ---- first half per group: pixels up, colour + DCT + quantisation, bit offsets, stream sizes back ----
    uint64_t launches = 0;
    for (int g = 0; g < G; ++g) {
        cudaStream_t st = s.gstream[g];
        const int k0 = gfirst[g], k1 = gfirst[g + 1];
        B2J_CUDA(cudaStreamWaitEvent(st, s.ev_in, 0));
        if (g)
            B2J_CUDA(cudaStreamWaitEvent(st, s.ev_meta, 0));
        if (s.profile)
            cudaEventRecord(s.ev[0], st);
        // page-locked inputs go up directly; pageable ones pass through the pinned staging buffer (filled by a few threads)
        std::vector<int> todo;
        for (int k = k0; k < k1; ++k)
            if (staged[k])
                todo.push_back(k);
        parallel_for(int(todo.size()), 8, [&](int j) {
            const int k = todo[j];
            std::memcpy(static_cast<uint8_t*>(s.h_pixels.p) + host_pix_off[k], host_src[k], pixel_bytes(images[k]));
        });
        for (int k = k0; k < k1; ++k)
            if (host_src[k])
                B2J_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(s.d_pixels.p) + host_pix_off[k],
                    staged[k] ? static_cast<uint8_t*>(s.h_pixels.p) + host_pix_off[k] : host_src[k], pixel_bytes(images[k]), cudaMemcpyHostToDevice,
                    st));
        EncBatchDev b;
        EncLaunchDims d;
        group_view(g, b, d);
        launches += launch_enc_dct(b, d, st);
        if (s.profile)
            cudaEventRecord(s.ev[1], st);
        launches += launch_enc_scan(b, d, st);
        B2J_CUDA(cudaGetLastError());
        B2J_CUDA(cudaMemcpyAsync(hsizes + k0, b.img_stream_bytes, sizeof(uint32_t) * (k1 - k0), cudaMemcpyDeviceToHost, st));
        B2J_CUDA(cudaEventRecord(s.ev_half[g], st));
    }
    const double t_half = now();

    // ---- the rest is driven by the groups' events, whichever comes first:
    //   stream sizes on the host  -> layout of the group's streams, Huffman emit, stuffing, stuffed sizes back
    //   stuffed sizes on the host -> headers (host) + entropy-coded bytes (device -> caller's buffer) + EOI
    //   bytes on the host         -> copy out of the bounce buffer where the destination is pageable
    // Page-locked destinations receive the bytes directly; pageable ones go through a pinned bounce buffer and a threaded
    // host copy (an asynchronous copy into pageable memory would be staged by the driver one image at a time).
    EncImageDyn* dyn = static_cast<EncImageDyn*>(s.h_dyn.p);
    uint32_t* map_stuff_all = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(s.h_dyn.p) + dyn_arr_bytes);
    uint64_t stream_bytes_total = 0, total_out = 0;
    std::vector<char> ok(m, 0), pinned_out(m, 0);
    bool any_bounce = false;
    for (int k = 0; k < m; ++k) {
        pinned_out[k] = is_pinned(out[batch_index[k]]);
        any_bounce = any_bounce || !pinned_out[k];
    }
    std::vector<uint64_t> bounce_off(m, 0);
    if (any_bounce) { // reserve by capacity: the stuffed sizes are not known yet
        uint64_t bytes = 0;
        for (int k = 0; k < m; ++k)
            if (!pinned_out[k]) {
                bounce_off[k] = bytes;
                bytes += align_up(out_capacity[batch_index[k]], 256);
            }
        if ((rc = s.ensure(s.h_out, bytes, true)))
            return rc;
    }
    auto second_half = [&](int g) -> int {
        cudaStream_t st = s.gstream[g];
        const int k0 = gfirst[g], k1 = gfirst[g + 1];
        uint32_t* map_stuff = map_stuff_all + k0 + g;
        uint64_t stream_bytes = 0, out_bytes = 0;
        uint32_t chunks = 0;
        for (int k = k0; k < k1; ++k) {
            dyn[k].stream_off = stream_bytes;
            dyn[k].stream_bytes = hsizes[k];
            dyn[k].chunk_first = chunks;
            dyn[k].out_off = out_bytes;
            map_stuff[k - k0] = chunks;
            stream_bytes += align_up(size_t(hsizes[k]) + 16, 256);
            out_bytes += align_up(2 * size_t(hsizes[k]) + 2 * size_t(images[k].num_segments) + 16, 256);
            chunks += (hsizes[k] + kEncStuffChunk - 1) / kEncStuffChunk;
        }
        map_stuff[k1 - k0] = chunks;
        stream_bytes_total += stream_bytes;
        int r = 0;
        if ((r = s.ensure(s.d_stream[g], stream_bytes, false)) || (r = s.ensure(s.d_out[g], out_bytes, false)) ||
            (r = s.ensure(s.d_chunks[g], size_t(chunks) * 4 + 4, false)))
            return r;
        EncBatchDev b;
        EncLaunchDims d;
        group_view(g, b, d);
        d.stuff_ctas = chunks;
        B2J_CUDA(cudaMemcpyAsync(static_cast<EncImageDyn*>(s.d_dyn.p) + k0, dyn + k0, sizeof(EncImageDyn) * (k1 - k0), cudaMemcpyHostToDevice, st));
        B2J_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(s.d_dyn.p) + dyn_arr_bytes + sizeof(uint32_t) * (k0 + g), map_stuff,
            sizeof(uint32_t) * (k1 - k0 + 1), cudaMemcpyHostToDevice, st));
        if (s.profile)
            cudaEventRecord(s.ev[2], st);
        launches += launch_enc_emit(b, d, st);
        if (s.profile)
            cudaEventRecord(s.ev[3], st);
        launches += launch_enc_stuff(b, d, st);
        if (s.profile)
            cudaEventRecord(s.ev[4], st);
        B2J_CUDA(cudaGetLastError());
        B2J_CUDA(cudaMemcpyAsync(hsizes + m + k0, b.img_out_bytes, sizeof(uint32_t) * (k1 - k0), cudaMemcpyDeviceToHost, st));
        B2J_CUDA(cudaEventRecord(s.ev_sizes[g], st));
        return 0;
    };
    auto copy_out = [&](int g) -> int {
        cudaStream_t st = s.gstream[g];
        for (int k = gfirst[g]; k < gfirst[g + 1]; ++k) {
            const int bi = batch_index[k];
            const size_t hdr = headers[k].size(), body = hsizes[m + k], need = hdr + body + 2;
            out_sizes[bi] = need;
            if (need > out_capacity[bi]) {
                statuses[bi] = B200JPEG_STATUS_BUFFER_TOO_SMALL;
                continue;
            }
            ok[k] = 1;
            total_out += need;
            uint8_t* dst = pinned_out[k] ? out[bi] + hdr : static_cast<uint8_t*>(s.h_out.p) + bounce_off[k];
            B2J_CUDA(cudaMemcpyAsync(dst, static_cast<uint8_t*>(s.d_out[g].p) + dyn[k].out_off, body, cudaMemcpyDeviceToHost, st));
            std::memcpy(out[bi], headers[k].data(), hdr);
            out[bi][hdr + body] = 0xFF;
            out[bi][hdr + body + 1] = 0xD9;
        }
        B2J_CUDA(cudaEventRecord(s.ev_done[g], st));
        return 0;
    };
    auto unbounce = [&](int g) {
        if (!any_bounce)
            return;
        const int k0 = gfirst[g];
        parallel_for(gfirst[g + 1] - k0, 8, [&](int j) {
            const int k = k0 + j;
            if (ok[k] && !pinned_out[k])
                std::memcpy(out[batch_index[k]] + headers[k].size(), static_cast<uint8_t*>(s.h_out.p) + bounce_off[k], hsizes[m + k]);
        });
    };
    std::vector<int> state(G, 0); // 0: waits for the stream sizes, 1: for the stuffed sizes, 2: for the bytes, 3: finished
    double t_dev = 0;
    for (int finished = 0; finished < G;) {
        bool progress = false;
        for (int g = 0; g < G; ++g) {
            if (state[g] == 3)
                continue;
            cudaEvent_t e = state[g] == 0 ? s.ev_half[g] : state[g] == 1 ? s.ev_sizes[g] : s.ev_done[g];
            const cudaError_t q = cudaEventQuery(e);
            if (q == cudaErrorNotReady)
                continue;
            B2J_CUDA(q);
            progress = true;
            if (state[g] == 0) {
                if ((rc = second_half(g)))
                    return rc;
            } else if (state[g] == 1) {
                if ((rc = copy_out(g)))
                    return rc;
                if (g == G - 1)
                    t_dev = now();
            } else {
                unbounce(g);
                ++finished;
            }
            ++state[g];
        }
        if (!progress)
            std::this_thread::yield();
    }
    if (s.profile)
        for (int i = 0; i < 4; ++i)
            cudaEventElapsedTime(&s.stage_ms[i], s.ev[i], s.ev[i + 1]);
    if (trace)
        for (int g = 0; g < G; ++g) {
            float a = 0, b = 0, c = 0;
            cudaEventElapsedTime(&a, s.ev[5], s.ev_half[g]);
            cudaEventElapsedTime(&b, s.ev[5], s.ev_sizes[g]);
            cudaEventElapsedTime(&c, s.ev[5], s.ev_done[g]);
            std::fprintf(stderr, "b200jpeg encode trace: group %d (%d images): coefficients + sizes %.3f, streams stuffed %.3f, bytes on host %.3f ms\n", g,
                gfirst[g + 1] - gfirst[g], a, b, c);
        }
    if (trace)
        std::fprintf(stderr, "b200jpeg encode trace (%d groups): plan %.3f, first halves enqueued %.3f, last group's copy enqueued %.3f, all done %.3f ms\n",
            G, t_plan - t_begin, t_half - t_plan, t_dev - t_half, now() - t_dev);
    s.stats[0] = launches;
    s.stats[1] = total_blocks;
    s.stats[2] = in_bytes;
    s.stats[3] = total_out;
    s.stats[4] = stream_bytes_total;
    s.stats[5] = uint64_t(s.stage_ms[0] * 1e3f);
    s.stats[6] = uint64_t((s.stage_ms[1] + s.stage_ms[2]) * 1e3f);
    s.stats[7] = uint64_t(s.stage_ms[3] * 1e3f);
    return B200JPEG_STATUS_SUCCESS;
}

int Encoder::set_allocators(const Allocators& a)
{
    if (impl_->allocated)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    impl_->alloc = a;
    return B200JPEG_STATUS_SUCCESS;
}

void Encoder::get_stats(uint64_t* s8) const
{
    for (int i = 0; i < 8; ++i)
        s8[i] = impl_->stats[i];
}

} // namespace b200jpeg
