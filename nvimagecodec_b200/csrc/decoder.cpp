// decoder.cpp -- host side of the b200jpeg batch decoder: header parsing, batch planning, pinned staging, H2D, kernel
// launches on the decoder's own stream, status read-back.  Mirrors the role of DecoderImpl::decodeBatch in the reference
// (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:444-490) with the plumbing pattern of the reference's batched GPU
// plugin (extensions/nvjpeg/hw_decoder.cpp:492-680): gather bitstreams, one batched launch sequence, sync, report.
#include "decoder.h"
#include <sched.h>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#if defined(__x86_64__) || defined(_M_X64)
#include <emmintrin.h>
#endif
#include "../../include/b200jpeg.h"
#include "log.h"

namespace b200jpeg {

namespace {
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
#define B2J_CUDA(call)                                                                                         \
    do {                                                                                                       \
        cudaError_t e__ = (call);                                                                              \
        if (e__ != cudaSuccess) {                                                                              \
            log_message(kLogError, "CUDA error %s at %s:%d", cudaGetErrorString(e__), __FILE__, __LINE__); \
            return B200JPEG_STATUS_CUDA_ERROR;                                                                 \
        }                                                                                                      \
    } while (0)
} // namespace

DecoderOptions parse_decoder_options(const char* options, const char* plugin_id)
{
    DecoderOptions o;
    if (!options)
        return o;
    std::string s(options);
    size_t pos = 0;
    while (pos < s.size()) {
        while (pos < s.size() && s[pos] == ' ')
            ++pos;
        size_t end = s.find(' ', pos);
        if (end == std::string::npos)
            end = s.size();
        std::string tok = s.substr(pos, end - pos);
        pos = end;
        if (tok.empty())
            continue;
        std::string module, name = tok, value;
        const size_t colon = tok.find(':');
        if (colon != std::string::npos) {
            module = tok.substr(0, colon);
            name = tok.substr(colon + 1);
        }
        const size_t eq = name.find('=');
        if (eq == std::string::npos)
            continue;
        value = name.substr(eq + 1);
        name = name.substr(0, eq);
        if (!module.empty() && module != plugin_id)
            continue;
        if (name == "fancy_upsampling")
            o.fancy_upsampling = std::atoi(value.c_str()) != 0;
        else if (name == "subseq_bytes") {
            int v = std::atoi(value.c_str());
            if (v >= 32 && v <= 4096 && (v % 4) == 0)
                o.subseq_bytes = v;
        } else if (name == "tail_launches") {
            o.tail_launches = std::max(0, std::min(32, std::atoi(value.c_str())));
        } else if (name == "tail_spread") {
            o.tail_spread = std::atoi(value.c_str()) != 0;
        } else if (name == "stream_priority") {
            o.stream_priority = std::atoi(value.c_str()) != 0;
        } else if (name == "copy_threads")
            o.copy_threads = std::max(1, std::atoi(value.c_str()));
        else if (name == "groups")
            o.groups = std::max(1, std::atoi(value.c_str()));
        else if (name == "groups_host")
            o.groups_host = std::max(0, std::atoi(value.c_str()));
        else if (name == "stage_stream")
            o.stage_stream = std::atoi(value.c_str()) != 0;
        else if (name == "dump_coef")
            o.dump_coef = std::atoi(value.c_str()) != 0;
        else if (name == "profile")
            o.profile = std::atoi(value.c_str()) != 0;
        else if (name == "fused")
            o.fused = std::atoi(value.c_str()) != 0;
        else if (name == "fast_idct")
            o.fast_idct = std::atoi(value.c_str()) != 0;
    }
    return o;
}

Decoder::Decoder(int device_id, const DecoderOptions& opt) : device_(device_id), opt_(opt) {}

Decoder::~Decoder()
{
    if (device_ >= 0)
        cudaSetDevice(device_);
    DevBuf* dbufs[] = {&d_blkpos_, &d_stamp_, &d_raw_, &d_clean_, &d_coef_, &d_dc_, &d_planes_, &d_meta_, &d_states_, &d_seg_, &d_chunks_, &d_exit_, &d_entry_,
        &d_rec_, &d_abs_, &d_wl0_, &d_wl1_, &d_sync_, &d_out_, &d_fusedrec_, &d_dcrec_, &d_scanrec_};
    for (cudaStream_t st : streams_)
        if (st)
            cudaStreamSynchronize(st);
    for (DevBuf* b : dbufs)
        alloc_.release(false, b->p, b->cap, streams_[0]);
    DevBuf* hbufs[] = {&h_raw_, &h_meta_, &h_states_};
    for (DevBuf* b : hbufs)
        alloc_.release(true, b->p, b->cap, streams_[0]);
    if (streams_[0])
        cudaStreamSynchronize(streams_[0]); // stream-ordered frees
    for (cudaEvent_t e : ev_user_)
        cudaEventDestroy(e);
    for (cudaEvent_t e : ev_done_)
        if (e)
            cudaEventDestroy(e);
    for (cudaEvent_t e : ev_stage_)
        if (e)
            cudaEventDestroy(e);
    if (ev_meta_)
        cudaEventDestroy(ev_meta_);
    for (cudaStream_t st : streams_)
        if (st)
            cudaStreamDestroy(st);
}

int Decoder::init()
{
    int count = 0;
    B2J_CUDA(cudaGetDeviceCount(&count));
    if (device_ < 0 || device_ >= count)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    B2J_CUDA(cudaSetDevice(device_));
    cudaDeviceProp prop;
    B2J_CUDA(cudaGetDeviceProperties(&prop, device_));
    if (prop.major != 10) {
        log_message(kLogWarning, "device %d is sm_%d%d; this library only carries sm_100a code", device_, prop.major, prop.minor);
        return B200JPEG_STATUS_ARCH_MISMATCH;
    }
    // option stream_priority=1: earlier groups run at higher priority, i.e. the block scheduler finishes group g's kernels
    // before it spends SMs on group g + 1 (equal priorities interleave the groups' CTAs)
    int prio_lo = 0, prio_hi = 0;
    B2J_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    for (int g = 0; g < kMaxGroups; ++g) {
        const int prio = opt_.stream_priority ? std::max(prio_hi, prio_lo - (kMaxGroups - 1 - g)) : prio_lo;
        B2J_CUDA(cudaStreamCreateWithPriority(&streams_[g], cudaStreamNonBlocking, prio));
    }
    for (cudaEvent_t& e : ev_done_)
        B2J_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    B2J_CUDA(cudaEventCreateWithFlags(&ev_meta_, cudaEventDisableTiming));
    B2J_CUDA(cudaDeviceGetAttribute(&sm_count_, cudaDevAttrMultiProcessorCount, device_));
    if (idct_configure() != 0) {
        log_message(kLogError, "cannot configure the IDCT kernel: %s", cudaGetErrorString(cudaGetLastError()));
        return B200JPEG_STATUS_CUDA_ERROR;
    }
    if (opt_.profile)
        for (cudaEvent_t& e : ev_stage_)
            B2J_CUDA(cudaEventCreate(&e));
    return B200JPEG_STATUS_SUCCESS;
}

int Decoder::ensure(DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap)
        return 0;
    if (b.p) {
        for (cudaStream_t st : streams_)
            B2J_CUDA(cudaStreamSynchronize(st));
        alloc_.release(false, b.p, b.cap, streams_[0]);
        scratch_bytes_ -= b.cap;
        b.p = nullptr;
        b.cap = 0;
    }
    size_t want = align_up(bytes + bytes / 8, 1 << 20);
    if (!alloc_.alloc(false, &b.p, &want, streams_[0]))
        return B200JPEG_STATUS_ALLOC_FAILED;
    B2J_CUDA(cudaStreamSynchronize(streams_[0])); // a stream-ordered allocation is used on every group's stream
    b.cap = want;
    scratch_bytes_ += want;
    return 0;
}

int Decoder::set_allocators(const Allocators& a)
{
    if (scratch_bytes_ != 0 || h_raw_.p || h_meta_.p || h_states_.p)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    alloc_ = a;
    return B200JPEG_STATUS_SUCCESS;
}

int Decoder::ensure_pinned(DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap)
        return 0;
    if (b.p) {
        for (cudaStream_t st : streams_)
            B2J_CUDA(cudaStreamSynchronize(st));
        alloc_.release(true, b.p, b.cap, streams_[0]);
        b.p = nullptr;
        b.cap = 0;
    }
    size_t want = align_up(bytes + bytes / 8, 1 << 20);
    if (!alloc_.alloc(true, &b.p, &want, streams_[0]))
        return B200JPEG_STATUS_ALLOC_FAILED;
    B2J_CUDA(cudaStreamSynchronize(streams_[0]));
    b.cap = want;
    return 0;
}

int Decoder::plan_batch(int n, const uint8_t* const* data, const size_t* sizes, const OutputDesc* outs, int32_t* statuses)
{
    NvtxRange nvtx("b200jpeg plan_batch");
    ready_ = false;
    groups_.clear();
    data_ = data;
    if (n < 0 || (n > 0 && (!data || !sizes || !outs || !statuses)))
        return B200JPEG_STATUS_INVALID_PARAMETER;
    B2J_CUDA(cudaSetDevice(device_));
    batch_size_ = n;
    parsed_.resize(n);
    pre_status_.assign(n, B200JPEG_STATUS_SUCCESS);
    planned_to_batch_.clear();
    host_outs_.clear();
    out_streams_.clear();
    any_default_stream_ = false;

    PlanOptions popt;
    popt.subseq_bytes = opt_.subseq_bytes;
    popt.fancy_upsampling = opt_.fancy_upsampling;
    plan_begin(plan_, popt);
    size_t dev_out_bytes = 0, dev_out_end = 0;
    std::vector<size_t> out_offsets;
    for (int i = 0; i < n; ++i) {
        const ParseResult r = parse_jpeg(data[i], sizes[i], parsed_[i]);
        if (r != PARSE_OK) {
            pre_status_[i] = B200JPEG_STATUS_BAD_JPEG;
            continue;
        }
        if (!is_supported(parsed_[i]) || !resolve_format(outs[i].format, parsed_[i].ncomp)) {
            pre_status_[i] = B200JPEG_STATUS_UNSUPPORTED;
            continue;
        }
        if (!outs[i].buffer) {
            pre_status_[i] = B200JPEG_STATUS_INVALID_PARAMETER;
            continue;
        }
        OutputSpec os;
        os.buffer = outs[i].buffer;
        os.row_stride = outs[i].row_stride;
        os.format = outs[i].format;
        int out_w = parsed_[i].width, out_h = parsed_[i].height;
        if (outs[i].roi_w > 0 || outs[i].roi_h > 0) { // same validity rule as the reference (jpeg_mem.cpp IsCropWindowValid:94-100)
            if (outs[i].roi_w <= 0 || outs[i].roi_h <= 0 || outs[i].roi_x < 0 || outs[i].roi_y < 0 ||
                int64_t(outs[i].roi_x) + outs[i].roi_w > parsed_[i].width || int64_t(outs[i].roi_y) + outs[i].roi_h > parsed_[i].height) {
                pre_status_[i] = B200JPEG_STATUS_INVALID_PARAMETER;
                continue;
            }
            os.roi_x = outs[i].roi_x;
            os.roi_y = outs[i].roi_y;
            os.roi_w = out_w = outs[i].roi_w;
            os.roi_h = out_h = outs[i].roi_h;
            if (outs[i].orient) { // orientation and region together are not offered
                pre_status_[i] = B200JPEG_STATUS_UNSUPPORTED;
                continue;
            }
        }
        os.orient = outs[i].orient & 15;
        if (os.orient & 1) // odd quarter turns: the buffer is height x width
            std::swap(out_w, out_h);
        { // rows must not overlap (a short stride would make the colour kernel's stores, or the copy back, run into each other)
            const int fmt0 = resolve_format(outs[i].format, parsed_[i].ncomp);
            const int64_t row_bytes = int64_t(out_w) * ((fmt0 == 4 || fmt0 == 6) ? 3 : 1);
            if (outs[i].row_stride < row_bytes) {
                pre_status_[i] = B200JPEG_STATUS_INVALID_PARAMETER;
                continue;
            }
        }
        if (outs[i].has_stream) {
            if (std::find(out_streams_.begin(), out_streams_.end(), outs[i].stream) == out_streams_.end())
                out_streams_.push_back(outs[i].stream);
        } else
            any_default_stream_ = true;
        if (outs[i].on_host) {
            const int fmt = resolve_format(outs[i].format, parsed_[i].ncomp);
            const bool interleaved = (fmt == 4 || fmt == 6);
            HostOutput ho;
            ho.image = int(plan_.images.size());
            ho.host = outs[i].buffer;
            ho.host_stride = outs[i].row_stride;
            ho.row_bytes = size_t(out_w) * (interleaved ? 3 : 1);
            ho.rows = size_t(out_h) * (interleaved ? 1 : (fmt == 7 ? 1 : 3));
            // The device-side copy mirrors the host layout whenever that is tight (rows back to back, and images back to
            // back as in a batch tensor): the copy back is then one linear transfer per run of outputs.  A copy of short
            // rows with slack (cudaMemcpy2D) reaches about two thirds of the linear PCIe rate.
            const bool tight = size_t(ho.host_stride) == ho.row_bytes;
            ho.dev_pitch = tight ? ho.row_bytes : align_up(ho.row_bytes, 16);
            bool follows = false;
            if (tight && !host_outs_.empty()) {
                const HostOutput& pv = host_outs_.back();
                follows = pv.dev_pitch == pv.row_bytes && static_cast<uint8_t*>(pv.host) + pv.row_bytes * pv.rows == static_cast<uint8_t*>(ho.host) &&
                          pv.dev_off + pv.row_bytes * pv.rows == dev_out_end;
            }
            ho.dev_off = follows ? dev_out_end : dev_out_bytes;
            dev_out_end = ho.dev_off + ho.dev_pitch * ho.rows;
            dev_out_bytes = align_up(dev_out_end, 256);
            const size_t pitch = ho.dev_pitch;
            host_outs_.push_back(ho);
            os.buffer = nullptr; // patched below once d_out_ is allocated
            os.row_stride = int64_t(pitch);
        }
        plan_add_image(plan_, parsed_[i], os, popt);
        planned_to_batch_.push_back(i);
    }
    plan_finish(plan_);
    for (int i = 0; i < n; ++i)
        statuses[i] = pre_status_[i];
    const int m = int(plan_.images.size());
    if (m == 0) {
        ready_ = true;
        return B200JPEG_STATUS_SUCCESS;
    }

    // ---- scratch ----
    int rc = 0;
    if ((rc = ensure(d_raw_, plan_.raw_bytes)) || (rc = ensure(d_clean_, plan_.clean_bytes)) ||
        ((opt_.dump_coef || !plan_.prog_images.empty()) && (rc = ensure(d_coef_, plan_.coef_elems * sizeof(int16_t)))) ||
        (rc = ensure(d_dc_, plan_.coef_elems / 64 * sizeof(int16_t) + 64)) ||
        (rc = ensure(d_blkpos_, plan_.coef_elems / 64 * sizeof(uint32_t) + 64)) || (rc = ensure(d_stamp_, 4 * size_t(plan_.total_subseq))) ||
        (rc = ensure(d_planes_, plan_.plane_bytes)) ||
        (rc = ensure(d_states_, sizeof(ImageState) * m)) || (rc = ensure(d_seg_, sizeof(uint32_t) * (plan_.total_segs + 1))) ||
        (rc = ensure(d_chunks_, sizeof(uint64_t) * plan_.total_chunks)) || (rc = ensure(d_exit_, 8 * size_t(plan_.total_subseq))) ||
        (rc = ensure(d_entry_, 8 * size_t(plan_.total_subseq))) || (rc = ensure(d_rec_, 4 * size_t(plan_.total_subseq))) ||
        (rc = ensure(d_abs_, 4 * size_t(plan_.total_subseq))) || (rc = ensure(d_wl0_, 4 * size_t(plan_.total_subseq))) ||
        (rc = ensure(d_wl1_, 4 * size_t(plan_.total_subseq))) || (rc = ensure(d_sync_, sizeof(SyncCounters) * kMaxGroups)) ||
        (rc = ensure(d_fusedrec_, 8 * size_t(plan_.map_fused.back()) + 8)) || (rc = ensure(d_dcrec_, 16 * size_t(plan_.map_dc.back()) + 16)) ||
        (rc = ensure(d_scanrec_, 8 * size_t(plan_.map_scan.back()) + 8)) ||
        (dev_out_bytes && (rc = ensure(d_out_, dev_out_bytes))))
        return rc;
    for (const HostOutput& ho : host_outs_)
        plan_.images[ho.image].out = static_cast<uint8_t*>(d_out_.p) + ho.dev_off;

    // ---- metadata block: descriptors, tables, CTA maps -> one pinned buffer, one H2D ----
    const size_t off_img = 0;
    const size_t off_huff = align_up(off_img + sizeof(ImageDesc) * m, 256);
    const size_t off_quant = align_up(off_huff + sizeof(DevHuffTable) * plan_.huff.size(), 256);
    const size_t off_m0 = align_up(off_quant + plan_.quant.size() * sizeof(uint16_t), 256);
    const size_t map_bytes = align_up(sizeof(uint32_t) * (m + 1), 256);
    const size_t off_cimg = off_m0 + 7 * map_bytes;
    const size_t off_scans = off_cimg + align_up(sizeof(uint32_t) * (plan_.total_subseq / kSubseqCta + 1), 256);
    const size_t off_prog = off_scans + align_up(sizeof(ScanDesc) * plan_.scans.size(), 256);
    const size_t meta_bytes = off_prog + align_up(sizeof(uint32_t) * plan_.prog_images.size(), 256);
    if ((rc = ensure(d_meta_, meta_bytes)) || (rc = ensure_pinned(h_meta_, meta_bytes)) ||
        (rc = ensure_pinned(h_raw_, plan_.raw_bytes)) ||
        (rc = ensure_pinned(h_states_, sizeof(ImageState) * m + sizeof(SyncCounters) * kMaxGroups)))
        return rc;
    uint8_t* hm = static_cast<uint8_t*>(h_meta_.p);
    std::memcpy(hm + off_img, plan_.images.data(), sizeof(ImageDesc) * m);
    std::memcpy(hm + off_huff, plan_.huff.data(), sizeof(DevHuffTable) * plan_.huff.size());
    std::memcpy(hm + off_quant, plan_.quant.data(), plan_.quant.size() * sizeof(uint16_t));
    std::memcpy(hm + off_m0 + 0 * map_bytes, plan_.map_unstuff.data(), sizeof(uint32_t) * (m + 1));
    std::memcpy(hm + off_m0 + 1 * map_bytes, plan_.map_subseq.data(), sizeof(uint32_t) * (m + 1));
    std::memcpy(hm + off_m0 + 2 * map_bytes, plan_.map_idct.data(), sizeof(uint32_t) * (m + 1));
    std::memcpy(hm + off_m0 + 3 * map_bytes, plan_.map_color.data(), sizeof(uint32_t) * (m + 1));
    std::memcpy(hm + off_m0 + 4 * map_bytes, plan_.map_fused.data(), sizeof(uint32_t) * (m + 1));
    std::memcpy(hm + off_m0 + 5 * map_bytes, plan_.map_dc.data(), sizeof(uint32_t) * (m + 1));
    std::memcpy(hm + off_m0 + 6 * map_bytes, plan_.map_scan.data(), sizeof(uint32_t) * (m + 1));
    for (int i = 0; i < m; ++i) { // image of every subsequence CTA (the work-list rounds look items up by subsequence index)
        uint32_t* cimg = reinterpret_cast<uint32_t*>(hm + off_cimg) + plan_.images[i].subseq_first / kSubseqCta;
        for (uint32_t c = 0; c < plan_.images[i].subseq_cta_count; ++c)
            cimg[c] = uint32_t(i);
    }

    if (!plan_.scans.empty())
        std::memcpy(hm + off_scans, plan_.scans.data(), sizeof(ScanDesc) * plan_.scans.size());
    if (!plan_.prog_images.empty())
        std::memcpy(hm + off_prog, plan_.prog_images.data(), sizeof(uint32_t) * plan_.prog_images.size());
    B2J_CUDA(cudaMemcpyAsync(d_meta_.p, h_meta_.p, meta_bytes, cudaMemcpyHostToDevice, streams_[0]));
    B2J_CUDA(cudaEventRecord(ev_meta_, streams_[0]));

    uint8_t* dm = static_cast<uint8_t*>(d_meta_.p);
    dev_ = BatchDev{};
    dev_.images = reinterpret_cast<const ImageDesc*>(dm + off_img);
    dev_.states = static_cast<ImageState*>(d_states_.p);
    dev_.num_images = m;
    dev_.raw = static_cast<const uint8_t*>(d_raw_.p);
    dev_.clean = static_cast<uint8_t*>(d_clean_.p);
    dev_.huff = reinterpret_cast<const DevHuffTable*>(dm + off_huff);
    dev_.quant = reinterpret_cast<const uint16_t*>(dm + off_quant);
    dev_.seg_start = static_cast<uint32_t*>(d_seg_.p);
    dev_.chunk_counts = static_cast<uint64_t*>(d_chunks_.p);
    dev_.sub_exit = static_cast<unsigned long long*>(d_exit_.p);
    dev_.sub_entry = static_cast<unsigned long long*>(d_entry_.p);
    dev_.sub_rec = static_cast<uint32_t*>(d_rec_.p);
    dev_.sub_abs = static_cast<uint32_t*>(d_abs_.p);
    dev_.worklist[0] = static_cast<uint32_t*>(d_wl0_.p);
    dev_.worklist[1] = static_cast<uint32_t*>(d_wl1_.p);
    dev_.sync = static_cast<SyncCounters*>(d_sync_.p);
    dev_.coef = static_cast<int16_t*>(d_coef_.p);
    dev_.dump_coef = opt_.dump_coef ? 1 : 0;
    dev_.blk_pos = static_cast<uint32_t*>(d_blkpos_.p);
    dev_.sub_stamp = static_cast<uint32_t*>(d_stamp_.p);
    dev_.dc = static_cast<int16_t*>(d_dc_.p);
    dev_.planes = static_cast<uint8_t*>(d_planes_.p);
    dev_.cta_map_unstuff = reinterpret_cast<const uint32_t*>(dm + off_m0 + 0 * map_bytes);
    dev_.cta_map_subseq = reinterpret_cast<const uint32_t*>(dm + off_m0 + 1 * map_bytes);
    dev_.cta_map_fused = reinterpret_cast<const uint32_t*>(dm + off_m0 + 4 * map_bytes);
    dev_.fused_rec = static_cast<unsigned long long*>(d_fusedrec_.p);
    dev_.cta_map_dc = reinterpret_cast<const uint32_t*>(dm + off_m0 + 5 * map_bytes);
    dev_.dc_rec = static_cast<unsigned long long*>(d_dcrec_.p);
    dev_.cta_map_scan = reinterpret_cast<const uint32_t*>(dm + off_m0 + 6 * map_bytes);
    dev_.scan_rec = static_cast<unsigned long long*>(d_scanrec_.p);
    dev_.subseq_cta_img = reinterpret_cast<const uint32_t*>(dm + off_cimg);
    dev_.scans = reinterpret_cast<const ScanDesc*>(dm + off_scans);
    dev_.prog_images = reinterpret_cast<const uint32_t*>(dm + off_prog);
    dev_.image_base = 0;
    dev_.cta_map_idct = reinterpret_cast<const uint32_t*>(dm + off_m0 + 2 * map_bytes);
    dev_.cta_map_color = reinterpret_cast<const uint32_t*>(dm + off_m0 + 3 * map_bytes);
    dev_.subseq_bits = plan_.subseq_bits;
    dev_.num_tables = int(plan_.huff.size());
    dev_.lut_slots = 1;
    dev_.lut_ac_slots = 1;
    for (const ImageDesc& im : plan_.images) {
        dev_.lut_slots = std::max(dev_.lut_slots, im.ntabs);
        dev_.lut_ac_slots = std::max(dev_.lut_ac_slots, im.nac);
    }
    // stream staging in shared memory: per thread its subsequence plus the words a symbol may run over, an odd count
    dev_.tail_mode = (opt_.tail_spread ? 1 : 0);
    dev_.stream_words = 0;
    if (opt_.stage_stream && opt_.subseq_bytes <= 256)
#ifndef B2J_STREAM_EXTRA
#define B2J_STREAM_EXTRA 3
#endif
        dev_.stream_words = (plan_.subseq_bits / 32 + B2J_STREAM_EXTRA) | 1; // a symbol starting in the last word reads two words past it
    use_fused_ = opt_.fused && dev_.stream_words != 0;
    if (dev_.lut_slots != cfg_lut_slots_ || dev_.stream_words != cfg_stream_words_ || dev_.subseq_bits != cfg_subseq_bits_) {
        if (huff_configure(dev_, device_, &tail_capacity_) != 0 || tail_capacity_ <= 0 || (use_fused_ && fused_configure(dev_) != 0)) {
            log_message(kLogError, "cannot configure the Huffman kernels: %s", cudaGetErrorString(cudaGetLastError()));
            return B200JPEG_STATUS_CUDA_ERROR;
        }
        cfg_lut_slots_ = dev_.lut_slots;
        cfg_stream_words_ = dev_.stream_words;
        cfg_subseq_bits_ = dev_.subseq_bits;
    }
    dev_.total_subseq = plan_.total_subseq;
    // ---- groups of consecutive images, balanced by entropy-coded bytes; each group runs on its own stream ----
    // With host outputs the device->host copies (9x the compressed bytes) bound the call: the first group is small so that
    // the copy engine starts early, and every later group is twice as large so that it is ready before the copies of its
    // predecessors drain.  With device outputs the groups are equal.
    int G = opt_.profile ? 1 : std::max(1, std::min(opt_.groups, kMaxGroups));
    if (!opt_.profile && !host_outs_.empty() && opt_.groups_host > 0)
        G = std::min(opt_.groups_host, kMaxGroups);
    G = std::min(G, m);
    const bool growing = !host_outs_.empty() && G > 1;
    uint64_t weight_total = 0, weight_sum = 0;
    for (int g = 0; g < G; ++g)
        weight_total += growing ? (uint64_t(1) << g) + (g == G - 1 ? 1 : 0) : 1;
    int per_group = std::max(1, tail_capacity_ / G); // the groups' cooperative kernels together fill the device exactly
    int first = 0;
    for (int g = 0; g < G; ++g) {
        int last = first;
        weight_sum += growing ? (uint64_t(1) << g) : 1;
        if (g == G - 1) {
            last = m;
        } else {
            const uint64_t target = plan_.raw_bytes * weight_sum / weight_total;
            while (last < m - (G - 1 - g) && (last == first || plan_.images[last].raw_off + plan_.images[last].stage_len / 2 <= target))
                ++last;
        }
        Group gr;
        gr.first = first;
        gr.count = last - first;
        gr.tail_grid = per_group;
        gr.dev = dev_;
        gr.dev.images = dev_.images + first;
        gr.dev.states = dev_.states + first;
        gr.dev.num_images = gr.count;
        gr.dev.cta_map_unstuff = dev_.cta_map_unstuff + first;
        gr.dev.cta_map_subseq = dev_.cta_map_subseq + first;
        gr.dev.image_base = first;
        gr.dev.cta_map_idct = dev_.cta_map_idct + first;
        gr.dev.cta_map_color = dev_.cta_map_color + first;
        gr.dev.base_unstuff = plan_.map_unstuff[first];
        gr.dev.base_subseq = plan_.map_subseq[first];
        gr.dev.cta_map_fused = dev_.cta_map_fused + first;
        gr.dev.base_fused = plan_.map_fused[first];
        gr.dims.fused_ctas = plan_.map_fused[last] - plan_.map_fused[first];
        gr.dev.cta_map_dc = dev_.cta_map_dc + first;
        gr.dev.base_dc = plan_.map_dc[first];
        gr.dims.dc_ctas = plan_.map_dc[last] - plan_.map_dc[first];
        gr.dev.cta_map_scan = dev_.cta_map_scan + first;
        gr.dev.base_scan = plan_.map_scan[first];
        gr.dims.scan_ctas = plan_.map_scan[last] - plan_.map_scan[first];
        gr.dev.base_idct = plan_.map_idct[first];
        gr.dev.base_color = plan_.map_color[first];
        gr.dev.worklist[0] = dev_.worklist[0] + plan_.images[first].subseq_first;
        gr.dev.worklist[1] = dev_.worklist[1] + plan_.images[first].subseq_first;
        gr.dev.sync = dev_.sync + g;
        gr.dims.num_images = gr.count;
        gr.dims.unstuff_ctas = plan_.map_unstuff[last] - plan_.map_unstuff[first];
        gr.dims.subseq_ctas = plan_.map_subseq[last] - plan_.map_subseq[first];
        gr.dims.idct_ctas = plan_.map_idct[last] - plan_.map_idct[first];
        gr.dims.color_ctas = plan_.map_color[last] - plan_.map_color[first];
        gr.prog_first = uint32_t(std::lower_bound(plan_.prog_images.begin(), plan_.prog_images.end(), uint32_t(first)) - plan_.prog_images.begin());
        gr.prog_count = uint32_t(std::lower_bound(plan_.prog_images.begin(), plan_.prog_images.end(), uint32_t(last)) - plan_.prog_images.begin()) -
                        gr.prog_first;
        gr.raw_lo = plan_.images[first].raw_off;
        gr.raw_hi = last < m ? plan_.images[last].raw_off : plan_.raw_bytes;
        groups_.push_back(gr);
        first = last;
    }
    for (size_t i = 0; i < host_outs_.size(); ++i)
        for (Group& gr : groups_)
            if (host_outs_[i].image >= gr.first && host_outs_[i].image < gr.first + gr.count)
                gr.host_outs.push_back(int(i));
    ready_ = true;
    return B200JPEG_STATUS_SUCCESS;
}

// entropy-coded bytes of one group -> pinned staging -> device (asynchronous on the group's stream)
// ---- staging pool ----
// Copy into the pinned staging buffer with streaming stores: the bytes are read next by the GPU's DMA engine, not by a
// CPU, and an upload that has to pull freshly written lines out of the cores' caches runs at half the PCIe rate.
static void stage_copy(uint8_t* dst, const uint8_t* src, size_t len)
{
#if defined(__x86_64__) || defined(_M_X64)
    if ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0 && len >= 64) {
        size_t i = 0;
        for (; i + 64 <= len; i += 64) {
            const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i));
            const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 16));
            const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 32));
            const __m128i d = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 48));
            _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i), a);
            _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 16), b);
            _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 32), c);
            _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 48), d);
        }
        if (i < len)
            std::memcpy(dst + i, src + i, len - i);
        _mm_sfence();
        return;
    }
#endif
    std::memcpy(dst, src, len);
}

StagePool::StagePool(int workers)
{
    for (int i = 0; i < workers; ++i)
        th_.emplace_back([this] { worker(); });
}

StagePool::~StagePool()
{
    {
        std::lock_guard<std::mutex> lk(mu_);
        stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : th_)
        t.join();
}

bool StagePool::claim_one()
{
    const size_t n = items_->size();
    if (next_.load(std::memory_order_relaxed) >= n)
        return false;
    const size_t i = next_.fetch_add(1, std::memory_order_relaxed);
    if (i >= n)
        return false;
    const Item& it = (*items_)[i];
    stage_copy(it.dst, it.src, it.len);
    done_[i].store(1, std::memory_order_release);
    return true;
}

void StagePool::worker()
{
    uint64_t seen = 0;
    for (;;) {
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] { return stop_ || epoch_ != seen; });
            if (stop_)
                return;
            seen = epoch_;
        }
        while (claim_one()) {
        }
        acks_.fetch_add(1, std::memory_order_acq_rel); // every worker passes through every job, however late it wakes up
    }
}

void StagePool::quiesce()
{
    while (acks_.load(std::memory_order_acquire) != epoch_ * th_.size())
        std::this_thread::yield();
}

void StagePool::start(const std::vector<Item>* items)
{
    quiesce();
    if (items->size() > done_cap_) {
        done_cap_ = items->size() * 2;
        done_.reset(new std::atomic<uint8_t>[done_cap_]);
    }
    for (size_t i = 0; i < items->size(); ++i)
        done_[i].store(0, std::memory_order_relaxed);
    checked_ = 0;
    {
        std::lock_guard<std::mutex> lk(mu_);
        items_ = items;
        next_.store(0, std::memory_order_release);
        ++epoch_;
    }
    cv_.notify_all();
}

void StagePool::wait_prefix(size_t upto)
{
    for (;;) {
        while (checked_ < upto && done_[checked_].load(std::memory_order_acquire))
            ++checked_;
        if (checked_ >= upto)
            return;
        if (!claim_one())
            std::this_thread::yield();
    }
}

namespace {
struct StageGuard
{
    StagePool* pool;
    size_t n;
    ~StageGuard() { pool->wait_prefix(n); }
};
} // namespace

// The entropy-coded bytes of every planned image, cut into pieces of at most 256 KB, in group order.
void Decoder::stage_start()
{
    constexpr size_t kPiece = 256 << 10;
    uint8_t* hr = static_cast<uint8_t*>(h_raw_.p);
    if (!pool_) {
        // as many copy threads as the option allows and as this process may run on: a job with one process per GPU gives
        // every process its share of the host's cores (sched_setaffinity / taskset / cgroup), and eight processes that each
        // wake eight threads on 32 cores spend their time switching (e2e_device_out at 8 GPUs: 7.3 ms per batch and rank)
        int hw = int(std::thread::hardware_concurrency());
        cpu_set_t set;
        if (sched_getaffinity(0, sizeof(set), &set) == 0 && CPU_COUNT(&set) > 0)
            hw = CPU_COUNT(&set);
        pool_.reset(new StagePool(std::max(0, std::min(opt_.copy_threads, hw > 0 ? hw : 1) - 1)));
    }
    pool_->quiesce(); // no worker of the previous job looks at the item list any more
    stage_items_.clear();
    stage_group_end_.assign(groups_.size(), 0);
    for (size_t g = 0; g < groups_.size(); ++g) {
        const Group& gr = groups_[g];
        for (int i = gr.first; i < gr.first + gr.count; ++i) {
            const ImageDesc& im = plan_.images[i];
            const int bi = planned_to_batch_[i];
            const uint8_t* src = data_[bi] + parsed_[bi].scan_offset;
            for (size_t off = 0; off < im.stage_len; off += kPiece)
                stage_items_.push_back({src + off, hr + im.raw_off + off, std::min(kPiece, size_t(im.stage_len) - off)});
        }
        stage_group_end_[g] = stage_items_.size();
    }
    pool_->start(&stage_items_);
}

int Decoder::stage_group(int g)
{
    NvtxRange nvtx("b200jpeg stage_group");
    Group& gr = groups_[g];
    pool_->wait_prefix(stage_group_end_[g]);
    const size_t bytes = gr.raw_hi - gr.raw_lo;
    B2J_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(d_raw_.p) + gr.raw_lo, static_cast<uint8_t*>(h_raw_.p) + gr.raw_lo, bytes, cudaMemcpyHostToDevice,
        streams_[g]));
    return 0;
}

int Decoder::prepare(int n, const uint8_t* const* data, const size_t* sizes, const OutputDesc* outs, int32_t* statuses)
{
    const int rc = plan_batch(n, data, sizes, outs, statuses);
    if (rc)
        return rc;
    stage_start();
    StageGuard guard{pool_.get(), stage_items_.size()}; // the caller's bytes are read by other threads until every item is done
    for (size_t g = 0; g < groups_.size(); ++g)
        if (const int r = stage_group(int(g)))
            return r;
    for (size_t g = 0; g < groups_.size(); ++g)
        B2J_CUDA(cudaStreamSynchronize(streams_[g])); // the caller's bytes are not needed any more
    return B200JPEG_STATUS_SUCCESS;
}

// every kernel of one group on the group's stream, then its statuses (and host outputs) back to the host
int Decoder::launch_group(int g)
{
    NvtxRange nvtx("b200jpeg launch_group");
    Group& gr = groups_[g];
    cudaStream_t st = streams_[g];
    const int m0 = gr.first, mc = gr.count;
    const bool prof = opt_.profile && groups_.size() == 1;
    auto mark = [&](int i) {
        if (prof)
            cudaEventRecord(ev_stage_[i], st);
    };
    if (trace_)
        cudaEventRecord(ev_trace_[1 + 3 * g], st);
    if (g != 0)
        B2J_CUDA(cudaStreamWaitEvent(st, ev_meta_, 0)); // descriptors / tables went up on stream 0
    for (size_t u = 0; u < user_streams_.size(); ++u)
        B2J_CUDA(cudaStreamWaitEvent(st, ev_user_[u], 0)); // and whatever the caller had in flight on the output buffers
    mark(kStageUnstuff);
    B2J_CUDA(cudaMemsetAsync(static_cast<ImageState*>(d_states_.p) + m0, 0, sizeof(ImageState) * mc, st));
    B2J_CUDA(cudaMemsetAsync(gr.dev.sync, 0, sizeof(SyncCounters), st));
    B2J_CUDA(cudaMemsetAsync(gr.dev.dc_rec + 2 * size_t(gr.dev.base_dc), 0, 16 * size_t(gr.dims.dc_ctas), st)); // the chunks of an image hand each other the DC predictors
    B2J_CUDA(cudaMemsetAsync(gr.dev.scan_rec + gr.dev.base_scan, 0, 8 * size_t(gr.dims.scan_ctas), st));           // ... and the slot counts
    stat_launches_ += launch_unstuff(gr.dev, gr.dims, st);
    if (use_fused_) { // the whole entropy-decode stage in one launch; its CTAs hand each other one word per CTA, zero = not there yet
        B2J_CUDA(cudaMemsetAsync(gr.dev.fused_rec + gr.dev.base_fused, 0, 8 * size_t(gr.dims.fused_ctas), st));
        mark(kStageSync);
        stat_launches_ += launch_huff_fused(gr.dev, gr.dims, st);
        mark(kStageTail);
        mark(kStageWrite);
    } else {
        mark(kStageSync);
        stat_launches_ += launch_huff_sync(gr.dev, gr.dims, st);
        mark(kStageTail);
        const int ns = launch_huff_tail(gr.dev, gr.dims, st, gr.tail_grid, opt_.tail_launches);
        if (ns < 0) {
            log_message(kLogError, "cooperative launch failed: %s", cudaGetErrorString(cudaGetLastError()));
            return B200JPEG_STATUS_CUDA_ERROR;
        }
        stat_launches_ += ns;
        mark(kStageWrite);
        stat_launches_ += launch_huff_write(gr.dev, gr.dims, st);
    }
    mark(kStageDc);
    stat_launches_ += launch_dc(gr.dev, gr.dims, st);
    stat_launches_ += launch_progressive(gr.dev, gr.prog_first, gr.prog_count, st); // progressive images of the group: scans -> coefficients
    mark(kStageIdct);
    stat_launches_ += launch_idct(gr.dev, gr.dims, st);
    mark(kStageColor);
    stat_launches_ += launch_color(gr.dev, gr.dims, st);
    mark(kNumStages);
    stage_valid_ = prof;
    B2J_CUDA(cudaGetLastError());
    if (trace_)
        cudaEventRecord(ev_trace_[2 + 3 * g], st);
    const int m = int(plan_.images.size());
    B2J_CUDA(cudaMemcpyAsync(static_cast<ImageState*>(h_states_.p) + m0, static_cast<ImageState*>(d_states_.p) + m0, sizeof(ImageState) * mc,
        cudaMemcpyDeviceToHost, st));
    B2J_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(h_states_.p) + sizeof(ImageState) * m + sizeof(SyncCounters) * g, gr.dev.sync,
        sizeof(SyncCounters), cudaMemcpyDeviceToHost, st));
    // host-memory outputs: copy back.  Outputs that are contiguous on both sides (rows, and then whole images of a batch
    // tensor) travel as one linear copy: every copy costs a few microseconds of PCIe idle time.
    // (Spreading many separate copies over several streams was measured and changes nothing: a 0.5 MB copy into its own
    // pinned allocation takes about 21 us wherever it is queued.)
    for (size_t k = 0; k < gr.host_outs.size();) {
        const HostOutput& ho = host_outs_[gr.host_outs[k]];
        const uint8_t* src = static_cast<const uint8_t*>(d_out_.p) + ho.dev_off;
        if (size_t(ho.host_stride) != ho.dev_pitch || ho.row_bytes != ho.dev_pitch) { // rows with slack: never write the caller's slack bytes
            B2J_CUDA(cudaMemcpy2DAsync(ho.host, size_t(ho.host_stride), src, ho.dev_pitch, ho.row_bytes, ho.rows, cudaMemcpyDeviceToHost, st));
            ++k;
            continue;
        }
        size_t len = ho.dev_pitch * ho.rows;
        size_t k2 = k + 1;
        for (; k2 < gr.host_outs.size(); ++k2) {
            const HostOutput& nx = host_outs_[gr.host_outs[k2]];
            if (size_t(nx.host_stride) != nx.dev_pitch || nx.row_bytes != nx.dev_pitch || nx.dev_off != ho.dev_off + len ||
                static_cast<uint8_t*>(nx.host) != static_cast<uint8_t*>(ho.host) + len)
                break;
            len += nx.dev_pitch * nx.rows;
        }
        B2J_CUDA(cudaMemcpyAsync(ho.host, src, len, cudaMemcpyDeviceToHost, st));
        k = k2;
    }
    if (trace_)
        cudaEventRecord(ev_trace_[3 + 3 * g], st);
    // the outputs' streams continue after this group (reference idiom: extensions/nvjpeg/hw_decoder.cpp:649-655)
    B2J_CUDA(cudaEventRecord(ev_done_[g], st));
    for (cudaStream_t us : user_streams_)
        B2J_CUDA(cudaStreamWaitEvent(us, ev_done_[g], 0));
    return 0;
}

int Decoder::begin_run(cudaStream_t user_stream, int32_t* statuses)
{
    if (!ready_)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    B2J_CUDA(cudaSetDevice(device_));
    for (int i = 0; i < batch_size_; ++i)
        statuses[i] = pre_status_[i];
    stat_launches_ = 0;
    user_streams_ = out_streams_;
    if ((any_default_stream_ || user_streams_.empty()) && std::find(user_streams_.begin(), user_streams_.end(), user_stream) == user_streams_.end())
        user_streams_.push_back(user_stream);
    while (ev_user_.size() < user_streams_.size()) {
        cudaEvent_t e = nullptr;
        B2J_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ev_user_.push_back(e);
    }
    if (!groups_.empty())
        for (size_t u = 0; u < user_streams_.size(); ++u)
            B2J_CUDA(cudaEventRecord(ev_user_[u], user_streams_[u]));
    if (trace_)
        cudaEventRecord(ev_trace_[0], streams_[0]);
    return 0;
}

int Decoder::finish(int32_t* statuses)
{
    NvtxRange nvtx("b200jpeg finish (wait for the device)");
    in_flight_ = false;
    const int m = int(plan_.images.size());
    for (int i = 0; i < batch_size_; ++i)
        statuses[i] = pre_status_[i];
    for (size_t g = 0; g < groups_.size(); ++g)
        B2J_CUDA(cudaStreamSynchronize(streams_[g]));
    if (m == 0)
        return B200JPEG_STATUS_SUCCESS;
    const ImageState* hs = static_cast<const ImageState*>(h_states_.p);
    const SyncCounters* sc = reinterpret_cast<const SyncCounters*>(static_cast<const uint8_t*>(h_states_.p) + sizeof(ImageState) * m);
    stat_rounds_ = 0;
    stat_redone_ = 0;
    for (size_t g = 0; g < groups_.size(); ++g) {
        stat_rounds_ = std::max<uint64_t>(stat_rounds_, 2 + sc[g].rounds);
        stat_redone_ += sc[g].redone;
        if (trace_enabled()) {
            std::fprintf(stderr, "b200jpeg: group %zu work-list rounds (items @ us since the kernel's start):", g);
            for (uint32_t r = 0; r < sc[g].rounds && r < uint32_t(kTraceRounds); ++r)
                std::fprintf(stderr, " %u@%.1f", sc[g].round_items[r], sc[g].round_ns[r] * 1e-3);
            std::fprintf(stderr, "\n");
        }
    }
    for (int i = 0; i < m; ++i) {
        const int bi = planned_to_batch_[i];
        if (hs[i].error || hs[i].done_blocks != plan_.images[i].total_blocks) {
            statuses[bi] = B200JPEG_STATUS_CORRUPT_STREAM;
            if (trace_enabled())
                std::fprintf(stderr, "b200jpeg: image %d: error flags 0x%x, %u of %u blocks\n", bi, hs[i].error, hs[i].done_blocks,
                    plan_.images[i].total_blocks);
        }
    }
    return B200JPEG_STATUS_SUCCESS;
}

// device work only: the entropy-coded bytes are already in device memory (prepare)
int Decoder::run(cudaStream_t user_stream, int32_t* statuses)
{
    if (const int rc = begin_run(user_stream, statuses))
        return rc;
    for (size_t g = 0; g < groups_.size(); ++g)
        if (const int rc = launch_group(int(g)))
            return rc;
    return finish(statuses);
}

// whole decode, pipelined over the groups: while group g is on the GPU the host stages group g+1
int Decoder::submit(int n, const uint8_t* const* data, const size_t* sizes, const OutputDesc* outs, cudaStream_t user_stream, int32_t* statuses)
{
    if (in_flight_)
        return B200JPEG_STATUS_INVALID_PARAMETER; // the batch before was not waited for
    const bool trace = trace_enabled(); // developer aid: host-side timeline on stderr
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = trace ? now() : 0;
    if (trace && !ev_trace_[0])
        for (cudaEvent_t& e : ev_trace_)
            cudaEventCreate(&e);
    trace_ = trace;
    int rc = plan_batch(n, data, sizes, outs, statuses);
    if (rc)
        return rc;
    if ((rc = begin_run(user_stream, statuses)))
        return rc;
    in_flight_ = true;
    const double t1 = trace ? now() : 0;
    double t_stage = 0, t_launch = 0;
    stage_start();
    StageGuard guard{pool_.get(), stage_items_.size()};
    for (size_t g = 0; g < groups_.size(); ++g) {
        const double a = trace ? now() : 0;
        if ((rc = stage_group(int(g))))
            return rc;
        const double b = trace ? now() : 0;
        if ((rc = launch_group(int(g))))
            return rc;
        if (trace) {
            t_stage += b - a;
            t_launch += now() - b;
        }
    }
    if (trace)
        std::fprintf(stderr, "b200jpeg trace: plan %.3f ms, stage %.3f ms, launch %.3f ms (%zu groups)\n", t1 - t0, t_stage, t_launch,
            groups_.size());
    return B200JPEG_STATUS_SUCCESS;
}

int Decoder::wait(int32_t* statuses)
{
    if (!in_flight_)
        return B200JPEG_STATUS_INVALID_PARAMETER;
    cudaSetDevice(device_); // wait may run on a thread of its own
    const bool trace = trace_;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t2 = trace ? now() : 0;
    const int rc = finish(statuses);
    if (trace) {
        for (size_t g = 0; g < groups_.size(); ++g) {
            float a = 0, b = 0, c = 0;
            cudaEventElapsedTime(&a, ev_trace_[0], ev_trace_[1 + 3 * g]);
            cudaEventElapsedTime(&b, ev_trace_[0], ev_trace_[2 + 3 * g]);
            cudaEventElapsedTime(&c, ev_trace_[0], ev_trace_[3 + 3 * g]);
            std::fprintf(stderr, "b200jpeg trace: group %zu (%d images): input on device %.3f, kernels done %.3f, outputs on host %.3f ms\n", g,
                groups_[g].count, a, b, c);
        }
        std::fprintf(stderr, "b200jpeg trace: wait %.3f ms\n", now() - t2);
    }
    return rc;
}

int Decoder::decode(int n, const uint8_t* const* data, const size_t* sizes, const OutputDesc* outs, cudaStream_t user_stream, int32_t* statuses)
{
    const int rc = submit(n, data, sizes, outs, user_stream, statuses);
    if (rc) {
        if (in_flight_) { // something may be running: drain it before reporting the failure
            for (cudaStream_t st : streams_)
                cudaStreamSynchronize(st);
            in_flight_ = false;
        }
        return rc;
    }
    return wait(statuses);
}

void Decoder::get_stats(uint64_t* s) const
{
    s[0] = stat_launches_;
    s[1] = stat_rounds_;
    s[2] = 0;
    for (const ImageDesc& im : plan_.images)
        s[2] += im.stage_len;
    s[3] = plan_.total_blocks;
    s[4] = plan_.total_subseq;
    s[5] = scratch_bytes_;
    s[6] = stat_redone_;
    s[7] = plan_.out_bytes;
    s[8] = 0;
    for (int i = 9; i < 16; ++i)
        s[i] = 0;
}

int Decoder::get_stage_times(float* ms, int n)
{
    if (!opt_.profile || !stage_valid_)
        return -1;
    for (int i = 0; i < n && i < kNumStages; ++i)
        if (cudaEventElapsedTime(&ms[i], ev_stage_[i], ev_stage_[i + 1]) != cudaSuccess)
            return -1;
    return kNumStages;
}

int64_t Decoder::debug_fetch(int what, int image, int comp, void* dst, size_t cap)
{
    int planned = -1;
    for (size_t i = 0; i < planned_to_batch_.size(); ++i)
        if (planned_to_batch_[i] == image)
            planned = int(i);
    if (planned < 0 || !dst)
        return -1;
    const ImageDesc& im = plan_.images[planned];
    const void* src = nullptr;
    size_t bytes = 0;
    if (what == 0) { // coefficients: AC from the coefficient array, DC (after prediction) from the DC array
        if (!opt_.dump_coef && !im.progressive)
            return -4; // needs the decoder option dump_coef=1 (progressive streams keep their coefficients anyway)
        const size_t nb = im.total_blocks;
        bytes = nb * 128;
        if (bytes > cap)
            return -2;
        if (cudaSetDevice(device_) != cudaSuccess ||
            cudaMemcpy(dst, static_cast<int16_t*>(d_coef_.p) + im.coef_off, bytes, cudaMemcpyDeviceToHost) != cudaSuccess)
            return -3;
        std::vector<int16_t> dcv(nb);
        if (cudaMemcpy(dcv.data(), static_cast<int16_t*>(d_dc_.p) + im.coef_off / 64, nb * 2, cudaMemcpyDeviceToHost) != cudaSuccess)
            return -3;
        int16_t* o = static_cast<int16_t*>(dst);
        if (!im.progressive) // baseline: the DC values live in their own array
            for (size_t i = 0; i < nb; ++i)
                o[i * 64] = dcv[i];
        return int64_t(bytes);
    } else if (what == 1) {
        if (comp < 0 || comp >= im.ncomp)
            return -1;
        src = static_cast<uint8_t*>(d_planes_.p) + im.plane_off[comp];
        bytes = size_t(im.plane_pitch[comp]) * im.hib[comp] * 8;
    } else if (what == 3) { // developer builds (B2J_FUSED_TRACE): eight time stamps per CTA of k_huff_fused, all CTAs of the batch
        src = d_exit_.p;
        bytes = std::min(cap, size_t(plan_.map_fused.back()) * 64);
    } else if (what == 2) {
        src = static_cast<uint8_t*>(d_clean_.p) + im.clean_off;
        bytes = static_cast<ImageState*>(h_states_.p)[planned].clean_len;
    } else
        return -1;
    if (bytes > cap)
        return -2;
    if (cudaSetDevice(device_) != cudaSuccess || cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost) != cudaSuccess)
        return -3;
    return int64_t(bytes);
}

} // namespace b200jpeg
