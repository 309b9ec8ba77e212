// tests/emu/emu_decode.cpp -- TEST INFRASTRUCTURE.
// CPU emulation of the b200jpeg decode pipeline: the *same* per-thread bodies the CUDA kernels run
// (nvimagecodec_b200/csrc/jpeg_math.cuh) and the same batch plan, executed by plain loops over "threads".  The
// block-cooperative kernels (unstuffing, scans, DC prediction) are replaced by serial statements of what they compute.
// Lets the "not gpu" tests check parser, planner, Huffman self-synchronisation, IDCT, upsampling and colour logic
// bit-for-bit against the oracle without a GPU.  Never linked into the product.
#include <cstdio>
#include <cstring>
#include <vector>
#include "../../nvimagecodec_b200/csrc/batch_plan.h"
#include "../../nvimagecodec_b200/csrc/jpeg_math.cuh"

using namespace b200jpeg;

extern "C" int emu_decode_batch(int n, const uint8_t* const* data, const size_t* sizes, uint8_t* const* outs, const int64_t* strides,
    const int* formats, int fancy, int subseq_bytes, int max_rounds, int jacobi, int* statuses, int* rounds_used,
    int16_t* const* coef_out /* optional per image, MCU order */)
{
    PlanOptions opt;
    opt.subseq_bytes = subseq_bytes;
    opt.fancy_upsampling = fancy != 0;
    BatchPlan plan;
    plan_begin(plan, opt);
    std::vector<ParsedJpeg> parsed(n);
    std::vector<int> idx;
    for (int i = 0; i < n; ++i) {
        statuses[i] = 0;
        ParseResult r = parse_jpeg(data[i], sizes[i], parsed[i]);
        if (r != PARSE_OK) {
            statuses[i] = 2;
            continue;
        }
        if (!is_supported(parsed[i]) || !resolve_format(formats[i], parsed[i].ncomp)) {
            statuses[i] = 3;
            continue;
        }
        OutputSpec o;
        o.buffer = outs[i];
        o.row_stride = strides[i];
        o.format = formats[i];
        plan_add_image(plan, parsed[i], o, opt);
        idx.push_back(i);
    }
    plan_finish(plan);
    const int m = int(plan.images.size());
    std::vector<uint8_t> raw(plan.raw_bytes, 0), clean(plan.clean_bytes, 0), planes(plan.plane_bytes, 0);
    std::vector<int16_t> coef(plan.coef_elems, 0);
    std::vector<uint32_t> seg(plan.total_segs + 1, 0), rec(plan.total_subseq, 0), sabs(plan.total_subseq, 0), redo(64, 0);
    std::vector<unsigned long long> ex(plan.total_subseq, 0), en(plan.total_subseq, 0);
    std::vector<ImageState> states(m);
    std::memset(states.data(), 0, sizeof(ImageState) * m);
    BatchDev b;
    std::memset(&b, 0, sizeof(b));
    b.images = plan.images.data();
    b.states = states.data();
    b.num_images = m;
    b.raw = raw.data();
    b.clean = clean.data();
    b.huff = plan.huff.data();
    b.quant = plan.quant.data();
    b.seg_start = seg.data();
    b.sub_exit = ex.data();
    b.sub_entry = en.data();
    b.sub_rec = rec.data();
    b.sub_abs = sabs.data();
    b.redo_counter = redo.data();
    b.coef = coef.data();
    b.planes = planes.data();
    b.subseq_bits = plan.subseq_bits;

    // ---- unstuffing + restart marker removal (what k_unstuff_count/scan/write compute) ----
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        const uint8_t* src = data[idx[i]] + parsed[idx[i]].scan_offset;
        std::memcpy(raw.data() + im.raw_off, src, im.raw_len);
        uint8_t* dst = clean.data() + im.clean_off;
        uint32_t k = 0, markers = 0;
        uint32_t* S = seg.data() + im.seg_first;
        for (uint32_t p = 0; p < im.raw_len; ++p) {
            const int cur = src[p], pv = p ? src[p - 1] : 0, nx = p + 1 < im.raw_len ? src[p + 1] : 0;
            bool remove = false;
            if (cur == 0x00 && pv == 0xFF) remove = true;
            else if (cur == 0xFF && nx == 0xFF) remove = true;
            else if (cur == 0xFF && nx >= 0xD0 && nx <= 0xD7) {
                remove = true;
                if (int(markers) + 1 < im.num_segments) {
                    S[markers + 1] = k;
                    if (nx != 0xD0 + int(markers & 7)) states[i].error |= kErrBadMarker;
                } else
                    states[i].error |= kErrBadMarker;
                ++markers;
            } else if (cur >= 0xD0 && cur <= 0xD7 && pv == 0xFF) remove = true;
            else if (cur == 0xFF && nx != 0 && p + 1 < im.raw_len) states[i].error |= kErrBadMarker;
            if (!remove) dst[k++] = uint8_t(cur);
        }
        states[i].clean_len = k;
        states[i].marker_count = markers;
        S[0] = 0;
        for (int g = int(markers) + 1; g <= im.num_segments; ++g) S[g] = k;
        if (int(markers) != im.num_segments - 1) states[i].error |= kErrBadMarker;
    }

    // ---- synchronisation rounds ----
    int rounds = 0;
    for (int round = 0; round < max_rounds; ++round) {
        std::vector<unsigned long long> snap;
        if (jacobi) snap = ex;
        uint32_t redone = 0;
        for (int i = 0; i < m; ++i) {
            const ImageDesc& im = plan.images[i];
            const uint32_t clean_len = states[i].clean_len;
            for (uint32_t k = 0; k < im.subseq_cta_count * kSubseqCta; ++k) {
                const uint32_t t = im.subseq_first + k;
                const bool idle = uint64_t(k) * plan.subseq_bits >= uint64_t(clean_len) * 8;
                SubseqState s;
                unsigned long long E;
                if (round == 0) {
                    if (idle) {
                        ex[t] = en[t] = (unsigned long long)clean_len * 8;
                        rec[t] = 0;
                        continue;
                    }
                    s.p = k * plan.subseq_bits;
                    s.slot = 0;
                    s.n = 0;
                    s.abs = (k == 0);
                    E = s.p;
                } else {
                    if (idle || k == 0) continue;
                    E = jacobi ? snap[t - 1] : ex[t - 1];
                    if (E == en[t]) continue;
                    s.p = uint32_t(E);
                    s.slot = uint32_t(E >> 32);
                    s.n = 0;
                    s.abs = false;
                    ++redone;
                }
                decode_subsequence<false>(b, im, &states[i], k, clean_len, s);
                ex[t] = (unsigned long long)s.p | ((unsigned long long)s.slot << 32);
                en[t] = E;
                rec[t] = (s.n & 0x7FFFFFFFu) | (s.abs ? 0x80000000u : 0);
            }
        }
        rounds = round + 1;
        if (round > 0 && redone == 0) break;
    }
    if (rounds_used) *rounds_used = rounds;

    // ---- segmented scan of slot counts (k_subseq_scan) ----
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        uint32_t A = 0;
        for (uint32_t k = 0; k < im.subseq_cta_count * kSubseqCta; ++k) {
            const uint32_t r = rec[im.subseq_first + k];
            A = (r & 0x80000000u) ? (r & 0x7FFFFFFFu) : A + r;
            sabs[im.subseq_first + k] = A | 0x80000000u;
        }
    }
    // ---- write pass (k_huff_write) ----
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        const uint32_t clean_len = states[i].clean_len;
        for (uint32_t k = 0; k < im.subseq_cta_count * kSubseqCta; ++k) {
            const uint32_t t = im.subseq_first + k;
            if (uint64_t(k) * plan.subseq_bits >= uint64_t(clean_len) * 8) continue;
            SubseqState s;
            if (k == 0) {
                s.p = 0; s.slot = 0; s.n = 0;
            } else {
                const unsigned long long E = ex[t - 1];
                if (E != en[t]) states[i].error |= kErrNotConverged;
                s.p = uint32_t(E); s.slot = uint32_t(E >> 32); s.n = sabs[t - 1] & 0x7FFFFFFFu;
            }
            s.abs = true;
            decode_subsequence<true>(b, im, &states[i], k, clean_len, s);
        }
    }
    // ---- DC prediction (k_dc_scan) ----
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        int pred[3] = {0, 0, 0};
        const uint32_t mcus = uint32_t(im.mcus_x) * im.mcus_y;
        for (uint32_t mc = 0; mc < mcus; ++mc) {
            if (im.restart_interval && mc % im.restart_interval == 0) pred[0] = pred[1] = pred[2] = 0;
            for (int j = 0; j < im.bpm; ++j) {
                int16_t* d = coef.data() + im.coef_off + (uint64_t(mc) * im.bpm + j) * 64;
                pred[im.blk_comp[j]] += *d;
                *d = int16_t(pred[im.blk_comp[j]]);
            }
        }
        if (coef_out && coef_out[idx[i]]) std::memcpy(coef_out[idx[i]], coef.data() + im.coef_off, size_t(im.total_slots) * 2);
    }
    // ---- IDCT + colour ----
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        for (uint32_t u = 0; u < im.idct_cta_count * kIdctCta; ++u) idct_thread(b, im, u);
        for (uint32_t u = 0; u < im.color_cta_count * kColorCta; ++u) color_thread(b, im, u);
        if (states[i].error) statuses[idx[i]] = 4;
        else if (states[i].done_slots != im.total_slots) statuses[idx[i]] = 4;
    }
    return 0;
}
