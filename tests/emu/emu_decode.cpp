// tests/emu/emu_decode.cpp -- TEST INFRASTRUCTURE.
// CPU emulation of the b200jpeg decode pipeline: the *same* per-thread bodies the CUDA kernels run
// (nvimagecodec_b200/csrc/jpeg_math.cuh) and the same batch plan, executed by plain loops over "threads".  The
// block-cooperative kernels (unstuffing, scans, DC prediction) are replaced by serial statements of what they compute.
// Lets the "not gpu" tests check parser, planner, Huffman self-synchronisation, IDCT, upsampling and colour logic
// bit-for-bit against the oracle without a GPU.  Never linked into the product.
#include <cstdio>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <vector>
#include "../../nvimagecodec_b200/csrc/batch_plan.h"
#include "../../nvimagecodec_b200/csrc/huffman.cuh"
#include "../../nvimagecodec_b200/csrc/jpeg_math.cuh"
#include "../../nvimagecodec_b200/csrc/progressive.cuh"

using namespace b200jpeg;

extern "C" int emu_decode_batch(int n, const uint8_t* const* data, const size_t* sizes, uint8_t* const* outs, const int64_t* strides,
    const int* formats, int fancy, int subseq_bytes, int max_rounds, int jacobi, int* statuses, int* rounds_used,
    int16_t* const* coef_out /* optional per image, MCU order */, unsigned long long* stats /* optional [8] */,
    const int* rois /* optional: x, y, width, height per image; width 0 = whole image */, const int* orients /* optional: ImageDesc::orient per image */)
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
        if (rois && rois[4 * i + 2] > 0) {
            o.roi_x = rois[4 * i];
            o.roi_y = rois[4 * i + 1];
            o.roi_w = rois[4 * i + 2];
            o.roi_h = rois[4 * i + 3];
        }
        if (orients)
            o.orient = orients[i];
        plan_add_image(plan, parsed[i], o, opt);
        idx.push_back(i);
    }
    plan_finish(plan);
    const int m = int(plan.images.size());
    std::vector<uint8_t> raw(plan.raw_bytes, 0), clean(plan.clean_bytes, 0), planes(plan.plane_bytes, 0);
    std::vector<int16_t> coef(plan.coef_elems, 0x5A5A), dcv(plan.coef_elems / 64 + 64, 0x5A5A); // poisoned: the kernels must zero what they own
    std::vector<uint32_t> seg(plan.total_segs + 1, 0), rec(plan.total_subseq, 0), sabs(plan.total_subseq, 0);
    std::vector<unsigned long long> ex(plan.total_subseq, 0), en(plan.total_subseq, 0);
    std::vector<uint32_t> blkpos(plan.coef_elems / 64 + 64, 0xDEADBEEFu); // poisoned
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
    b.coef = coef.data();
    b.dc = dcv.data();
    b.planes = planes.data();
    b.subseq_bits = plan.subseq_bits;
    b.total_subseq = plan.total_subseq;
    const uint32_t S = uint32_t(plan.subseq_bits);

    auto make_env = [&](int i, bool shared_layout, std::vector<uint16_t>& s_lut) {
        const ImageDesc& im = plan.images[i];
        HuffEnv env;
        env.words = reinterpret_cast<const uint32_t*>(clean.data() + im.clean_off);
        if (shared_layout) { // what stage_tables() builds in shared memory
            s_lut.assign(size_t(kMaxTabSlots) * kLutEntries, 0);
            for (int s = 0; s < im.ntabs; ++s)
                std::memcpy(&s_lut[size_t(s) * kLutEntries], plan.huff[im.tab_pool[s]].lut, kLutEntries * 2);
            env.lut = s_lut.data();
            env.tabsel = im.tabsel_s;
        } else {
            env.lut = reinterpret_cast<const uint16_t*>(plan.huff.data());
            env.tabsel = im.tabsel_g;
        }
        env.pool = plan.huff.data();
        env.seg = seg.data() + im.seg_first;
        env.num_segments = im.num_segments;
        env.seg_blocks = im.seg_blocks;
        env.total_blocks = im.total_blocks;
        env.clean_bits = states[i].clean_len * 8u;
        env.bpm = im.bpm;
        return env;
    };
    // sync_decode of decode_kernels.cu
    std::vector<uint8_t> clean_sw; // filled after unstuffing
    auto sync_decode = [&](const HuffEnv& env, uint32_t t, uint32_t k, unsigned long long E, bool exact) {
        HuffState s;
        s.p = uint32_t(E);
        const uint32_t slot = uint32_t(E >> 32);
        s.z = slot & 63u;
        s.blk = slot >> 6;
        s.nblk = 0;
        s.abs = exact;
        const uint32_t z_entry = s.z;
        if (!clean_sw.empty()) { // the kernels' staged (byte-swapped) stream: exercises the fast synchronisation loop
            HuffEnv env2 = env;
            env2.words = reinterpret_cast<const uint32_t*>(clean_sw.data()) + (env.words - reinterpret_cast<const uint32_t*>(clean.data()));
            decode_span<false, true>(env2, s, (k + 1u) * S, nullptr);
        } else
            decode_span<false>(env, s, (k + 1u) * S, nullptr);
        const unsigned long long X = pack_state(s.p, s.blk * 64u + s.z);
        const bool changed = X != ex[t];
        ex[t] = X;
        en[t] = E;
        const uint32_t n = s.abs ? s.nblk * 64u + s.z : s.nblk * 64u + s.z - z_entry;
        rec[t] = (n & 0x7FFFFFFFu) | (s.abs ? 0x80000000u : 0u);
        return changed;
    };
    // ---- unstuffing + restart marker removal (what k_unstuff_count/scan/write compute) ----
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        const uint8_t* src = data[idx[i]] + parsed[idx[i]].scan_offset;
        std::memcpy(raw.data() + im.raw_off, src, im.stage_len);
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
            else if (cur == 0xFF && nx != 0 && p + 1 < im.raw_len) break; // any other marker ends the entropy-coded data
            if (!remove) dst[k++] = uint8_t(cur);
        }
        states[i].clean_len = k;
        states[i].marker_count = markers;
        S[0] = 0;
        for (int g = int(markers) + 1; g <= im.num_segments; ++g) S[g] = k;
        if (int(markers) != im.num_segments - 1) states[i].error |= kErrBadMarker;
    }

    // ---- synchronisation: round 0 (guessed state), round 1 (predecessor's exit), then work-list rounds ----
    clean_sw.assign(clean.size() + 64, 0);
    std::memcpy(clean_sw.data(), clean.data(), clean.size());
    for (size_t w = 0; w + 4 <= clean_sw.size(); w += 4) {
        uint32_t v;
        std::memcpy(&v, clean_sw.data() + w, 4);
        v = __builtin_bswap32(v);
        std::memcpy(clean_sw.data() + w, &v, 4);
    }
    std::vector<uint32_t> wl, wl_next;
    unsigned long long redone = 0, skipped = 0, chain_redo = 0;
    int rounds = 2;
    const bool fused = (jacobi & 0xFF) == 2;
    const uint32_t H = (jacobi >> 8) ? uint32_t(jacobi >> 8) : uint32_t(kFusedHalo), OWN = uint32_t(kSubseqCta) - H;
    if (!fused) {
    for (int round = 0; round < 2; ++round) {
        const std::vector<unsigned long long> snap = ex; // worst-case race: every thread sees its predecessor's old exit
        for (int i = 0; i < m; ++i) {
            const ImageDesc& im = plan.images[i];
            std::vector<uint16_t> s_lut;
            const HuffEnv env = make_env(i, true, s_lut);
            const uint32_t clean_bits = states[i].clean_len * 8u;
            for (uint32_t k = 0; k < im.subseq_cta_count * kSubseqCta; ++k) {
                const uint32_t t = im.subseq_first + k;
                const bool idle = uint64_t(k) * S >= uint64_t(clean_bits);
                if (round == 0) {
                    if (idle) {
                        ex[t] = en[t] = pack_state(clean_bits, 0);
                        rec[t] = 0;
                        continue;
                    }
                    ex[t] = ~0ull;
                    sync_decode(env, t, k, pack_state(k * S, 0), k == 0);
                } else {
                    if (idle || k == 0)
                        continue;
                    const unsigned long long E = snap[t - 1];
                    if (E == en[t])
                        continue;
                    const bool changed = sync_decode(env, t, k, E, false);
                    if (changed && uint64_t(k + 1) * S < uint64_t(clean_bits))
                        wl.push_back(t + 1);
                }
            }
        }
    }
    // work-list rounds (tail_round of huffman_kernels.cu): Jacobi until the list is empty; from round kRunAheadFrom on one
    // thread per run of listed items walks the run and on into what follows, periodic stretches without decoding
    std::vector<uint32_t> stamp(plan.total_subseq, 0);
    for (uint32_t t : wl)
        stamp[t] = 1;
    rounds = 2;
    while (!wl.empty() && rounds < max_rounds) {
        const int r = rounds - 2;
        const bool run_ahead = r >= kRunAheadFrom;
        const std::vector<unsigned long long> snap = ex;
        wl_next.clear();
        for (uint32_t t0 : wl) {
            uint32_t t = t0;
            if (run_ahead && stamp[t - 1] >= uint32_t(r) + 1u)
                continue; // not the head of its run of listed items
            int i = 0;
            while (i + 1 < m && plan.images[i + 1].subseq_first <= t)
                ++i;
            const ImageDesc& im = plan.images[i];
            std::vector<uint16_t> unused;
            const HuffEnv env = make_env(i, false, unused); // the tail kernel reads the tables from the global pool
            const uint32_t* gwords = reinterpret_cast<const uint32_t*>(clean.data() + im.clean_off);
            const uint32_t spw = S >> 5;
            const uint64_t clean_bits = uint64_t(states[i].clean_len) * 8u;
            unsigned long long E = snap[t - 1];
            bool prev_shift = false, in_run = true;
            for (;;) {
                const uint32_t k = t - im.subseq_first;
                bool changed = false;
                unsigned long long X;
                if (E == en[t]) {
                    X = ex[t];
                } else if (prev_shift && im.num_segments == 1 && uint64_t(k + 1) * S + 64u <= clean_bits && subseq_repeats(gwords, (k - 1u) * spw, spw)) {
                    X = E + S;
                    changed = X != ex[t];
                    ex[t] = X;
                    en[t] = E;
                    rec[t] = rec[t - 1];
                    ++skipped;
                    ++redone;
                } else {
                    changed = sync_decode(env, t, k, E, false);
                    X = ex[t];
                    ++redone;
                }
                if (uint64_t(k + 1) * S >= clean_bits)
                    break;
                if (!run_ahead) {
                    if (changed) {
                        wl_next.push_back(t + 1);
                        stamp[t + 1] = uint32_t(r) + 2u;
                    }
                    break;
                }
                const bool next_listed = stamp[t + 1] >= uint32_t(r) + 1u;
                if (next_listed && !in_run) {
                    if (changed) {
                        wl_next.push_back(t + 1);
                        stamp[t + 1] = uint32_t(r) + 2u;
                    }
                    break;
                }
                if (!next_listed) {
                    if (!changed)
                        break;
                    in_run = false;
                }
                prev_shift = (X == E + S);
                E = X;
                ++t;
            }
        }
        wl.swap(wl_next);
        ++rounds;
    }
    // ---- segmented scan of slot counts + consistency check (k_subseq_scan) ----
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        uint32_t A = 0x80000000u;
        for (uint32_t k = 0; k < im.subseq_cta_count * kSubseqCta; ++k) {
            const uint32_t t = im.subseq_first + k;
            if (k > 0 && en[t] != ex[t - 1] && uint64_t(k) * S < uint64_t(states[i].clean_len) * 8u)
                states[i].error |= kErrNotConverged;
            const uint32_t r = rec[t];
            A = (r & 0x80000000u) ? r : A + r;
            sabs[t] = A;
        }
    }
    }
    // ---- block pass (k_huff_write): per block the DC difference and the position of its first AC symbol; the
    // subsequences in reverse order to show that the threads do not depend on each other ----
    std::vector<uint8_t> clean_sw2 = clean_sw;
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        std::vector<uint16_t> s_lut;
        HuffEnv env = make_env(i, true, s_lut);
        env.words = reinterpret_cast<const uint32_t*>(clean_sw2.data()) + (env.words - reinterpret_cast<const uint32_t*>(clean.data()));
        const uint32_t clean_bits = states[i].clean_len * 8u;
        if (fused) {
            // k_huff_fused: per CTA (kFusedOwn own subsequences + halo) rounds in "shared memory", validation against the
            // predecessor's published state, segmented scan, block pass
            BlkOut out;
            out.dc = dcv.data() + im.coef_off / 64;
            out.pos = blkpos.data() + im.coef_off / 64;
            const uint32_t live = uint32_t((uint64_t(clean_bits) + S - 1u) / S);
            const uint32_t spw = S >> 5;
            const uint32_t* bwords = env.words; // byte-swapped stream of the image
            uint32_t pred_state = 0, pred_abs = 0x80000000u;
            const uint32_t nctas = (im.subseq_count + OWN - 1) / OWN;
            for (uint32_t c = 0; c < nctas; ++c) {
                const uint32_t h = c ? H : 0u, k0 = c * OWN - h;
                const uint32_t nv = live > k0 ? std::min(OWN + h, live - k0) : 0u;
                if (nv <= h)
                    break;
                std::vector<uint32_t> xs(kSubseqCta, 0xFFFFFFFFu), rc(kSubseqCta, 0);
                std::vector<uint32_t> list, next;
                bool validated = c == 0;
                for (uint32_t round = 0;; ++round) {
                    std::vector<int> items;
                    if (round == 0) for (uint32_t i2 = 0; i2 < nv; ++i2) items.push_back(int(i2));
                    else if (round == 1) for (uint32_t i2 = 1; i2 < nv; ++i2) items.push_back(int(i2));
                    else for (uint32_t v : list) items.push_back(int(v));
                    std::vector<uint32_t> entries(items.size(), 0);
                    std::vector<uint8_t> start(kSubseqCta + 1, 0);
                    for (size_t q = 0; q < items.size(); ++q) {
                        if (items[q] > 0) entries[q] = xs[items[q] - 1];
                        if (round >= 2) start[items[q]] = 1;
                    }
                    next.clear();
                    const bool run_heads = round >= uint32_t(kFusedRunHeads);
                    for (size_t q = 0; q < items.size(); ++q) {
                        int li = items[q];
                        if (run_heads && li > 0 && start[li - 1])
                            continue;
                        uint32_t E = entries[q];
                        bool prev_shift = false, in_run = true;
                        for (;;) {
                            const uint32_t kk = k0 + uint32_t(li);
                            uint32_t X;
                            if (prev_shift && im.num_segments == 1 && uint32_t(li) + 2u < uint32_t(kSubseqCta) && uint64_t(kk + 1) * S + 64u <= uint64_t(clean_bits)) {
                                bool same = true;
                                for (uint32_t j = 0; j < spw + 2u && same; ++j)
                                    same = bwords[(kk - 1u) * spw + j] == bwords[kk * spw + j];
                                prev_shift = same;
                            } else
                                prev_shift = false;
                            bool changed;
                            if (prev_shift) {
                                X = E;
                                changed = X != xs[li];
                                xs[li] = X;
                                rc[li] = rc[li - 1];
                                ++skipped;
                            } else {
                                HuffState s;
                                if (round == 0) { s.p = kk * S; s.z = 0; s.blk = 0; s.abs = kk == 0; }
                                else { s.p = rel_pos(E, kk - 1u, S); s.z = (E >> 16) & 63u; s.blk = E >> 22; s.abs = false; }
                                s.nblk = 0;
                                const uint32_t z_entry = s.z;
                                decode_span<false, true>(env, s, (kk + 1u) * S, nullptr);
                                X = pack_rel(s.p, kk, S, s.blk * 64u + s.z);
                                changed = X != xs[li];
                                xs[li] = X;
                                const uint32_t n = s.abs ? s.nblk * 64u + s.z : s.nblk * 64u + s.z - z_entry;
                                rc[li] = (n & 0x7FFFFFFFu) | (s.abs ? 0x80000000u : 0u);
                            }
                            if (round == 0) break;
                            if (round >= 2) ++redone;
                            if (uint32_t(li) + 1u >= nv) break;
                            const bool next_listed = round >= 2 && start[li + 1];
                            bool push = false, stop;
                            if (round == 1) { push = changed; stop = true; }
                            else if (!run_heads) { push = changed && next_listed; stop = !changed || next_listed; }
                            else if (next_listed && !in_run) { push = changed; stop = true; }
                            else { stop = !next_listed && !changed; if (!next_listed) in_run = false; }
                            if (push) next.push_back(uint32_t(li + 1));
                            if (stop) break;
                            prev_shift = (X == E);
                            E = X;
                            ++li;
                        }
                    }
                    list.swap(next);
                    if (round >= 1 && list.empty()) {
                        rounds = std::max(rounds, int(round) + 1);
                        if (validated) break;
                        validated = true;
                        if (pred_state == xs[h - 1]) break;
                        xs[h - 1] = pred_state;
                        list.assign(1, h);
                        ++chain_redo;
                    }
                }
                uint32_t A = pred_abs;
                std::vector<uint32_t> exc(kSubseqCta, 0);
                for (uint32_t i2 = h; i2 < nv; ++i2) {
                    exc[i2] = A;
                    A = (rc[i2] & 0x80000000u) ? rc[i2] : A + rc[i2];
                }
                pred_state = xs[nv - 1];
                pred_abs = A | 0x80000000u;
                for (int64_t i64 = int64_t(nv) - 1; i64 >= int64_t(h); --i64) { // reverse order: the threads do not depend on each other
                    const uint32_t i2 = uint32_t(i64), k = k0 + i2;
                    HuffState s;
                    if (k == 0) { s.p = 0; s.z = 0; s.blk = 0; s.nblk = 0; }
                    else {
                        const uint32_t E = xs[i2 - 1];
                        s.p = rel_pos(E, k - 1u, S); s.z = (E >> 16) & 63u; s.blk = E >> 22;
                        s.nblk = (exc[i2] & 0x7FFFFFFFu) >> 6;
                    }
                    s.abs = true;
                    const uint32_t nblk_in = s.nblk;
                    uint32_t err = decode_span<true, true>(env, s, (k + 1u) * S, &out);
                    if (s.nblk == im.total_blocks && s.z == 0 && nblk_in < im.total_blocks) states[i].done_blocks = s.nblk;
                    states[i].error |= err;
                }
            }
        } else
        for (int64_t k64 = int64_t(im.subseq_cta_count) * kSubseqCta - 1; k64 >= 0; --k64) {
            const uint32_t k = uint32_t(k64), t = im.subseq_first + k;
            if (uint64_t(k) * S >= uint64_t(clean_bits)) continue;
            HuffState s;
            if (k == 0) {
                s.p = 0; s.z = 0; s.blk = 0; s.nblk = 0;
            } else {
                const unsigned long long E = ex[t - 1];
                const uint32_t slot = uint32_t(E >> 32);
                s.p = uint32_t(E); s.z = slot & 63u; s.blk = slot >> 6;
                s.nblk = (sabs[t - 1] & 0x7FFFFFFFu) >> 6;
            }
            s.abs = true;
            BlkOut out;
            out.dc = dcv.data() + im.coef_off / 64;
            out.pos = blkpos.data() + im.coef_off / 64;
            const uint32_t nblk_in = s.nblk;
            uint32_t err = decode_span<true, true>(env, s, (k + 1u) * S, &out);
            if (s.nblk == im.total_blocks && s.z == 0 && nblk_in < im.total_blocks) states[i].done_blocks = s.nblk;
            if (err && std::getenv("EMU_DEBUG")) std::fprintf(stderr, "emu: block pass err %u at subseq %u (p0 %u) nblk_in %u -> %u z %u\n", err, k, k * S, nblk_in, s.nblk, s.z);
            states[i].error |= err;
        }
        // dense coefficients: what every IDCT thread decodes into its shared-memory slot (also the debug tap)
        if (states[i].done_blocks == im.total_blocks) {
            std::vector<uint16_t> lut(size_t(3) * kLutEntries, 0);
            for (int sidx = 0; sidx < im.nac; ++sidx)
                std::memcpy(&lut[size_t(sidx) * kLutEntries], plan.huff[im.ac_pool[sidx]].lut, kLutEntries * 2);
            uint8_t zzd[kZigzagDump];
            for (int k = 0; k < kZigzagDump; ++k) zzd[k] = zigzag_dump_entry(h_zigzag, k);
            for (uint32_t bb = 0; bb < im.total_blocks; ++bb) {
                int16_t blk[kBlkSlots] = {};
                const int c = im.blk_comp[bb % uint32_t(im.bpm)];
                const uint32_t p = blkpos[im.coef_off / 64 + bb];
                uint32_t err = kErrOverrun;
                if (p < clean_bits)
                    err = decode_block_ac<false>(reinterpret_cast<const uint32_t*>(clean.data() + im.clean_off), (im.raw_len + 12u) >> 2, p, lut.data(),
                        uint32_t(im.ac_slot[c]) * kLutEntries, plan.huff.data() + im.ac_pool[im.ac_slot[c]], zzd, blk);
                if (err && std::getenv("EMU_DEBUG") && !(states[i].error & err)) std::fprintf(stderr, "emu: block %u decode err %u (pos %u)\n", bb, err, p);
                states[i].error |= err;
                std::memcpy(coef.data() + im.coef_off + uint64_t(bb) * 64, blk, 128);
            }
        }
    }
    if (rounds_used) *rounds_used = rounds;
    if (stats) {
        stats[0] = uint64_t(rounds);
        stats[1] = redone;
        stats[2] = plan.total_subseq;
        stats[3] = skipped;
        stats[4] = chain_redo;
    }

    // ---- progressive images (k_prog_decode): the scans one after the other into the dense coefficients ----
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        if (!im.progressive)
            continue;
        std::memset(coef.data() + im.coef_off, 0, size_t(im.total_blocks) * 128);
        uint32_t err = 0;
        // k_prog_decode runs scans side by side when ScanDesc::deps allows: emulate the most adverse legal order (always the
        // LAST scan of the file whose dependencies are complete), so that a missing dependency shows up as wrong coefficients
        std::vector<uint8_t> done(im.scan_count, 0);
        for (uint32_t step = 0; step < im.scan_count && !err; ++step) {
            int pick = -1;
            for (int sidx = int(im.scan_count) - 1; sidx >= 0 && pick < 0; --sidx) {
                if (done[sidx])
                    continue;
                const ScanDesc& sc = plan.scans[im.scan_first + sidx];
                bool ready = true;
                for (int e = 0; e < sidx; ++e)
                    if ((im.scan_count > 32 || ((sc.deps >> e) & 1u)) && !done[e])
                        ready = false;
                if (ready)
                    pick = sidx;
            }
            const ScanDesc& sc = plan.scans[im.scan_first + pick];
            // look-ahead tables as the kernel builds them (prog_fast_entry)
            std::vector<uint16_t> fast(size_t(3) << kProgFastBits, 0);
            const int ntab = sc.ss != 0 ? 1 : (sc.ah == 0 ? sc.ncomp : 0);
            for (int t = 0; t < ntab; ++t)
                for (uint32_t e = 0; e < (1u << kProgFastBits); ++e)
                    fast[(size_t(t) << kProgFastBits) + e] = prog_fast_entry(plan.huff.data() + (sc.ss != 0 ? sc.ac_tab : sc.dc_tab[t]), e);
            err = prog_decode_scan(im, sc, raw.data() + im.raw_off, plan.huff.data(), coef.data() + im.coef_off, h_zigzag, ntab ? fast.data() : nullptr);
            done[pick] = 1;
        }
        states[i].error |= err; // (the baseline passes saw an empty stream and must not have flagged anything)
        if (!err)
            states[i].done_blocks = im.total_blocks;
    }
    // ---- DC prediction (k_dc_scan) ----
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        if (im.progressive) { // the DC values are in the coefficients already
            for (uint32_t bb = 0; bb < im.total_blocks; ++bb)
                dcv[im.coef_off / 64 + bb] = coef[im.coef_off + uint64_t(bb) * 64];
            if (coef_out && coef_out[idx[i]])
                std::memcpy(coef_out[idx[i]], coef.data() + im.coef_off, size_t(im.total_blocks) * 128);
            continue;
        }
        int pred[3] = {0, 0, 0};
        const uint32_t mcus = uint32_t(im.mcus_x) * im.mcus_y;
        int16_t* dcp = dcv.data() + im.coef_off / 64;
        for (uint32_t mc = 0; mc < mcus; ++mc) {
            if (im.restart_interval && mc % im.restart_interval == 0) pred[0] = pred[1] = pred[2] = 0;
            for (int j = 0; j < im.bpm; ++j) {
                int16_t* d = dcp + uint64_t(mc) * im.bpm + j;
                pred[im.blk_comp[j]] += *d;
                *d = int16_t(pred[im.blk_comp[j]]);
            }
        }
        if (coef_out && coef_out[idx[i]]) {
            int16_t* o = coef_out[idx[i]];
            std::memcpy(o, coef.data() + im.coef_off, size_t(im.total_blocks) * 128);
            for (uint32_t bb = 0; bb < im.total_blocks; ++bb) o[uint64_t(bb) * 64] = dcp[bb];
        }
    }
    // ---- IDCT + colour ----
    for (int i = 0; i < m; ++i) {
        const ImageDesc& im = plan.images[i];
        for (uint32_t u = 0; u < im.idct_cta_count * kIdctCta; ++u) {
            const IdctBlockRef ref = idct_locate(im, u);
            if (ref.comp < 0) continue;
            uint32_t out[16];
            idct_block(coef.data() + im.coef_off + uint64_t(ref.src_blk) * 64, dcv[im.coef_off / 64 + ref.src_blk],
                plan.quant.data() + im.qt[ref.comp] * 64, out);
            for (int r = 0; r < 8; ++r) std::memcpy(planes.data() + ref.dst + uint64_t(r) * ref.pitch, &out[2 * r], 8);
        }
        const int hx = color_hx(im), vy = color_vy(im);
        const uint32_t tiles = uint32_t(im.color_rows) * uint32_t(im.color_xtiles);
        for (uint32_t tile = 0; tile < tiles; ++tile) {
            const int q = int(tile / uint32_t(im.color_xtiles)) - (vy == 2 ? 1 : 0);
            const int xbase = int(tile % uint32_t(im.color_xtiles)) * kColorTilePx;
            for (int lane = 0; lane < 32; ++lane) {
                const int x0 = xbase + lane * 16;
                if (x0 >= im.width) continue;
                uint32_t px[32];
                int yrow[2] = {-1, -1};
                if (vy == 2 && hx == 4) { color_unit<4, 2>(b, im, q, x0, px, yrow); store_unit_simple<2>(im, x0, px, yrow); }
                else if (vy == 2 && hx == 2) { color_unit<2, 2>(b, im, q, x0, px, yrow); store_unit_simple<2>(im, x0, px, yrow); }
                else if (vy == 2) { color_unit<1, 2>(b, im, q, x0, px, yrow); store_unit_simple<2>(im, x0, px, yrow); }
                else if (hx == 4) { color_unit<4, 1>(b, im, q, x0, px, yrow); store_unit_simple<1>(im, x0, px, yrow); }
                else if (hx == 2) { color_unit<2, 1>(b, im, q, x0, px, yrow); store_unit_simple<1>(im, x0, px, yrow); }
                else { color_unit<1, 1>(b, im, q, x0, px, yrow); store_unit_simple<1>(im, x0, px, yrow); }
            }
        }
        if (std::getenv("EMU_DEBUG"))
            std::fprintf(stderr, "emu: image %d error flags %u, done_blocks %u of %u\n", i, states[i].error, states[i].done_blocks, im.total_blocks);
        if (states[i].error) statuses[idx[i]] = 4;
        else if (states[i].done_blocks != im.total_blocks) statuses[idx[i]] = 4;
    }
    return 0;
}
