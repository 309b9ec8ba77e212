// batch_plan.cpp -- see batch_plan.h
#include "batch_plan.h"
#include <algorithm>
#include <cstring>

namespace b200jpeg {

namespace {
inline uint64_t align_up(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }
inline uint32_t cdiv(uint64_t a, uint64_t b) { return uint32_t((a + b - 1) / b); }
} // namespace

// canonical code assignment as in jpeg_make_d_derived_tbl (external/libjpeg-turbo/jdhuff.c:143-285); the first-level
// table holds pre-digested entries (device_types.h huff_entry)
void build_dev_table(const HuffSpec& spec, bool is_dc, DevHuffTable& t)
{
    std::memset(&t, 0, sizeof(t));
    std::memcpy(t.vals, spec.vals, 256);
    t.is_dc = is_dc ? 1 : 0;
    int code = 0, p = 0;
    for (int l = 1; l <= 16; ++l) {
        const int n = spec.bits[l];
        if (n) {
            t.valoffset[l] = p - code;
            for (int i = 0; i < n; ++i, ++p, ++code) {
                if (l <= kLutBits && code < (1 << l)) {
                    const int first = code << (kLutBits - l);
                    const int cnt = 1 << (kLutBits - l);
                    const uint16_t e = uint16_t(huff_entry(l, spec.vals[p & 255], is_dc));
                    for (int j = 0; j < cnt; ++j)
                        t.lut[first + j] = e;
                }
            }
            t.maxcode[l] = code - 1;
        } else {
            t.maxcode[l] = -1;
        }
        code <<= 1;
    }
    t.maxcode[0] = -1;
    t.maxcode[17] = 0x7FFFFFFF;
    // second level: group the codes longer than kLutBits by their kLutBits-bit prefix; a prefix whose longest code has
    // kLutBits + n bits gets 2^n entries indexed by the n bits after the prefix
    int maxlen[kLutSize];
    std::memset(maxlen, 0, sizeof(maxlen));
    code = 0;
    for (int l = 1; l <= 16; ++l) {
        for (int i = 0; i < spec.bits[l]; ++i, ++code)
            if (l > kLutBits && (code >> (l - kLutBits)) < kLutSize && l > maxlen[code >> (l - kLutBits)])
                maxlen[code >> (l - kLutBits)] = l;
        code <<= 1;
    }
    int used = 0, off_of[kLutSize];
    for (int pre = 0; pre < kLutSize; ++pre) {
        off_of[pre] = -1;
        if (!maxlen[pre] || t.lut[pre])
            continue; // (a first-level entry there would mean the table is not prefix free: leave it to the walk)
        const int n = maxlen[pre] - kLutBits, size = 1 << n;
        if (used + size > kLut2Size)
            continue; // does not fit: entry stays 0, the kernel walks maxcode[]
        off_of[pre] = used;
        t.lut[pre] = uint16_t((n << 5) | ((used / 2) << 8));
        used += size;
    }
    code = 0;
    p = 0;
    for (int l = 1; l <= 16; ++l) {
        for (int i = 0; i < spec.bits[l]; ++i, ++p, ++code) {
            if (l <= kLutBits || (code >> (l - kLutBits)) >= kLutSize)
                continue;
            const int pre = code >> (l - kLutBits);
            if (off_of[pre] < 0)
                continue;
            const int n = maxlen[pre] - kLutBits, sub = code & ((1 << (l - kLutBits)) - 1);
            const int first = sub << (kLutBits + n - l), cnt = 1 << (kLutBits + n - l);
            const uint16_t e = uint16_t(huff_entry(l, spec.vals[p & 255], is_dc));
            for (int j = 0; j < cnt; ++j)
                t.lut[kLutSize + off_of[pre] + first + j] = e;
        }
        code <<= 1;
    }
}

int resolve_format(int f, int ncomp)
{
    (void)ncomp;
    switch (f) {
    // *_UNCHANGED is 3-channel RGB in the reference plugin even for gray streams
    // (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:336-345,400-412)
    case 1: return 3; // P_UNCHANGED -> planar RGB
    case 2: return 4; // I_UNCHANGED -> interleaved RGB
    case 3: case 4: case 5: case 6: case 7: return f;
    default: return 0;
    }
}

void plan_begin(BatchPlan& plan, const PlanOptions& opt)
{
    plan = BatchPlan();
    plan.subseq_bits = opt.subseq_bytes * 8;
    plan.map_unstuff.push_back(0);
    plan.map_subseq.push_back(0);
    plan.map_fused.push_back(0);
    plan.map_dc.push_back(0);
    plan.map_scan.push_back(0);
    plan.map_idct.push_back(0);
    plan.map_color.push_back(0);
}

struct SeenTable
{
    HuffSpec spec;
    bool is_dc;
};

static int intern_huff(BatchPlan& plan, std::vector<SeenTable>& seen, const HuffSpec& s, bool is_dc)
{
    for (size_t i = 0; i < seen.size(); ++i)
        if (seen[i].is_dc == is_dc && !std::memcmp(seen[i].spec.bits, s.bits, 17) && !std::memcmp(seen[i].spec.vals, s.vals, 256))
            return int(i);
    seen.push_back(SeenTable{s, is_dc});
    plan.huff.emplace_back();
    build_dev_table(s, is_dc, plan.huff.back());
    return int(seen.size() - 1);
}

static int intern_quant(BatchPlan& plan, const uint16_t* q)
{
    const size_t n = plan.quant.size() / 64;
    for (size_t i = 0; i < n; ++i)
        if (!std::memcmp(&plan.quant[i * 64], q, 128))
            return int(i);
    plan.quant.insert(plan.quant.end(), q, q + 64);
    return int(n);
}

// one spec list per plan; kept in a side table keyed by the plan address (plans are built single-threaded)
static thread_local std::vector<SeenTable> g_seen;

void plan_add_image(BatchPlan& plan, const ParsedJpeg& pj, const OutputSpec& out, const PlanOptions& opt)
{
    if (plan.images.empty())
        g_seen.clear();
    ImageDesc im;
    std::memset(&im, 0, sizeof(im));
    im.width = pj.width;
    im.height = pj.height;
    im.ncomp = pj.ncomp;
    im.out_fmt = resolve_format(out.format, pj.ncomp);
    im.ncomp_out = (im.out_fmt == 7) ? 1 : pj.ncomp;
    im.hmax = im.vmax = 1;
    for (int c = 0; c < pj.ncomp; ++c) {
        im.hs[c] = pj.ncomp == 1 ? 1 : pj.hs[c]; // a single-component scan is never interleaved (jdinput.c:166-186)
        im.vs[c] = pj.ncomp == 1 ? 1 : pj.vs[c];
        if (im.hs[c] > im.hmax) im.hmax = im.hs[c];
        if (im.vs[c] > im.vmax) im.vmax = im.vs[c];
    }
    for (int c = pj.ncomp; c < 3; ++c)
        im.hs[c] = im.vs[c] = 1;
    im.mcus_x = (pj.width + 8 * im.hmax - 1) / (8 * im.hmax);
    im.mcus_y = (pj.height + 8 * im.vmax - 1) / (8 * im.vmax);
    int j = 0;
    int dc_tab[3] = {0, 0, 0}, ac_tab[3] = {0, 0, 0};
    for (int c = 0; c < pj.ncomp; ++c) {
        im.blk_first[c] = j;
        for (int k = 0; k < im.hs[c] * im.vs[c]; ++k)
            im.blk_comp[j++] = uint8_t(c);
        im.wib[c] = im.mcus_x * im.hs[c];
        im.hib[c] = im.mcus_y * im.vs[c];
        im.ds_w[c] = (pj.width * im.hs[c] + im.hmax - 1) / im.hmax;
        im.ds_h[c] = (pj.height * im.vs[c] + im.vmax - 1) / im.vmax;
        im.qt[c] = intern_quant(plan, pj.quant[pj.tq[c]]);
        if (pj.sof != 0xC2) {
            dc_tab[c] = intern_huff(plan, g_seen, pj.dc[pj.td[c]], true);
            ac_tab[c] = intern_huff(plan, g_seen, pj.ac[pj.ta[c]], false);
        }
    }
    im.progressive = pj.sof == 0xC2 ? 1 : 0;
    if (im.progressive) { // the scans, with the tables each of them uses interned into the pool
        im.scan_first = uint32_t(plan.scans.size());
        im.scan_count = uint32_t(pj.scans.size());
        for (const ScanInfo& si : pj.scans) {
            ScanDesc sd;
            std::memset(&sd, 0, sizeof(sd));
            sd.off = uint32_t(si.offset - pj.scan_offset);
            sd.len = uint32_t(si.bytes);
            sd.ncomp = si.ncomp;
            sd.ss = si.ss;
            sd.se = si.se;
            sd.ah = si.ah;
            sd.al = si.al;
            sd.restart_interval = si.restart_interval;
            for (int i = 0; i < si.ncomp; ++i) {
                sd.comp[i] = si.comp[i];
                if (si.ss == 0 && si.ah == 0)
                    sd.dc_tab[i] = intern_huff(plan, g_seen, si.dc[si.td[i]], true);
            }
            if (si.ss != 0)
                sd.ac_tab = intern_huff(plan, g_seen, si.ac[si.ta[0]], false);
            // which earlier scans this one has to wait for: those that share a component and whose band overlaps (a scan writes
            // 16-bit coefficients of its own band only, a refinement scan reads nothing outside its band)
            const size_t idx = plan.scans.size() - im.scan_first;
            if (pj.scans.size() > 32) {
                sd.deps = 0xFFFFFFFFu;
            } else {
                for (size_t e = 0; e < idx; ++e) {
                    const ScanDesc& pe = plan.scans[im.scan_first + e];
                    bool shared = false;
                    for (int a = 0; a < sd.ncomp; ++a)
                        for (int b2 = 0; b2 < pe.ncomp; ++b2)
                            shared = shared || sd.comp[a] == pe.comp[b2];
                    if (shared && sd.ss <= pe.se && pe.ss <= sd.se)
                        sd.deps |= 1u << e;
                }
            }
            plan.scans.push_back(sd);
        }
        plan.prog_images.push_back(uint32_t(plan.images.size()));
    }
    im.bpm = j;
    // table selectors: distinct pool tables get shared-memory slots in order of first use
    im.ntabs = 0;
    auto slot_of = [&](int pool_idx) {
        for (int s = 0; s < im.ntabs; ++s)
            if (im.tab_pool[s] == pool_idx)
                return s;
        im.tab_pool[im.ntabs] = pool_idx;
        return im.ntabs++;
    };
    for (int b = 0; b < im.bpm; ++b) {
        const int c = im.blk_comp[b];
        im.tabsel_g[b] = TabSel{uint32_t(dc_tab[c]) * kHuffTableU16, uint32_t(ac_tab[c]) * kHuffTableU16, uint32_t(dc_tab[c]), uint32_t(ac_tab[c])};
        im.tabsel_s[b] = TabSel{uint32_t(slot_of(dc_tab[c])) * kLutEntries, uint32_t(slot_of(ac_tab[c])) * kLutEntries, uint32_t(dc_tab[c]),
            uint32_t(ac_tab[c])};
    }
    // the reconstruction kernel stages the image's distinct AC tables only
    im.nac = 0;
    for (int c = 0; c < pj.ncomp; ++c) {
        int s = 0;
        for (; s < im.nac; ++s)
            if (im.ac_pool[s] == ac_tab[c])
                break;
        if (s == im.nac)
            im.ac_pool[im.nac++] = ac_tab[c];
        im.ac_slot[c] = s;
    }
    const uint32_t mcus = uint32_t(im.mcus_x) * uint32_t(im.mcus_y);
    im.total_blocks = mcus * uint32_t(im.bpm);
    im.restart_interval = pj.restart_interval;
    if (!im.progressive && pj.restart_interval > 0 && uint32_t(pj.restart_interval) < mcus) { // (a progressive stream's restart intervals belong to its scans)
        im.num_segments = int(cdiv(mcus, pj.restart_interval));
        im.seg_blocks = uint32_t(pj.restart_interval) * uint32_t(im.bpm);
    } else {
        im.num_segments = 1;
        im.seg_blocks = im.total_blocks;
        im.restart_interval = 0;
    }
    im.seg_first = plan.total_segs;
    plan.total_segs += uint32_t(im.num_segments) + 1;

    im.raw_off = plan.raw_bytes;
    im.stage_len = uint32_t(pj.scan_bytes);
    im.raw_len = im.progressive ? 0u : im.stage_len; // the baseline Huffman kernels see nothing of a progressive stream
    plan.raw_bytes = align_up(plan.raw_bytes + im.stage_len + 1, 16);
    im.clean_off = plan.clean_bytes;
    plan.clean_bytes = align_up(plan.clean_bytes + im.raw_len + 16, 16);

    im.chunk_first = plan.total_chunks;
    im.chunk_count = cdiv(im.raw_len, kChunkBytes);
    if (im.chunk_count == 0) im.chunk_count = 1;
    plan.total_chunks += im.chunk_count;
    plan.map_unstuff.push_back(plan.total_chunks);

    im.subseq_count = cdiv(uint64_t(im.raw_len) * 8, plan.subseq_bits);
    if (im.subseq_count == 0) im.subseq_count = 1;
    im.subseq_cta_count = cdiv(im.subseq_count, kSubseqCta);
    im.subseq_first = plan.total_subseq;
    im.subseq_cta_first = plan.map_subseq.back();
    plan.total_subseq += im.subseq_cta_count * kSubseqCta;
    plan.map_subseq.push_back(im.subseq_cta_first + im.subseq_cta_count);
    im.fused_cta_first = plan.map_fused.back();
    im.fused_cta_count = cdiv(im.subseq_count, kFusedOwn);
    plan.map_fused.push_back(im.fused_cta_first + im.fused_cta_count);
    im.dc_cta_first = plan.map_dc.back();
    im.dc_cta_count = std::max<uint32_t>(1, cdiv(uint64_t(im.mcus_x) * im.mcus_y, kDcChunkMcus));
    plan.map_dc.push_back(im.dc_cta_first + im.dc_cta_count);
    im.scan_cta_first = plan.map_scan.back();
    im.scan_cta_count = cdiv(uint64_t(im.subseq_cta_count) * kSubseqCta, kScanChunk);
    plan.map_scan.push_back(im.scan_cta_first + im.scan_cta_count);

    im.coef_off = plan.coef_elems;
    plan.coef_elems = align_up(plan.coef_elems + uint64_t(im.total_blocks) * 64, 64);
    uint64_t blocks_out = 0;
    for (int c = 0; c < pj.ncomp; ++c) {
        im.plane_pitch[c] = im.wib[c] * 8;
        im.plane_off[c] = plan.plane_bytes;
        plan.plane_bytes = align_up(plan.plane_bytes + uint64_t(im.plane_pitch[c]) * im.hib[c] * 8 + 16, 128);
        if (c < im.ncomp_out)
            blocks_out += uint64_t(im.wib[c]) * im.hib[c];
    }
    im.idct_cta_first = plan.map_idct.back();
    (void)blocks_out;
    im.idct_cta_count = cdiv(im.total_blocks, kIdctCta); // blocks in MCU scan order, the order of the stream
    plan.map_idct.push_back(im.idct_cta_first + im.idct_cta_count);
    const int vy = pj.ncomp == 1 ? 1 : im.vmax / im.vs[1];
    im.color_rows = vy == 2 ? im.ds_h[1] + 1 : pj.height; // row pairs (2q+1, 2q+2), q = -1 .. ds_h-1, when chroma is v-subsampled
    im.color_xtiles = int(cdiv(pj.width, kColorTilePx));
    im.color_cta_first = plan.map_color.back();
    im.color_cta_count = cdiv(uint64_t(im.color_rows) * im.color_xtiles, kColorCta / 32);
    plan.map_color.push_back(im.color_cta_first + im.color_cta_count);

    im.out = static_cast<uint8_t*>(out.buffer);
    im.out_stride = out.row_stride;
    im.fancy = opt.fancy_upsampling ? 1 : 0;
    const bool whole = out.roi_w <= 0 || out.roi_h <= 0;
    im.roi_x = whole ? 0 : out.roi_x;
    im.roi_y = whole ? 0 : out.roi_y;
    im.roi_w = whole ? pj.width : out.roi_w;
    im.roi_h = whole ? pj.height : out.roi_h;
    im.orient = out.orient & 15;

    plan.total_blocks += uint64_t(mcus) * im.bpm;
    plan.total_pixels += uint64_t(pj.width) * pj.height;
    const int planes = (im.out_fmt == 7) ? 1 : 3;
    plan.out_bytes += uint64_t(im.roi_w) * im.roi_h * planes;
    plan.images.push_back(im);
}

void plan_finish(BatchPlan& plan)
{
    plan.raw_bytes = align_up(plan.raw_bytes + 64, 256);
    plan.clean_bytes = align_up(plan.clean_bytes + 64, 256);
    plan.coef_elems = align_up(plan.coef_elems + 64, 128);
    plan.plane_bytes = align_up(plan.plane_bytes + 64, 256);
}

} // namespace b200jpeg
