// jpeg_parse.cpp -- see jpeg_parse.h
#include "jpeg_parse.h"
#include <cstring>

namespace b200jpeg {

// zigzag position -> natural (row-major) position; 16 guard entries like jutils.c:59 so a corrupt run cannot index out
const uint8_t kZigzagToNatural[64 + 16] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20,
    13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54,
    47, 55, 62, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};

namespace {
inline unsigned be16(const uint8_t* p) { return (unsigned(p[0]) << 8) | p[1]; }

// End of the entropy-coded data that starts at `from`: the position of the FF of the first marker that is neither stuffing
// (FF00), fill (FF FF) nor RSTn -- what jpeg_fill_bit_buffer stops at (ljt/jdhuff.c:287-391); n if there is none.
size_t entropy_end(const uint8_t* d, size_t n, size_t from)
{
    size_t p = from;
    while (p + 1 < n) {
        const uint8_t* q = static_cast<const uint8_t*>(std::memchr(d + p, 0xFF, n - 1 - p));
        if (!q)
            return n;
        p = size_t(q - d);
        const int x = d[p + 1];
        if (x == 0x00 || x == 0xFF || (x >= 0xD0 && x <= 0xD7)) {
            p += (x == 0xFF) ? 1 : 2;
            continue;
        }
        return p;
    }
    return n;
}

// The progression rules of ljt/jdphuff.c start_pass_phuff_decoder:96-211 (taken strictly: what libjpeg only warns about is
// declined here), and completeness: at the end every coefficient of every component has been sent down to bit 0.
bool check_progression(const ParsedJpeg& o)
{
    int bits[3][64];
    for (auto& c : bits)
        for (int& b : c)
            b = -1;
    for (const ScanInfo& sc : o.scans) {
        if (sc.ss < 0 || sc.ss > sc.se || sc.se > 63 || sc.al < 0 || sc.al > 13 || sc.ah < 0 || sc.ah > 13)
            return false;
        if (sc.ss == 0 && sc.se != 0)
            return false; // DC and AC coefficients never share a scan
        if (sc.ss != 0 && sc.ncomp != 1)
            return false; // AC scans carry one component
        if (sc.ah != 0 && sc.al != sc.ah - 1)
            return false;
        for (int i = 0; i < sc.ncomp; ++i) {
            const int c = sc.comp[i];
            if (sc.ss != 0 && bits[c][0] < 0)
                return false; // AC before the component's DC
            for (int k = sc.ss; k <= sc.se; ++k) {
                const int expected = bits[c][k] < 0 ? 0 : bits[c][k];
                if (sc.ah != expected || (bits[c][k] < 0 && sc.ah != 0) || (bits[c][k] >= 0 && sc.ah == 0))
                    return false;
                bits[c][k] = sc.al;
            }
        }
    }
    for (int c = 0; c < o.ncomp; ++c)
        for (int k = 0; k < 64; ++k)
            if (bits[c][k] != 0)
                return false;
    return !o.scans.empty();
}
} // namespace

ParseResult parse_jpeg(const uint8_t* d, size_t n, ParsedJpeg& o)
{
    o = ParsedJpeg();
    if (!d || n < 4 || d[0] != 0xFF || d[1] != 0xD8)
        return PARSE_BAD;
    size_t p = 2;
    bool have_sof = false;
    while (true) {
        // next_marker: skip any non-FF garbage is an error for us; skip FF fill bytes
        if (p + 2 > n || d[p] != 0xFF)
            return PARSE_BAD;
        while (p < n && d[p] == 0xFF)
            ++p;
        if (p >= n)
            return PARSE_BAD;
        const int m = d[p++];
        if (m == 0x01 || (m >= 0xD0 && m <= 0xD8))
            continue; // parameterless
        if (m == 0xD9) {
            if (o.scans.empty())
                return PARSE_BAD; // EOI before SOS
            o.progression_ok = check_progression(o);
            return PARSE_OK;
        }
        if (p + 2 > n)
            return PARSE_BAD;
        const size_t L = be16(d + p);
        if (L < 2 || p + L > n)
            return PARSE_BAD;
        const uint8_t* s = d + p + 2;
        size_t sl = L - 2;
        switch (m) {
        case 0xDB: // DQT
            while (sl > 0) {
                const int pq = s[0] >> 4, tq = s[0] & 15;
                const size_t need = 1 + (pq ? 128 : 64);
                if (tq > 3 || pq > 1 || sl < need)
                    return PARSE_BAD;
                for (int i = 0; i < 64; ++i)
                    o.quant[tq][kZigzagToNatural[i]] = pq ? uint16_t(be16(s + 1 + 2 * i)) : s[1 + i];
                o.quant_present[tq] = true;
                s += need;
                sl -= need;
            }
            break;
        case 0xC4: // DHT
            while (sl > 0) {
                if (sl < 17)
                    return PARSE_BAD;
                const int tc = s[0] >> 4, th = s[0] & 15;
                if (tc > 1 || th > 3)
                    return PARSE_BAD;
                HuffSpec& t = tc ? o.ac[th] : o.dc[th];
                int cnt = 0;
                t.bits[0] = 0;
                for (int i = 1; i <= 16; ++i)
                    cnt += (t.bits[i] = s[i]);
                if (cnt > 256 || size_t(17 + cnt) > sl)
                    return PARSE_BAD;
                std::memset(t.vals, 0, sizeof(t.vals));
                std::memcpy(t.vals, s + 17, cnt);
                t.count = cnt;
                t.present = true;
                s += 17 + cnt;
                sl -= 17 + cnt;
            }
            break;
        case 0xC0: case 0xC1: case 0xC2: case 0xC3: case 0xC5: case 0xC6: case 0xC7:
        case 0xC9: case 0xCA: case 0xCB: case 0xCD: case 0xCE: case 0xCF: { // SOFn
            if (have_sof || sl < 6)
                return PARSE_BAD;
            o.sof = m;
            o.precision = s[0];
            o.height = be16(s + 1);
            o.width = be16(s + 3);
            o.ncomp = s[5];
            if (o.width == 0 || o.height == 0 || o.ncomp == 0 || sl < size_t(6 + 3 * o.ncomp))
                return PARSE_BAD;
            if (o.ncomp > 3) {
                have_sof = true; // CMYK/YCCK etc.: valid, unsupported
                break;
            }
            for (int c = 0; c < o.ncomp; ++c) {
                o.comp_id[c] = s[6 + 3 * c];
                o.hs[c] = s[7 + 3 * c] >> 4;
                o.vs[c] = s[7 + 3 * c] & 15;
                o.tq[c] = s[8 + 3 * c];
                if (o.hs[c] < 1 || o.hs[c] > 4 || o.vs[c] < 1 || o.vs[c] > 4 || o.tq[c] > 3)
                    return PARSE_BAD;
            }
            have_sof = true;
            break;
        }
        case 0xDD: // DRI
            if (sl < 2)
                return PARSE_BAD;
            o.restart_interval = be16(s);
            break;
        case 0xE0:
            if (sl >= 5 && !std::memcmp(s, "JFIF", 5))
                o.jfif = true;
            break;
        case 0xEE:
            if (sl >= 12 && !std::memcmp(s, "Adobe", 5)) {
                o.adobe = true;
                o.adobe_transform = s[11];
            }
            break;
        case 0xDA: { // SOS
            if (!have_sof || sl < 1)
                return PARSE_BAD;
            const int ns = s[0];
            if (ns < 1 || ns > 4 || sl < size_t(1 + 2 * ns + 3))
                return PARSE_BAD;
            o.scan_ncomp = ns;
            bool in_order = (ns == o.ncomp) && o.ncomp <= 3;
            for (int i = 0; i < ns && in_order; ++i) {
                if (s[1 + 2 * i] != o.comp_id[i])
                    in_order = false;
                o.td[i] = s[2 + 2 * i] >> 4;
                o.ta[i] = s[2 + 2 * i] & 15;
                if (o.td[i] > 3 || o.ta[i] > 3)
                    return PARSE_BAD;
            }
            const uint8_t* tail = s + 1 + 2 * ns;
            if (o.sof == 0xC2 && o.ncomp <= 3) { // progressive: note the scan, skip its entropy-coded bytes, go on to the next marker
                ScanInfo sc;
                sc.ncomp = ns;
                if (ns > 3)
                    return PARSE_BAD;
                for (int i = 0; i < ns; ++i) {
                    int c = -1;
                    for (int k = 0; k < o.ncomp; ++k)
                        if (o.comp_id[k] == s[1 + 2 * i])
                            c = k;
                    if (c < 0 || (i > 0 && c <= sc.comp[i - 1]))
                        return PARSE_BAD; // unknown component, or not in frame order (ljt/jdmarker.c:336-354)
                    sc.comp[i] = c;
                    sc.td[i] = s[2 + 2 * i] >> 4;
                    sc.ta[i] = s[2 + 2 * i] & 15;
                    if (sc.td[i] > 3 || sc.ta[i] > 3)
                        return PARSE_BAD;
                    sc.dc[sc.td[i]] = o.dc[sc.td[i]];
                    sc.ac[sc.ta[i]] = o.ac[sc.ta[i]];
                }
                sc.ss = tail[0];
                sc.se = tail[1];
                sc.ah = tail[2] >> 4;
                sc.al = tail[2] & 15;
                sc.restart_interval = o.restart_interval;
                sc.offset = p + L;
                const size_t e = entropy_end(d, n, sc.offset);
                sc.bytes = e - sc.offset;
                if (o.scans.empty())
                    o.scan_offset = sc.offset;
                o.scan_bytes = e - o.scan_offset; // everything from the first scan's data to the end of the last one's
                o.scans.push_back(sc);
                o.scan_ncomp = ns;
                if (e >= n) { // the file ends inside this scan: no EOI
                    o.progression_ok = false;
                    return PARSE_OK;
                }
                p = e;
                continue;
            }
            o.single_interleaved_scan = in_order && tail[0] == 0 && tail[1] == 63 && tail[2] == 0;
            o.scan_offset = p + L;
            // entropy-coded segment ends at the last EOI of the file (search backwards, tolerate trailing bytes)
            size_t e = n;
            while (e >= o.scan_offset + 2 && !(d[e - 2] == 0xFF && d[e - 1] == 0xD9))
                --e;
            if (e < o.scan_offset + 2)
                o.scan_bytes = n - o.scan_offset; // no EOI: the reference warns and pads; we take everything
            else
                o.scan_bytes = e - 2 - o.scan_offset;
            return PARSE_OK;
        }
        default:
            break; // APPn, COM, DNL... skipped
        }
        p += L;
    }
}

int subsampling_of(const ParsedJpeg& p)
{
    if (p.ncomp == 1)
        return 7; // GRAY
    if (p.ncomp != 3 || p.hs[1] != 1 || p.vs[1] != 1 || p.hs[2] != 1 || p.vs[2] != 1)
        return -1;
    const int h = p.hs[0], v = p.vs[0];
    if (h == 1 && v == 1) return 0;
    if (h == 2 && v == 1) return 2;
    if (h == 2 && v == 2) return 3;
    if (h == 1 && v == 2) return 4;
    if (h == 4 && v == 1) return 5;
    if (h == 4 && v == 2) return 6;
    return -1;
}

bool is_supported(const ParsedJpeg& p)
{
    if (p.sof != 0xC0 && p.sof != 0xC1 && p.sof != 0xC2)
        return false; // lossless, arithmetic, differential
    if (p.precision != 8)
        return false;
    if (p.ncomp != 1 && p.ncomp != 3)
        return false;
    const bool progressive = p.sof == 0xC2;
    if (progressive ? !p.progression_ok : !p.single_interleaved_scan)
        return false; // a progression that is incomplete or irregular is the CPU decoder's business (it smooths what is missing)
    if (p.ncomp == 3) {
        const int css = subsampling_of(p);
        if (css != 0 && css != 2 && css != 3 && css != 4 && css != 5 && css != 6) // 4:4:4, 4:2:2, 4:2:0, 4:4:0, 4:1:1, 4:1:0
            return false;
        // colour space decision of jdapimin.c default_decompress_parms:117-196: JFIF => YCbCr; Adobe transform 0 => RGB;
        // no marker: ids 'R','G','B' => RGB, else YCbCr.  RGB-coded JPEGs are out of scope.
        if (p.adobe && !p.jfif && p.adobe_transform == 0)
            return false;
        if (!p.adobe && !p.jfif && p.comp_id[0] == 'R' && p.comp_id[1] == 'G' && p.comp_id[2] == 'B')
            return false;
    }
    if (progressive) {
        for (int c = 0; c < p.ncomp; ++c)
            if (!p.quant_present[p.tq[c]])
                return false;
        for (const ScanInfo& sc : p.scans)
            for (int i = 0; i < sc.ncomp; ++i) {
                if (sc.ss == 0 && sc.ah == 0 && !sc.dc[sc.td[i]].present)
                    return false;
                if (sc.ss != 0 && !sc.ac[sc.ta[i]].present)
                    return false;
            }
        return size_t(p.width) * size_t(p.height) <= (size_t(1) << 28);
    }
    for (int c = 0; c < p.ncomp; ++c) {
        if (!p.quant_present[p.tq[c]] || !p.dc[p.td[c]].present || !p.ac[p.ta[c]].present)
            return false;
        // 8-bit baseline coefficients: DC categories 0..11, AC sizes 0..10 (T.81 F.1.2); tables that list anything larger
        // cannot come from an 8-bit encoder (the 12-bit value field of a coefficient record relies on this)
        int ndc = 0, nac = 0;
        for (int l = 1; l <= 16; ++l) {
            ndc += p.dc[p.td[c]].bits[l];
            nac += p.ac[p.ta[c]].bits[l];
        }
        for (int i = 0; i < ndc && i < 256; ++i)
            if (p.dc[p.td[c]].vals[i] > 11)
                return false;
        for (int i = 0; i < nac && i < 256; ++i)
            if ((p.ac[p.ta[c]].vals[i] & 15) > 10)
                return false;
    }
    if (size_t(p.width) * size_t(p.height) > (size_t(1) << 28))
        return false;
    return true;
}

} // namespace b200jpeg
