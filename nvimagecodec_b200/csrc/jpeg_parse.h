// jpeg_parse.h -- host-side JPEG marker parser for the b200jpeg decoder.
// Covers what the reference obtains from jpeg_read_header (external/libjpeg-turbo/jdmarker.c get_sof:242, get_sos:308,
// get_dht:437, get_dqt:511, get_dri:569, next_marker:892) plus the geometry of jdinput.c initial_setup:47 /
// per_scan_setup:156; the reference's own builtin probe (src/parsers/jpeg.cpp:202-361) stops at SOF and never reads
// DHT/DQT/DRI/SOS, so the plugin needs this.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace b200jpeg {

struct HuffSpec // one DHT table as transmitted
{
    bool present = false;
    uint8_t bits[17] = {0}; // bits[1..16]
    uint8_t vals[256] = {0};
    int count = 0;
};

// One scan of a progressive stream (SOF2): what get_sos reads (ljt/jdmarker.c:308-376) plus the tables and the restart
// interval in force when the scan starts (DHT / DRI may be sent again between scans) and where its entropy-coded bytes are.
struct ScanInfo
{
    int ncomp = 0;
    int comp[3] = {0, 0, 0};          // indices into the frame's components
    int td[3] = {0, 0, 0}, ta[3] = {0, 0, 0};
    int ss = 0, se = 0, ah = 0, al = 0;
    int restart_interval = 0;
    size_t offset = 0, bytes = 0;     // entropy-coded bytes (stuffed, RSTn included) inside the file
    HuffSpec dc[4], ac[4];            // only the tables the scan refers to are filled in
};

struct ParsedJpeg
{
    int width = 0, height = 0, precision = 0, ncomp = 0;
    int sof = 0; // marker byte 0xC0..0xCF
    int comp_id[3] = {0, 0, 0};
    int hs[3] = {1, 1, 1}, vs[3] = {1, 1, 1}, tq[3] = {0, 0, 0}, td[3] = {0, 0, 0}, ta[3] = {0, 0, 0};
    int restart_interval = 0;
    bool jfif = false, adobe = false;
    int adobe_transform = -1;
    uint16_t quant[4][64]; // natural (de-zigzagged) order
    bool quant_present[4] = {false, false, false, false};
    HuffSpec dc[4], ac[4];
    size_t scan_offset = 0, scan_bytes = 0; // entropy-coded segment [offset, offset+bytes) (EOI stripped)
    int scan_ncomp = 0;
    bool single_interleaved_scan = false; // SOS lists all frame components in frame order, Ss=0 Se=63 Ah=Al=0
    std::vector<ScanInfo> scans;          // progressive streams only: every scan of the file, in file order
    bool progression_ok = false;          // progressive: the scans are a valid progression that ends with every coefficient at full precision
};

enum ParseResult
{
    PARSE_OK = 0,
    PARSE_BAD = 1,        // not a JPEG / truncated / inconsistent
    PARSE_UNSUPPORTED = 2 // valid but outside scope
};

// Walk markers up to and including SOS. Never reads entropy-coded data except to locate the trailing EOI.
ParseResult parse_jpeg(const uint8_t* data, size_t size, ParsedJpeg& out);

// Scope check of a successfully parsed stream (baseline/extended-sequential Huffman 8-bit, 1 or 3 components, YCbCr,
// luma sampling in {1,2}x{1,2} with 1x1 chroma, one interleaved scan, tables present).
bool is_supported(const ParsedJpeg& p);

// chroma subsampling as nvimgcodecChromaSubsampling_t value, -1 if other
int subsampling_of(const ParsedJpeg& p);

extern const uint8_t kZigzagToNatural[64 + 16];

} // namespace b200jpeg
