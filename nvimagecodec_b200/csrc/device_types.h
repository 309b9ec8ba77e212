// device_types.h -- structures shared by the host scheduler and the sm_100a kernels of b200jpeg.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define B2J_HD __host__ __device__ __forceinline__
#else
#define B2J_HD inline
#endif

namespace b200jpeg {

#ifndef B2J_LUT_BITS
#define B2J_LUT_BITS 11
#endif
constexpr int kLutBits = B2J_LUT_BITS; // first-level Huffman lookup width (build-time: variant builds)
constexpr int kLutSize = 1 << kLutBits;
constexpr int kLut2Size = 512;        // second-level entries (codes longer than kLutBits)
constexpr int kLutEntries = kLutSize + kLut2Size;
#ifndef B2J_SUBSEQ_CTA
#define B2J_SUBSEQ_CTA 256
#endif
constexpr int kSubseqCta = B2J_SUBSEQ_CTA; // threads per CTA in the subsequence kernels (one subsequence per thread)
constexpr int kTailCta = 256;         // threads per CTA of the work-list rounds (any subsequence, any thread)
constexpr int kFusedHalo = 8;         // fused Huffman kernel: subsequences before a CTA's own range that it decodes as well, so that the
                                      // decoder state at the start of its own range is (almost always) known without its predecessor
constexpr int kFusedOwn = kSubseqCta - kFusedHalo; // subsequences a CTA of the fused kernel owns
constexpr int kFusedRunHeads = 6;     // fused kernel, rounds from this one on: of a run of consecutive listed subsequences only the head is
                                      // worked on, by one thread that walks the whole run (streams that never lock on, e.g. flat image areas)
constexpr int kUnstuffThreads = 256;  // threads per CTA in the unstuffing kernels
constexpr int kUnstuffGranules = 4;   // 16-byte granules per thread (consecutive: 64 bytes of the stream per thread)
constexpr int kUnstuffBytesPerThread = 16 * kUnstuffGranules;
constexpr int kChunkBytes = kUnstuffThreads * kUnstuffBytesPerThread; // raw bytes per unstuff CTA (16 KB)
#ifndef B2J_IDCT_CTA
#define B2J_IDCT_CTA 256
#endif
constexpr int kIdctCta = B2J_IDCT_CTA; // threads per CTA, one 8x8 block per thread
constexpr int kColorCta = 128;       // threads per CTA: 4 warps, each warp one tile of 512 pixels x (1|2) rows
constexpr int kColorTilePx = 512;    // pixels per warp tile (16 per lane)
constexpr int kDcCta = 512;          // threads per CTA, one MCU per thread per iteration
constexpr int kDcBatch = 8;          // iterations a warp of the DC kernel keeps in flight
constexpr int kScanChunk = 256 * 8;   // subsequences per CTA of k_subseq_scan (256 threads, eight iterations per warp range)
constexpr int kDcChunkMcus = kDcCta * kDcBatch; // MCUs per CTA of the DC kernel: big images are cut into chunks that hand each other the predictors
constexpr int kMaxBlocksPerMcu = 10;  // the limit of the JPEG standard: luma 4x2 + Cb + Cr (4:1:0)
constexpr int kMaxTabSlots = 6;       // distinct Huffman tables one image can reference (3 components x DC/AC)

// Decode table derived from one DHT table (reference: d_derived_tbl, external/libjpeg-turbo/jdhuff.h:27-51 and
// jpeg_make_d_derived_tbl, jdhuff.c:143-285).  We use an 11-bit first level instead of 8 and store pre-digested entries
//   entry = (code_length + size) | size << 5 | zadv << 9
// size = number of value bits that follow the code (so the low 5 bits are the total number of bits the symbol takes:
// at most 16 + 15); zadv = how far the zigzag position advances (DC: 1; AC run/size: run + 1; ZRL: 16; EOB: 64).  Codes longer than kLutBits: the first-level entry of their 11-bit prefix has code_length 0
// and holds n << 5 | (offset / 2) << 8; the entry of the code is lut[kLutSize + offset + next n bits] (same format, full
// length).  entry == 0: invalid code, or a table too irregular for kLut2Size second-level entries -> maxcode walk.
// The same symbols come out as with the reference's tables.
struct DevHuffTable
{
    uint16_t lut[kLutEntries];
    int32_t maxcode[18];   // largest code of each length, -1 if none
    int32_t valoffset[18]; // vals[] index = code + valoffset[len]
    uint8_t vals[256];
    int32_t is_dc;
    int32_t pad_[3];       // sizeof == 5536, a multiple of 16
};
constexpr uint32_t kHuffTableU16 = sizeof(DevHuffTable) / 2; // stride of the table pool in uint16 units

B2J_HD uint32_t huff_entry(int len, int sym, bool is_dc)
{
    const int s = sym & 15;
    int zadv = 1;
    if (!is_dc) {
        const int r = sym >> 4;
        zadv = s ? r + 1 : (r == 15 ? 16 : 64);
    }
    return uint32_t(len + s) | (uint32_t(s) << 5) | (uint32_t(zadv) << 9);
}

// Table selector of one block position inside the MCU: offsets (uint16 units from the LUT base the kernel uses) and pool
// indices (for the long-code walk) of the DC and AC tables of that block's component.
struct alignas(16) TabSel
{
    uint32_t dc_off, ac_off, dc_idx, ac_idx;
};

// One scan of a progressive stream (jpeg_parse.h ScanInfo with the tables interned into the batch's pool).
struct ScanDesc
{
    uint32_t off, len;        // entropy-coded bytes (stuffed, RSTn included), relative to the image's bytes in the raw buffer
    int32_t ncomp;
    int32_t comp[3];
    int32_t dc_tab[3];        // pool indices (DC first scans)
    int32_t ac_tab;           // pool index (AC scans: one component)
    int32_t ss, se, ah, al;
    int32_t restart_interval; // MCUs of this scan between RSTn markers, 0 = none
    uint32_t deps;            // bit j: scan j of the image (j < this one) touches coefficients this scan reads or writes (a component in common and
                              // overlapping bands) and has to be complete before this scan starts; scans without such a link run side by side
                              // (k_prog_decode).  Images with more than 32 scans: every earlier scan (all ones).
};

// Per-image descriptor, built on the host, read-only on the device.
struct ImageDesc
{
    int32_t width, height;
    int32_t ncomp;        // components in the JPEG (1 or 3)
    int32_t ncomp_out;    // components that must be reconstructed (1 for Y-only output)
    int32_t hs[3], vs[3];
    int32_t hmax, vmax;
    int32_t mcus_x, mcus_y;
    int32_t bpm;          // blocks per MCU
    int32_t restart_interval; // MCUs per restart segment (0 => single segment)
    int32_t num_segments;
    uint32_t seg_blocks;   // blocks per full segment (restart_interval * bpm, or total)
    uint32_t total_blocks; // mcus * bpm
    int32_t wib[3], hib[3];   // MCU-padded block grid per component
    int32_t ds_w[3], ds_h[3]; // downsampled_width / height (jdinput.c:130-135)
    int32_t blk_first[3];     // first block-in-MCU index of each component
    uint8_t blk_comp[12];     // component of block j in the MCU
    int32_t qt[3];            // quant table pool index per component

    // Huffman tables: per block-in-MCU selectors, relative to the global pool (tail kernel) and to the per-CTA shared
    // memory copy (main kernels); tab_pool lists the distinct pool tables to stage, slot s -> shared offset s * kLutEntries
    TabSel tabsel_g[kMaxBlocksPerMcu];
    TabSel tabsel_s[kMaxBlocksPerMcu];
    int32_t ntabs;
    int32_t tab_pool[kMaxTabSlots];

    // progressive streams (SOF2): the scans are decoded one after the other into dense coefficients (progressive.cuh), the
    // baseline Huffman kernels see an empty stream (raw_len == 0) and the reconstruction kernel reads the coefficients
    int32_t progressive;
    uint32_t scan_first, scan_count; // into BatchDev::scans
    uint32_t stage_len;   // bytes of the file that are copied to raw_off (== raw_len for baseline streams)

    // entropy-coded bytes
    uint64_t raw_off;     // byte offset in the batch raw buffer (16-byte aligned)
    uint32_t raw_len;
    uint64_t clean_off;   // byte offset in the unstuffed buffer (16-byte aligned)
    uint32_t chunk_first, chunk_count;   // unstuff CTAs of this image
    uint32_t subseq_first, subseq_count; // subsequence slots (multiple of kSubseqCta) / used slots
    uint32_t subseq_cta_first, subseq_cta_count;
    uint32_t fused_cta_first, fused_cta_count; // CTAs of the fused Huffman kernel (kFusedOwn subsequences each)
    uint32_t dc_cta_first, dc_cta_count;       // CTAs of the DC kernel (kDcChunkMcus MCUs each)
    uint32_t scan_cta_first, scan_cta_count;   // CTAs of k_subseq_scan (kScanChunk subsequences each)
    uint32_t seg_first;   // index into the segment-start pool; entries [seg_first, seg_first + num_segments]

    // Huffman tables of the reconstruction kernel: it decodes the AC symbols of a block itself and stages only AC tables
    int32_t nac;              // distinct AC tables of the image
    int32_t ac_pool[3];       // their pool indices; slot s sits at s * kLutEntries of the kernel's staged LUTs
    int32_t ac_slot[3];       // per component: slot of its AC table

    // intermediates
    uint64_t coef_off;        // int16 units, blocks in MCU scan order (coef_off / 64 = first block in the DC / block-position arrays)
    uint64_t plane_off[3];    // bytes
    int32_t plane_pitch[3];
    uint32_t idct_cta_first, idct_cta_count;
    uint32_t color_cta_first, color_cta_count;
    int32_t color_rows;       // row groups the colour kernel walks (height, or ds_h + 1 row pairs when vmax == 2)
    int32_t color_xtiles;     // warp tiles per row group

    // output
    uint8_t* out;
    int64_t out_stride;
    int32_t out_fmt;      // b200jpegFormat_t resolved: P_RGB/I_RGB/P_BGR/I_BGR/P_Y
    int32_t fancy;
    // region of interest: the output buffer holds pixels [roi_x, roi_x + roi_w) x [roi_y, roi_y + roi_h) of the image (the
    // whole image by default).  The reference decodes a region to exactly the pixels the full decode has there
    // (extensions/libjpeg_turbo/jpeg_mem.cpp:206-240 asks libjpeg for one spare pixel left and right so that the upsampling
    // filter at the region's edge sees its real neighbours).
    int32_t roi_x, roi_y, roi_w, roi_h;
    // EXIF orientation applied to the output (nvimgcodecOrientation_t, reference include/nvimgcodec.h; meaning fixed by
    // src/parsers/exif_orientation.h:37-60): the image is rotated counter-clockwise by 90 * (orient & 3) degrees, then mirrored
    // horizontally (bit 2, flip_x) and vertically (bit 3, flip_y).  The output buffer has the oriented size.  0: as stored.
    int32_t orient;
};

// Per-image state written by the device.
struct ImageState
{
    uint32_t clean_len;     // unstuffed bytes
    uint32_t marker_count;  // RSTn markers seen
    uint32_t error;         // bit flags, see kErr*
    uint32_t done_blocks;   // set by the thread that completes the last block (== total_blocks when the scan is complete)
    uint32_t reserved0;
    uint32_t trunc;         // raw_len - position of the first marker other than RSTn inside the scan data (0: none)
    uint32_t raw_eff;       // entropy-coded bytes actually taken: raw_len - trunc
    uint32_t pad_[1];
};

constexpr uint32_t kErrBadMarker = 1;     // wrong RSTn order / count
constexpr uint32_t kErrBadCode = 2;       // invalid Huffman code on the synchronised path
constexpr uint32_t kErrNotConverged = 4;  // subsequence synchronisation incomplete
constexpr uint32_t kErrProgressive = 16;  // a progressive scan ran out of data / hit a bad code / a marker is missing
constexpr uint32_t kErrOverrun = 8;       // decode ran past the end of a segment / image

// State of the Huffman decoder between two symbols, packed in 64 bits: bit position | (block-in-MCU * 64 + zigzag index) << 32
B2J_HD unsigned long long pack_state(uint32_t p, uint32_t slot) { return (unsigned long long)p | ((unsigned long long)slot << 32); }

constexpr int kRunAheadFrom = 4; // work-list rounds from this one on: one thread per run of consecutive listed subsequences
                                 // walks the run and on into what follows while exit states keep changing (the lists are
                                 // short by then, so nothing is serialised that would have run in parallel); with the
                                 // shortcut of huffman.cuh subseq_repeats a periodic stream is walked at memory speed
                                 // instead of one subsequence per round

constexpr int kTraceRounds = 16; // work-list rounds whose length and time are kept for B200JPEG_TRACE


struct SyncCounters // device-side bookkeeping of the synchronisation rounds
{
    uint32_t count[3];    // work-list lengths (rotating: round r reads [r % 3], appends to [(r + 1) % 3])
    uint32_t rounds;      // tail rounds executed
    uint32_t redone;      // subsequences re-decoded in tail rounds
    uint32_t ticket;      // fused kernel: CTAs take their position in launch order from here (a CTA only ever waits for one that started before it)
    uint32_t chain_redo;  // fused kernel: CTAs whose halo did not deliver the true entry state (they re-walked their range)
    uint32_t ticket_dc;   // DC kernel: the same for its chain of chunks
    uint32_t ticket_scan; // k_subseq_scan: likewise
    uint32_t pad2_[3];
    uint32_t round_items[kTraceRounds]; // diagnostics (B200JPEG_TRACE): list length of each work-list round ...
    uint32_t round_ns[kTraceRounds];    // ... and the time from the kernel's start to the end of that round, ns
};

struct BatchDev // pointers handed to every kernel
{
    const ImageDesc* images;
    ImageState* states;
    int32_t num_images;
    const uint8_t* raw;
    uint8_t* clean;
    const DevHuffTable* huff;
    const uint16_t* quant; // pool of 64-entry natural-order tables
    uint32_t* seg_start;   // pool
    uint64_t* chunk_counts; // per unstuff chunk: kept bytes | markers << 32  (then exclusive-scanned in place)
    unsigned long long* sub_exit;   // per subsequence: decoder state after its last symbol
    unsigned long long* sub_entry;  // state the last decode of that subsequence started from
    uint32_t* sub_rec;      // coefficient slots (64 per block) consumed by that decode; bit31: absolute (a restart boundary was crossed)
    uint32_t* sub_abs;      // inclusive scan of sub_rec: absolute slot count at the subsequence's exit
    uint32_t* worklist[2];  // subsequences whose entry changed (global subsequence indices)
    SyncCounters* sync;
    uint32_t* sub_stamp;    // work-list stamps of the synchronisation rounds: the subsequence is an item of work-list round r
                            // iff stamp >= r + 1 (r + 2 once it has also been listed for round r + 1)
    uint32_t* blk_pos;      // per block: bit position of its first AC symbol in the image's unstuffed stream (block pass)
    const ScanDesc* scans;  // scans of the progressive images of the batch
    const uint32_t* prog_images; // batch indices of the progressive images (the launch of k_prog_decode covers a range of this list)
    int32_t dense_coef;     // the IDCT kernel takes dense coefficients from coef[] (multi-scan streams) instead of decoding AC symbols
    int32_t dump_coef;      // debug: the IDCT kernel also writes the dense coefficients it reconstructed to coef[]
    int16_t* coef;          // debug tap (dump_coef): AC coefficients, natural order, 64 per block (slot 0 unused: the DC lives in dc[])
    int16_t* dc;            // per block: DC difference after the block pass, DC value after the DC scan
    uint8_t* planes;
    const uint32_t* cta_map_unstuff; // prefix arrays: first CTA of each image (+ total at [num_images])
    const uint32_t* cta_map_subseq;
    const uint32_t* cta_map_fused;   // fused Huffman kernel
    unsigned long long* fused_rec;   // fused Huffman kernel, per CTA: exit state of its last subsequence (huffman.cuh pack_rel) | (coefficient slots
                                     // consumed up to there | 1 << 31) << 32; zero until the CTA has published it
    const uint32_t* cta_map_dc;      // DC kernel
    const uint32_t* cta_map_scan;    // k_subseq_scan
    unsigned long long* scan_rec;    // k_subseq_scan, per chunk: absolute slot count at the chunk's end | 1 << 63; zero until published
    unsigned long long* dc_rec;      // DC kernel, two words per chunk: what the image's MCUs up to the chunk's end leave in the three predictors
                                     // (word 0: Y | Cb << 32; word 1: Cr | (1 << 31 there, bit 0: a restart boundary was passed) << 32); zero until published
    const uint32_t* subseq_cta_img;  // batch image index of every subsequence CTA (indexed by global subsequence / kSubseqCta)
    int32_t image_base;              // batch index of images[0] (a launch may cover a group of consecutive images)
    const uint32_t* cta_map_idct;
    const uint32_t* cta_map_color;
    // a launch may cover only a group of consecutive images: images / states / the CTA maps then point at the group's first
    // image (map values stay absolute) and the kernels add these bases to blockIdx.x
    uint32_t base_unstuff, base_subseq, base_idct, base_color, base_fused, base_dc, base_scan;
    int32_t subseq_bits;   // bits per subsequence
    uint32_t total_subseq;
    int32_t num_tables;    // Huffman tables in the pool
    int32_t lut_slots;     // LUT slots the decoding kernels reserve in shared memory (>= every image's ntabs)
    int32_t lut_ac_slots;  // AC-table slots the reconstruction kernel reserves (>= every image's nac)
    int32_t stream_words;  // 32-bit words of staged stream per thread (odd, >= subseq_bits / 32 + 3); 0 = read the stream from global memory
    int32_t tail_mode;     // work-list rounds: bit 0 = items dealt lane-major (few decoding lanes per warp once the list is short)
};

} // namespace b200jpeg
