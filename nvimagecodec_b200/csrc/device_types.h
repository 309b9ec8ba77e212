// device_types.h -- structures shared by the host scheduler and the sm_100a kernels of b200jpeg.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define B2J_HD __host__ __device__ __forceinline__
#else
#define B2J_HD inline
#endif

namespace b200jpeg {

constexpr int kLutBits = 10;          // first-level Huffman lookup width
constexpr int kSubseqCta = 128;       // threads per CTA in the subsequence kernels (one subsequence per thread)
constexpr int kUnstuffThreads = 256;  // threads per CTA in the unstuffing kernels
constexpr int kUnstuffBytesPerThread = 16;
constexpr int kChunkBytes = kUnstuffThreads * kUnstuffBytesPerThread; // raw bytes per unstuff CTA
constexpr int kIdctCta = 128;        // threads per CTA, one 8x8 block per thread
constexpr int kColorCta = 256;       // threads per CTA, 16 pixels per thread
constexpr int kDcCta = 512;          // threads per CTA, one MCU per thread per iteration
constexpr int kMaxBlocksPerMcu = 6;   // luma up to 2x2 + Cb + Cr

// Decode table derived from one DHT table (reference: d_derived_tbl, external/libjpeg-turbo/jdhuff.h:27-51 and
// jpeg_make_d_derived_tbl, jdhuff.c:143-285; we use a 10-bit first level instead of 8 -- same symbols come out).
struct DevHuffTable
{
    uint16_t lut[1 << kLutBits]; // (code length << 8) | symbol ; 0 => code longer than kLutBits (or invalid)
    int32_t maxcode[18];         // largest code of each length, -1 if none
    int32_t valoffset[18];       // vals[] index = code + valoffset[len]
    uint8_t vals[256];
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
    int32_t mcu_slots;    // bpm * 64
    int32_t restart_interval; // MCUs per restart segment (0 => single segment)
    int32_t num_segments;
    uint32_t seg_slots;   // coefficient slots per full segment (restart_interval * mcu_slots, or total)
    uint32_t total_slots; // mcus * mcu_slots
    int32_t wib[3], hib[3];   // MCU-padded block grid per component
    int32_t ds_w[3], ds_h[3]; // downsampled_width / height (jdinput.c:130-135)
    int32_t blk_first[3];     // first block-in-MCU index of each component
    uint8_t blk_comp[8];      // component of block j in the MCU
    int32_t qt[3];            // quant table pool index per component
    int32_t dc_tab[3], ac_tab[3]; // Huffman table pool index per component

    // entropy-coded bytes
    uint64_t raw_off;     // byte offset in the batch raw buffer (16-byte aligned)
    uint32_t raw_len;
    uint64_t clean_off;   // byte offset in the unstuffed buffer (16-byte aligned)
    uint32_t chunk_first, chunk_count;   // unstuff CTAs of this image
    uint32_t subseq_first, subseq_count; // subsequence slots (multiple of kSubseqCta) / used slots
    uint32_t subseq_cta_first, subseq_cta_count;
    uint32_t seg_first;   // index into the segment-start pool; entries [seg_first, seg_first + num_segments]

    // intermediates
    uint64_t coef_off;        // int16 units, blocks in MCU scan order
    uint64_t plane_off[3];    // bytes
    int32_t plane_pitch[3];
    uint32_t idct_cta_first, idct_cta_count;
    uint32_t color_cta_first, color_cta_count;
    uint32_t dc_cta;          // unused for now

    // output
    uint8_t* out;
    int64_t out_stride;
    int32_t out_fmt;      // b200jpegFormat_t resolved: P_RGB/I_RGB/P_BGR/I_BGR/P_Y
    int32_t fancy;
};

// Per-image state written by the device.
struct ImageState
{
    uint32_t clean_len;     // unstuffed bytes
    uint32_t marker_count;  // RSTn markers seen
    uint32_t error;         // bit flags, see kErr*
    uint32_t done_slots;    // set by the thread that completes the last block (== total_slots when the scan is complete)
};

constexpr uint32_t kErrBadMarker = 1;     // unexpected marker inside entropy-coded data / wrong RSTn order / count
constexpr uint32_t kErrBadCode = 2;       // invalid Huffman code on the synchronised path
constexpr uint32_t kErrNotConverged = 4;  // subsequence synchronisation incomplete (host re-runs more rounds)
constexpr uint32_t kErrOverrun = 8;       // decode ran past the end of a segment / image

// Subsequence record after a sync round.
struct SubseqRec
{
    uint32_t n_abs; // bit31: value is absolute (relative to image start), else relative to the entry; low 31 bits: slots
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
    unsigned long long* sub_exit;   // per subsequence: p | slot << 32
    unsigned long long* sub_entry;  // entry used by the last decode of that subsequence
    uint32_t* sub_rec;      // SubseqRec
    uint32_t* sub_abs;      // scanned absolute exit slot
    uint32_t* redo_counter; // [round]
    int16_t* coef;
    uint8_t* planes;
    const uint32_t* cta_map_unstuff; // prefix arrays: first CTA of each image (+ total at [num_images])
    const uint32_t* cta_map_subseq;
    const uint32_t* cta_map_idct;
    const uint32_t* cta_map_color;
    const uint32_t* cta_map_dc;
    int32_t subseq_bits;   // bits per subsequence
};

} // namespace b200jpeg
