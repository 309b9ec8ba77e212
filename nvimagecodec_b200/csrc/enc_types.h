// enc_types.h -- structures shared by the host side of the b200jpeg encoder (encoder.cpp) and its sm_100a kernels
// (encode_kernels.cu), plus the geometry helpers both use.
#pragma once
#include "device_types.h"

namespace b200jpeg {

constexpr int kEncCta = 128;            // threads per CTA in the FDCT, bit-count and emit kernels: one 8x8 block per thread
constexpr int kEncColorCta = 256;       // colour-conversion kernel: 4 pixels x vs rows per thread
constexpr int kEncMaxBlockBytes = 208;  // upper bound of one block's entropy-coded bits (64 x (16 + 10) bits) in bytes
constexpr int kEncStuffThreads = 256;   // stuffing kernels: 16 stream bytes per thread
constexpr int kEncStuffChunk = kEncStuffThreads * 16;

// Quantisation by reciprocal, per coefficient in natural order (reference: ljt/jcdctmgr.c compute_reciprocal:173-228 with
// the 32-bit DCTELEM of the non-SIMD build, quantize:419-452):  q = ((|x| + corr) * recip) >> (32 + shift)
struct EncQuantTable
{
    uint32_t recip[64];
    uint16_t corr[64];
    uint8_t shift[64];
    uint8_t qval[64]; // the 8-bit quant values, natural order (for the DQT segment / debugging)
    B2J_HD uint32_t get_recip(int n) const { return recip[n]; }
    B2J_HD uint32_t get_corr(int n) const { return corr[n]; }
    B2J_HD uint32_t get_shift(int n) const { return shift[n]; }
};

// Standard Huffman tables in encoder form (reference: ljt/jchuff.c jpeg_make_c_derived_tbl:232-312 applied to the T.81
// Annex K.3 tables of ljt/jstdhuff.c:54-144): entry = code << 8 | length, 0 = symbol not in the table.
struct EncHuffTables
{
    uint32_t dc[2][16];   // [luma|chroma][category]
    uint32_t ac[2][256];  // [luma|chroma][run << 4 | size]
};

struct EncImageDesc
{
    int32_t width, height;
    int32_t ncomp;          // 1 (gray) or 3 (YCbCr)
    int32_t hs, vs;         // luma sampling factors; chroma is 1x1
    int32_t mcus_x, mcus_y;
    int32_t bpm;            // blocks per MCU
    int32_t wib[3], hib[3]; // real (non-dummy) blocks per row / block rows of each component
    int32_t restart_interval;
    uint32_t seg_blocks;    // blocks per restart segment (whole image when there is none)
    uint32_t num_segments;
    uint32_t total_blocks;
    int32_t in_fmt;         // B200JPEG_FMT_I_RGB / P_RGB / P_Y
    const uint8_t* pixels;  // device memory
    int64_t stride;         // bytes per row; planar: plane p at pixels + p * stride * height
    int32_t qt;             // index of the luma table in the quant pool (chroma = qt + 1)
    int32_t mcus_per_tile;  // (CPU emulation of the tests: MCUs handled together, a run inside one MCU row)
    int32_t tiles_per_row;
    uint32_t color_cta_first; // CTA prefix of this image in the colour-conversion grid
    int32_t plane_pitch[3];   // Y / Cb / Cr planes after colour conversion + downsampling, padded to whole MCUs:
    uint64_t plane_off[3];    // bytes per row (multiple of 16) and offset from the plane buffer base
    uint32_t emit_cta_first; // CTA prefix in the emit grid (ceil(total_blocks / kEncCta) CTAs)
    uint64_t blk_first;     // first block of this image in the batch-wide block arrays (coef = blk * 64 int16)
    uint32_t seg_first;     // first segment in the batch-wide segment arrays
};

// Filled in after the bit-length scan (second upload): where the image's streams live.
struct EncImageDyn
{
    uint64_t stream_off;    // unstuffed entropy-coded stream, bytes from the stream buffer base (16-byte aligned)
    uint32_t stream_bytes;
    uint32_t chunk_first;   // first stuffing chunk (CTA) of this image
    uint64_t out_off;       // stuffed stream incl. RSTn markers, bytes from the output buffer base
};

struct EncBatchDev
{
    const EncImageDesc* images;
    const EncImageDyn* dyn;
    int32_t num_images;
    const EncQuantTable* quant;
    const EncHuffTables* huff;
    uint8_t* planes;          // component planes (EncImageDesc::plane_off / plane_pitch)
    int16_t* coef;            // quantised coefficients, zigzag order, 64 per block, MCU scan order
    int16_t* dc;              // per block: quantised DC
    uint16_t* acbits;         // per block: entropy-coded bits of the AC part
    uint64_t* nzmask;         // per block: bit k = the coefficient at zigzag position k is non-zero
    uint32_t* blk_off;        // per block: bit offset inside its tile (the kEncCta blocks of one emit CTA)
    uint32_t* tile_off;       // per emit CTA: bits of the image's blocks before the tile
    uint32_t* img_bits;       // per image: bits of all blocks
    uint32_t* seg_byte_off;   // per segment: byte offset inside the image's unstuffed stream
    uint32_t* img_stream_bytes; // per image
    uint8_t* stream;          // unstuffed streams
    uint32_t* chunk_out;      // per stuffing chunk: output bytes, then exclusive prefix inside the image
    uint32_t* img_out_bytes;  // per image: stuffed length
    uint8_t* out;             // stuffed streams
    // A launch covers a group of consecutive images: images / dyn / the per-image arrays point at the group's first image,
    // the CTA maps hold batch-wide prefixes and the launch's first CTA is cta_base_* (stuffing chunks: group-local).
    const uint32_t* cta_map_color; // prefix arrays [num_images + 1]
    const uint32_t* cta_map_emit;  // tiles of kEncCta blocks: the FDCT, bit-count and emit kernels
    const uint32_t* cta_map_stuff;
    uint32_t cta_base_color, cta_base_emit;
};

// component of block j inside the MCU, its first block index in the MCU and its block count
B2J_HD void enc_block_comp(const EncImageDesc& im, int j, int& comp, int& first_j, int& count)
{
    const int ny = im.ncomp == 1 ? 1 : im.hs * im.vs;
    if (j < ny) {
        comp = 0;
        first_j = 0;
        count = ny;
    } else {
        comp = j - ny + 1;
        first_j = j;
        count = 1;
    }
}

// The block whose DC predicts block b's (previous block of the same component in scan order), or -1 when the
// prediction is 0 (first block of the component in the image or in its restart segment).
// Reference: ljt/jchuff.c encode_mcu_huff:688-760 (last_dc_val per component, reset by emit_restart:663).
B2J_HD int64_t enc_dc_pred_block(const EncImageDesc& im, uint32_t b)
{
    const uint32_t mcu = b / uint32_t(im.bpm);
    const int j = int(b - mcu * uint32_t(im.bpm));
    int comp, first_j, count;
    enc_block_comp(im, j, comp, first_j, count);
    if (j > first_j)
        return int64_t(b) - 1;
    if (mcu == 0 || (im.restart_interval && mcu % uint32_t(im.restart_interval) == 0))
        return -1;
    return int64_t(mcu - 1) * im.bpm + first_j + count - 1;
}

} // namespace b200jpeg
