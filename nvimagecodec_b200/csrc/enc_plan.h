// enc_plan.h -- host-side planning of the b200jpeg encoder: geometry, quantisation / Huffman tables and the JFIF header
// bytes.  No CUDA dependency (also compiled into the CPU emulation used by the tests).
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>
#include "enc_types.h"

namespace b200jpeg {

// quality 1..100 -> table `which` (0 luma, 1 chroma), baseline-limited.
// Reference: ljt/jcparam.c jpeg_quality_scaling:132-154, jpeg_add_quant_table:30-63, std tables :75-94.
void enc_build_quant(int quality, int which, EncQuantTable& out);

// Standard Huffman tables as encoder look-ups (ljt/jstdhuff.c:54-144, ljt/jchuff.c:232-312).
void enc_build_huff(EncHuffTables& out);

// Geometry of one image.  subsampling: b200jpegSubsampling_t.  Returns false for parameters outside the encoder's scope.
bool enc_plan_image(int width, int height, int format, int subsampling, int restart_interval, EncImageDesc& im);

// SOI, JFIF APP0, DQT, SOF0, DHT, DRI, SOS exactly as libjpeg-turbo writes them for this configuration
// (ljt/jcmarker.c write_file_header:475, write_frame_header:500, write_scan_header:563).
void enc_write_headers(const EncImageDesc& im, const EncQuantTable& luma, const EncQuantTable& chroma, std::vector<uint8_t>& out);

} // namespace b200jpeg
