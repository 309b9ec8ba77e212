// batch_plan.h -- host-side batch scheduling for the b200jpeg decoder: turns parsed headers + output descriptions into
// the device descriptor tables (ImageDesc, Huffman/quant pools, CTA maps) and the scratch layout.  Pure C++ (no CUDA
// calls) so the same plan drives the kernels and the CPU emulation used in tests.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>
#include "device_types.h"
#include "jpeg_parse.h"

namespace b200jpeg {

struct OutputSpec
{
    void* buffer = nullptr; // device pointer the colour kernel writes to
    int64_t row_stride = 0;
    int format = 4;
    int roi_x = 0, roi_y = 0, roi_w = 0, roi_h = 0; // region of interest inside the image; roi_w == 0: the whole image
    int orient = 0; // ImageDesc::orient
};

struct PlanOptions
{
    int subseq_bytes = 128;
    bool fancy_upsampling = true;
};

struct BatchPlan
{
    std::vector<ImageDesc> images;
    std::vector<DevHuffTable> huff;
    std::vector<uint16_t> quant; // 64 per table
    std::vector<ScanDesc> scans;       // scans of the progressive images
    std::vector<uint32_t> prog_images; // planned indices of the progressive images
    std::vector<uint32_t> map_unstuff, map_subseq, map_idct, map_color, map_fused, map_dc, map_scan; // prefix CTA maps, size n+1
    uint64_t raw_bytes = 0, clean_bytes = 0, coef_elems = 0, plane_bytes = 0;
    uint32_t total_chunks = 0, total_subseq = 0, total_segs = 0;
    uint64_t total_blocks = 0, total_pixels = 0, out_bytes = 0;
    int subseq_bits = 1024;
    int max_sync_depth = 0;
};

void build_dev_table(const HuffSpec& spec, bool is_dc, DevHuffTable& out);

// Resolve UNCHANGED formats; returns 0 if (format, ncomp) is not something this codec writes.
int resolve_format(int requested, int ncomp);

// Append one image to the plan. `parsed` must satisfy is_supported().
void plan_add_image(BatchPlan& plan, const ParsedJpeg& parsed, const OutputSpec& out, const PlanOptions& opt);
void plan_begin(BatchPlan& plan, const PlanOptions& opt);
void plan_finish(BatchPlan& plan);

} // namespace b200jpeg
