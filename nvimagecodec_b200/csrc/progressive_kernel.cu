// progressive_kernel.cu -- P1 k_prog_decode: the scans of the progressive images of a batch -> dense coefficients.
// One warp per image; lane 0 walks the image's scans in file order (progressive.cuh: the refinement scans are serial by
// nature, symbol by symbol), the other lanes help where the work is plain memory traffic (clearing the coefficients).
// The images of the batch run side by side; the reconstruction kernel then reads the coefficients (ImageDesc::progressive).
#include <cuda_runtime.h>
#include "decode_kernels.h"
#include "huffman.cuh"
#include "progressive.cuh"

namespace b200jpeg {

__global__ void __launch_bounds__(32) k_prog_decode(BatchDev b, uint32_t first, uint32_t count)
{
    const uint32_t j = blockIdx.x;
    if (j >= count)
        return;
    const int img = int(b.prog_images[first + j]) - b.image_base;
    const ImageDesc& im = b.images[img];
    int16_t* coef = b.coef + im.coef_off;
    { // zero the image's coefficients: 128 bytes per block, 16 bytes per lane per step
        uint4* p = reinterpret_cast<uint4*>(coef);
        const uint64_t n16 = uint64_t(im.total_blocks) * 8;
        for (uint64_t i = threadIdx.x; i < n16; i += 32)
            p[i] = make_uint4(0, 0, 0, 0);
    }
    __syncwarp();
    if (threadIdx.x != 0)
        return;
    uint32_t err = 0;
    for (uint32_t s = 0; s < im.scan_count && !err; ++s)
        err = prog_decode_scan(im, b.scans[im.scan_first + s], b.raw + im.raw_off, b.huff, coef, c_zigzag);
    if (err)
        atomicOr(&b.states[img].error, err);
    else
        b.states[img].done_blocks = im.total_blocks;
}

int launch_progressive(const BatchDev& b, uint32_t first, uint32_t count, cudaStream_t st)
{
    if (count == 0)
        return 0;
    k_prog_decode<<<count, 32, 0, st>>>(b, first, count);
    return 1;
}

} // namespace b200jpeg
