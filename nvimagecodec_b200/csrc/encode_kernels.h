// encode_kernels.h -- launch interface of encode_kernels.cu
#pragma once
#include <cuda_runtime_api.h>
#include "enc_types.h"

namespace b200jpeg {

struct EncLaunchDims
{
    int num_images = 0;
    uint32_t color_ctas = 0, emit_ctas = 0, stuff_ctas = 0;
};

// each returns the number of kernels it launched
int launch_enc_dct(const EncBatchDev& b, const EncLaunchDims& d, cudaStream_t st);   // colour + downsample, then FDCT + quantise
int launch_enc_scan(const EncBatchDev& b, const EncLaunchDims& d, cudaStream_t st);  // bit offsets of blocks and segments
int launch_enc_emit(const EncBatchDev& b, const EncLaunchDims& d, cudaStream_t st);  // Huffman coding into the unstuffed stream
int launch_enc_stuff(const EncBatchDev& b, const EncLaunchDims& d, cudaStream_t st); // byte stuffing + RSTn markers

} // namespace b200jpeg
