// decode_kernels.h -- launch interface of decode_kernels.cu
#pragma once
#include <cuda_runtime_api.h>
#include "device_types.h"

namespace b200jpeg {

struct LaunchDims
{
    int num_images = 0;
    uint32_t unstuff_ctas = 0, subseq_ctas = 0, idct_ctas = 0, color_ctas = 0;
};

// each returns the number of kernels it launched
int launch_unstuff(const BatchDev& b, const LaunchDims& d, cudaStream_t st);
int launch_huff_sync(const BatchDev& b, const LaunchDims& d, cudaStream_t st, int first_round, int num_rounds);
int launch_huff_write(const BatchDev& b, const LaunchDims& d, cudaStream_t st);
int launch_dc(const BatchDev& b, const LaunchDims& d, cudaStream_t st);
int launch_idct(const BatchDev& b, const LaunchDims& d, cudaStream_t st);
int launch_color(const BatchDev& b, const LaunchDims& d, cudaStream_t st);

enum Stage { kStageUnstuff = 0, kStageSync, kStageWrite, kStageDc, kStageIdct, kStageColor, kNumStages };

constexpr int kMaxSyncRounds = 256; // size of the redo counter array

} // namespace b200jpeg
