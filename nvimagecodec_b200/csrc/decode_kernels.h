// decode_kernels.h -- launch interface of decode_kernels.cu
#pragma once
#include <cuda_runtime_api.h>
#include "device_types.h"

namespace b200jpeg {

struct LaunchDims
{
    int num_images = 0;
    uint32_t unstuff_ctas = 0, subseq_ctas = 0, idct_ctas = 0, color_ctas = 0, fused_ctas = 0, dc_ctas = 0, scan_ctas = 0;
};

// each returns the number of kernels it launched
int launch_unstuff(const BatchDev& b, const LaunchDims& d, cudaStream_t st);
// rounds 0 and 1 over every subsequence
int launch_huff_sync(const BatchDev& b, const LaunchDims& d, cudaStream_t st);
// the cooperative work-list kernel (grid size from huff_configure); -1 on launch failure
int launch_huff_tail(const BatchDev& b, const LaunchDims& d, cudaStream_t st, int tail_grid, int plain_rounds);
// set the dynamic shared memory limits for b.lut_slots / b.stream_words and size the cooperative tail grid; -1 on error
int huff_configure(const BatchDev& b, int device, int* tail_grid);
int launch_huff_write(const BatchDev& b, const LaunchDims& d, cudaStream_t st);
// the whole entropy-decode stage in one launch (k_huff_fused); needs the stream staged in shared memory (b.stream_words != 0),
// b.sync->ticket == 0 and the group's words of b.fused_rec zeroed
int launch_huff_fused(const BatchDev& b, const LaunchDims& d, cudaStream_t st);
int fused_configure(const BatchDev& b); // dynamic shared memory limit of k_huff_fused; -1 on error
// needs b.sync->ticket_dc == 0 and the group's words of b.dc_rec zeroed
int launch_dc(const BatchDev& b, const LaunchDims& d, cudaStream_t st);
int launch_idct(const BatchDev& b, const LaunchDims& d, cudaStream_t st);
// the scans of progressive images [first, first + count) of b.prog_images -> dense coefficients
int launch_progressive(const BatchDev& b, uint32_t first, uint32_t count, cudaStream_t st);
int idct_configure(); // dynamic shared memory limit of k_idct; -1 on error
int launch_color(const BatchDev& b, const LaunchDims& d, cudaStream_t st);

enum Stage { kStageUnstuff = 0, kStageSync, kStageTail, kStageWrite, kStageDc, kStageIdct, kStageColor, kNumStages };


} // namespace b200jpeg
