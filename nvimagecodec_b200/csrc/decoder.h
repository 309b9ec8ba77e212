// decoder.h -- batch decoder object behind b200jpeg_decode_* (host scheduling, staging, launches).
#pragma once
#include <cuda_runtime_api.h>
#include <atomic>
#include <condition_variable>
#include <cstddef>
#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "alloc.h"
#include "batch_plan.h"
#include "decode_kernels.h"

namespace b200jpeg {

struct DecoderOptions
{
    bool fancy_upsampling = true; // reference default (libjpeg_turbo_decoder.cpp:250-276)
    int subseq_bytes = 128;
    int groups_host = 5;  // groups (of doubling size) when outputs go to host memory; 0 = use `groups`
    int copy_threads = 8; // host threads (the caller's included) that copy entropy-coded bytes into pinned staging
    bool stream_priority = false; // group g's stream outranks group g + 1's (measured: no gain for host outputs, and with
                                  // equal priorities the latency-bound kernels of one group hide behind the others' throughput kernels)
    int tail_launches = 1;  // work-list rounds launched as ordinary kernels (a thread per subsequence: the long first list in one pass) before the cooperative kernel takes the rest
    bool tail_spread = true;  // work-list items dealt lane-major: a short list leaves one decoding lane per warp
    int groups = 4;           // the batch is cut into this many groups of images that run on separate streams (pipelining)
    bool stage_stream = true; // stage the unstuffed stream in shared memory in the Huffman kernels (subseq_bytes <= 256)
    bool dump_coef = false; // debug: keep dense DCT coefficients for b200jpeg_debug_fetch(what = 0)
    bool profile = false; // record CUDA events between the kernel stages (b200jpeg_decoder_get_stage_times)
    bool fused = false;     // the entropy-decode stage as ONE kernel (k_huff_fused: synchronisation rounds inside the CTA over a halo, CTAs chained by
                            // look-back, block pass at the end) instead of k_huff_sync x2 + work-list rounds + k_subseq_scan + k_huff_write.
                            // Bit-exact and tested, but slower on B200 (C1: 0.87 ms against 0.74 ms): a CTA holds 57 KB of shared memory through its
                            // latency-bound rounds and the wait for its predecessor, which the pooled work-list rounds of the separate kernels do not
                            // (profiles/r02s_fused_trace.log).  Needs stage_stream.
    bool fast_idct = false; // the reference's fast_idct=1 selects JDCT_IFAST (libjpeg_turbo_decoder.cpp:272-273,352), whose
                            // pixels differ from islow and between libjpeg-turbo builds: not offered, create() refuses it
};

// Parse "name=value" tokens; tokens may be prefixed "b200jpeg_decoder:" or ":"; tokens addressed to another plugin are
// ignored (same grammar as the reference, extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:250-276).
DecoderOptions parse_decoder_options(const char* options, const char* plugin_id);

struct DevBuf
{
    void* p = nullptr;
    size_t cap = 0;
};

struct HostOutput // an output the caller wants in host memory
{
    int image = -1;      // planned image index
    void* host = nullptr;
    int64_t host_stride = 0;
    size_t dev_off = 0;
    size_t dev_pitch = 0; // row pitch of the device-side copy: the host's own layout when its rows are tight
    size_t row_bytes = 0;
    size_t rows = 0;
};

constexpr int kMaxGroups = 8;

// Persistent host threads that fill the pinned staging buffer.  The items of a job are claimed in order; the calling
// thread helps until the prefix it needs (one group of images) is complete, issues that group's upload and launches, and
// comes back for the next prefix while the workers keep copying.
class StagePool
{
  public:
    struct Item
    {
        const uint8_t* src;
        uint8_t* dst;
        size_t len;
    };
    explicit StagePool(int workers);
    ~StagePool();
    void start(const std::vector<Item>* items); // the vector must stay alive until wait_prefix(items->size()) returned
    void wait_prefix(size_t upto);
    void quiesce(); // every worker is done with the previous job

  private:
    bool claim_one();
    void worker();
    std::vector<std::thread> th_;
    std::mutex mu_;
    std::condition_variable cv_;
    uint64_t epoch_ = 0;
    bool stop_ = false;
    const std::vector<Item>* items_ = nullptr;
    std::atomic<size_t> next_{0};
    std::unique_ptr<std::atomic<uint8_t>[]> done_;
    size_t done_cap_ = 0, checked_ = 0;
    std::atomic<uint64_t> acks_{0};
};

struct Group // consecutive planned images that are launched together on one stream
{
    int first = 0, count = 0, tail_grid = 0;
    uint32_t prog_first = 0, prog_count = 0; // the group's progressive images: a range of BatchPlan::prog_images
    BatchDev dev{};
    LaunchDims dims{};
    size_t raw_lo = 0, raw_hi = 0;
    std::vector<int> host_outs; // indices into host_outs_
};

class Decoder
{
  public:
    Decoder(int device_id, const DecoderOptions& opt);
    ~Decoder();
    int init(); // b200jpegStatus_t

    struct OutputDesc
    {
        void* buffer;
        int64_t row_stride;
        int format;
        int on_host;
        int roi_x = 0, roi_y = 0, roi_w = 0, roi_h = 0; // region of interest; roi_w == 0: the whole image
        cudaStream_t stream = nullptr; // with has_stream: the stream this output is ordered on (else the call's stream)
        bool has_stream = false;
        int orient = 0; // ImageDesc::orient
    };

    // two-phase: prepare = parse + plan + H2D of the bitstreams; run = all device work of the prepared batch
    int prepare(int n, const uint8_t* const* data, const size_t* sizes, const OutputDesc* outs, int32_t* statuses);
    int run(cudaStream_t user_stream, int32_t* statuses);
    // one-shot, pipelined over the groups (host staging of group g+1 overlaps the device work of group g)
    int decode(int n, const uint8_t* const* data, const size_t* sizes, const OutputDesc* outs, cudaStream_t user_stream, int32_t* statuses);
    // the same in two steps: submit returns when the host work is done (headers parsed, entropy-coded bytes in pinned staging,
    // every copy and kernel enqueued, the outputs' streams made to wait for them); wait blocks until the device is done and
    // fills in the final statuses.  The caller's bytes are not read after submit returns.
    int submit(int n, const uint8_t* const* data, const size_t* sizes, const OutputDesc* outs, cudaStream_t user_stream, int32_t* statuses);
    int wait(int32_t* statuses);
    bool in_flight() const { return in_flight_; }
    void get_stats(uint64_t* s8) const;
    int get_stage_times(float* ms, int n); // per Stage, last run
    int64_t debug_fetch(int what, int image, int comp, void* dst, size_t cap);
    int device() const { return device_; }
    // memory callbacks for every later allocation; only before the first decode (buffers are returned to where they came from)
    int set_allocators(const Allocators& a);

  private:
    int ensure(DevBuf& b, size_t bytes);
    int ensure_pinned(DevBuf& b, size_t bytes);
    int plan_batch(int n, const uint8_t* const* data, const size_t* sizes, const OutputDesc* outs, int32_t* statuses);
    void stage_start();
    int stage_group(int g);
    int launch_group(int g);
    int begin_run(cudaStream_t user_stream, int32_t* statuses);
    int finish(int32_t* statuses);

    int device_;
    DecoderOptions opt_;
    Allocators alloc_;
    cudaStream_t streams_[kMaxGroups] = {};
    cudaEvent_t ev_meta_ = nullptr;
    std::vector<cudaStream_t> out_streams_;  // distinct streams named by the outputs of the planned batch
    std::vector<cudaStream_t> user_streams_; // the streams of the running batch: out_streams_ (+ the call's stream)
    std::vector<cudaEvent_t> ev_user_;       // one event per user stream: what the caller had in flight on the outputs
    cudaEvent_t ev_done_[kMaxGroups] = {};   // end of each group's device work
    bool any_default_stream_ = false;        // some output of the planned batch names no stream of its own
    bool in_flight_ = false;
    bool trace_ = false;
    cudaEvent_t ev_trace_[1 + 3 * kMaxGroups] = {}; // B200JPEG_TRACE: start, then per group: staged, kernels done, copied back
    cudaEvent_t ev_stage_[kNumStages + 1] = {};
    bool stage_valid_ = false;
    bool ready_ = false;

    BatchPlan plan_;
    std::vector<ParsedJpeg> parsed_;
    std::vector<int> planned_to_batch_; // planned image -> index in the caller's batch
    std::vector<int32_t> pre_status_;   // per batch entry status decided at prepare time
    int batch_size_ = 0;
    std::vector<HostOutput> host_outs_;

    DevBuf d_blkpos_, d_stamp_, d_raw_, d_clean_, d_coef_, d_dc_, d_planes_, d_meta_, d_states_, d_seg_, d_chunks_, d_exit_, d_entry_, d_rec_, d_abs_, d_wl0_,
        d_wl1_, d_sync_, d_out_, d_fusedrec_, d_dcrec_, d_scanrec_;
    DevBuf h_raw_, h_meta_, h_states_;
    BatchDev dev_{};
    LaunchDims dims_{};
    std::vector<Group> groups_;
    const uint8_t* const* data_ = nullptr; // caller's bitstreams during decode()/prepare()
    std::unique_ptr<StagePool> pool_;
    std::vector<StagePool::Item> stage_items_;
    std::vector<size_t> stage_group_end_; // per group: end of its items
    size_t scratch_bytes_ = 0;

    uint64_t stat_launches_ = 0, stat_rounds_ = 0, stat_redone_ = 0;
    int tail_capacity_ = 0, sm_count_ = 0, cfg_lut_slots_ = -1, cfg_stream_words_ = -1, cfg_subseq_bits_ = -1;
    bool use_fused_ = false;
};

} // namespace b200jpeg
