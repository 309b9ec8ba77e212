// log.h -- where the library's diagnostics go: stderr, or the callback the host application installs
// (b200jpeg_set_log_callback; the nvImageCodec extension forwards to the framework's logger, reference
// include/nvimgcodec.h nvimgcodecFrameworkDesc_t::log).
#pragma once
#include <cstdarg>

namespace b200jpeg {

enum LogSeverity { kLogInfo = 0, kLogWarning = 1, kLogError = 2 };
typedef void (*LogFn)(void* ctx, int severity, const char* message);
void set_log_callback(LogFn fn, void* ctx);
void log_message(int severity, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
bool trace_enabled(); // B200JPEG_TRACE in the environment (read once)

// NVTX range over a host-side stage (plan / stage / launch / wait ...), the way the reference marks its own
// (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:180,464, extensions/nvjpeg/hw_decoder.cpp:192,639): lines up with the
// kernels in an nsys timeline; costs nothing when no tool is attached.
struct NvtxRange
{
    explicit NvtxRange(const char* name);
    ~NvtxRange();
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

} // namespace b200jpeg
