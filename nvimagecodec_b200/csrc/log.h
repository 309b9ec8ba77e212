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

} // namespace b200jpeg
