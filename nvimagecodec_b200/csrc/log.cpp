// log.cpp -- see log.h
#include "log.h"
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <nvtx3/nvToolsExt.h>

namespace b200jpeg {

namespace {
std::mutex g_mu;
LogFn g_fn = nullptr;
void* g_ctx = nullptr;
} // namespace

void set_log_callback(LogFn fn, void* ctx)
{
    std::lock_guard<std::mutex> lk(g_mu);
    g_fn = fn;
    g_ctx = ctx;
}

void log_message(int severity, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    std::vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    LogFn fn;
    void* ctx;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        fn = g_fn;
        ctx = g_ctx;
    }
    if (fn)
        fn(ctx, severity, buf);
    else
        std::fprintf(stderr, "b200jpeg: %s\n", buf);
}

bool trace_enabled()
{
    static const bool on = std::getenv("B200JPEG_TRACE") != nullptr;
    return on;
}

NvtxRange::NvtxRange(const char* name) { nvtxRangePushA(name); }
NvtxRange::~NvtxRange() { nvtxRangePop(); }

} // namespace b200jpeg
