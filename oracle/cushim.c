/* oracle/cushim.c -- TEST INFRASTRUCTURE.
 * The reference core calls three CUDA *driver* entry points from src/device_guard.cpp:25,36,54
 * through lazily generated dlopen stubs (src/dynlink/dynlink_cuda.cpp, generated with python-libclang
 * which this image lacks).  This shim provides the same three symbols by forwarding to libcuda.so.1
 * when a driver exists, and reports "no driver" otherwise (never reached with DEVICE_CPU_ONLY). */
#include <cuda.h>
#include <dlfcn.h>
#include <stddef.h>

static void* drv(void)
{
    static void* h = NULL;
    static int tried = 0;
    if (!tried) {
        tried = 1;
        h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
        if (!h)
            h = dlopen("libcuda.so", RTLD_NOW | RTLD_GLOBAL);
    }
    return h;
}

#define FWD(name, proto, args)                                         \
    CUresult name proto                                                \
    {                                                                  \
        typedef CUresult(*fn_t) proto;                                 \
        void* h = drv();                                               \
        fn_t f = h ? (fn_t)dlsym(h, #name) : NULL;                     \
        if (!f)                                                        \
            return CUDA_ERROR_SHARED_OBJECT_SYMBOL_NOT_FOUND;          \
        return f args;                                                 \
    }

#undef cuInit
#undef cuCtxGetCurrent
#undef cuCtxSetCurrent
FWD(cuInit, (unsigned int flags), (flags))
FWD(cuCtxGetCurrent, (CUcontext * pctx), (pctx))
FWD(cuCtxSetCurrent, (CUcontext ctx), (ctx))

/* dynlink_cuda.h also declares these helpers (src/dynlink/dynlink_cuda.cpp:62-75). */
