// ofx_common.h -- error plumbing shared by the translation units of libofx.so.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cuda_runtime.h>

namespace ofx {

char *error_buffer();   // thread-local, 512 bytes (defined in ofx_volume.cu)

inline int fail(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(error_buffer(), 512, fmt, ap);
    va_end(ap);
    return 1;
}

}  // namespace ofx

#define OFX_CUDA(expr)                                                                         \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return ofx::fail("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
    } while (0)

#define OFX_LAUNCH_CHECK()                                                                     \
    do {                                                                                       \
        cudaError_t _e = cudaGetLastError();                                                   \
        if (_e != cudaSuccess)                                                                 \
            return ofx::fail("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
    } while (0)

#define OFX_REQUIRE(cond, ...)                                                                 \
    do {                                                                                       \
        if (!(cond)) return ofx::fail(__VA_ARGS__);                                            \
    } while (0)
