/*
 * host_common.cuh -- small host-side helpers shared by the boundary layer (bsw_api.cu) and the
 * staged-call engine (engine.cuh): error convention, grow-only device / pinned buffers, env knobs.
 *
 * Error convention = the reference's (R:docs-archive/WEEK1_WRAPPER_COMPLETE.md:239-242): print to
 * stderr and exit(EXIT_FAILURE).  There is no CPU fallback anywhere in this library.
 */
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#define CK(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            fprintf(stderr, "[bsw-b200] CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_),  \
                    __FILE__, __LINE__, cudaGetErrorString(e_));                              \
            exit(EXIT_FAILURE);                                                               \
        }                                                                                     \
    } while (0)

namespace bswh {

[[noreturn]] inline void die(const char *msg)
{
    fprintf(stderr, "[bsw-b200] fatal: %s\n", msg);
    exit(EXIT_FAILURE);
}

inline int env_int(const char *name, int dflt)
{
    const char *s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}

/* grow-only device buffer.  Growth frees and reallocates (cudaFree synchronises the device), so callers size
 * generously up front; `slack` = extra fraction added on growth. */
struct DevBuf {
    void *p = nullptr; size_t cap = 0;
    void reserve(size_t bytes)
    {
        if (bytes <= cap) return;
        if (p) CK(cudaFree(p));
        size_t want = bytes + bytes / 4 + 256;
        CK(cudaMalloc(&p, want)); cap = want;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

/* grow-only pinned host buffer */
struct PinBuf {
    void *p = nullptr; size_t cap = 0;
    void reserve(size_t bytes)
    {
        if (bytes <= cap) return;
        if (p) CK(cudaFreeHost(p));
        size_t want = bytes + bytes / 4 + 256;
        CK(cudaHostAlloc(&p, want, cudaHostAllocPortable)); cap = want;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

} /* namespace bswh */
