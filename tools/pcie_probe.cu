// tools/pcie_probe.cu -- measures what bounds the host<->device side of the boundary calls:
// contiguous vs strided (24 of 56 bytes) DMA, zero-copy kernel access to mapped pinned memory,
// pageable uploads, small round-trip latency, and the per-GPU / concurrent H2D + D2H matrix
// (VERDICT r01 tasks 3 and 4).  Prints one JSON object.  Build: see tools/Makefile.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <thread>
#include <atomic>
#include <chrono>
#include <string>
#include <algorithm>
#include <functional>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void zc_write(int2 *rec, int n)          // 24 result bytes per 56-byte record, straight into mapped host memory
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int2 *o = (int2 *)((char *)rec + (size_t)i * 56 + 32);
    o[0] = make_int2(i, 1); o[1] = make_int2(i, 2); o[2] = make_int2(i, 3);
}
__global__ void zc_read(const int2 *rec, int n, int2 *out)   // 24 input bytes per record from mapped host memory
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int2 *p = (const int2 *)((const char *)rec + (size_t)i * 56);
    int2 a = p[0], b = p[1], c = p[2];
    out[i] = make_int2(a.x + b.x + c.x, a.y + b.y + c.y);
}
__global__ void zc_read_bytes(const uint4 *src, uint4 *dst, size_t n16)   // contiguous 16-byte loads from mapped host memory
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
    for (; i < n16; i += st) dst[i] = src[i];
}
__global__ void tiny(int *p) { if (threadIdx.x == 0) p[0]++; }

struct Dev {
    int d; char *h = nullptr, *hp = nullptr, *dv = nullptr, *dc = nullptr; cudaStream_t s, s2; cudaEvent_t e0, e1;
};

static double timed(Dev &x, const std::function<void()> &fn, int reps = 3)
{
    double best = 1e30;
    for (int r = 0; r < reps; ++r) {
        CK(cudaDeviceSynchronize());
        auto t0 = std::chrono::steady_clock::now();
        fn();
        CK(cudaDeviceSynchronize());
        double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        best = std::min(best, s);
    }
    (void)x;
    return best;
}

int main(int argc, char **argv)
{
    int ngpu = 1; long n = 4000000;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--gpus") && i + 1 < argc) ngpu = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--pairs") && i + 1 < argc) n = atol(argv[++i]);
    }
    int have = 0; CK(cudaGetDeviceCount(&have));
    ngpu = std::min(ngpu, have);
    const size_t rec = 56, bytes = (size_t)n * rec, cbytes = (size_t)n * 24;
    std::vector<Dev> dv(ngpu);
    for (int d = 0; d < ngpu; ++d) {
        Dev &x = dv[d]; x.d = d;
        CK(cudaSetDevice(d));
        CK(cudaHostAlloc((void **)&x.h, bytes, cudaHostAllocPortable | cudaHostAllocMapped));
        memset(x.h, 1, bytes);
        x.hp = (char *)malloc(bytes); memset(x.hp, 2, bytes);
        CK(cudaMalloc((void **)&x.dv, bytes)); CK(cudaMalloc((void **)&x.dc, bytes));
        CK(cudaStreamCreateWithFlags(&x.s, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&x.s2, cudaStreamNonBlocking));
        CK(cudaEventCreate(&x.e0)); CK(cudaEventCreate(&x.e1));
    }
    printf("{\"pairs\": %ld, \"record_bytes\": 56, \"gpus\": %d", n, ngpu);
    {   // ---- single-GPU tests on device 0
        Dev &x = dv[0]; CK(cudaSetDevice(0));
        const double GB = 1e9;
        double t;
        t = timed(x, [&] { CK(cudaMemcpyAsync(x.dv, x.h, bytes, cudaMemcpyHostToDevice, x.s)); });
        printf(", \"h2d_contig_GBs\": %.2f", bytes / t / GB);
        t = timed(x, [&] { CK(cudaMemcpyAsync(x.h, x.dv, bytes, cudaMemcpyDeviceToHost, x.s)); });
        printf(", \"d2h_contig_GBs\": %.2f", bytes / t / GB);
        t = timed(x, [&] { CK(cudaMemcpyAsync(x.dv, x.h, bytes, cudaMemcpyHostToDevice, x.s)); CK(cudaMemcpyAsync(x.h, x.dc, bytes, cudaMemcpyDeviceToHost, x.s2)); });
        printf(", \"duplex_each_GBs\": %.2f", bytes / t / GB);
        t = timed(x, [&] { CK(cudaMemcpy2DAsync(x.dc, 24, x.h, 56, 24, n, cudaMemcpyHostToDevice, x.s)); });
        printf(", \"h2d_2d_24of56_payload_GBs\": %.2f, \"h2d_2d_ms\": %.3f", cbytes / t / GB, t * 1e3);
        t = timed(x, [&] { CK(cudaMemcpy2DAsync(x.h + 32, 56, x.dc, 24, 24, n, cudaMemcpyDeviceToHost, x.s)); });
        printf(", \"d2h_2d_24to56_payload_GBs\": %.2f, \"d2h_2d_ms\": %.3f", cbytes / t / GB, t * 1e3);
        t = timed(x, [&] { CK(cudaMemcpy2DAsync(x.dc, 32, x.h, 56, 32, n, cudaMemcpyHostToDevice, x.s)); });
        printf(", \"h2d_2d_32of56_payload_GBs\": %.2f", (double)n * 32 / t / GB);
        char *hm = nullptr; CK(cudaHostGetDevicePointer((void **)&hm, x.h, 0));
        t = timed(x, [&] { zc_write<<<(n + 255) / 256, 256, 0, x.s>>>((int2 *)hm, (int)n); });
        printf(", \"zerocopy_write_24of56_payload_GBs\": %.2f, \"zc_write_ms\": %.3f", cbytes / t / GB, t * 1e3);
        t = timed(x, [&] { zc_read<<<(n + 255) / 256, 256, 0, x.s>>>((const int2 *)hm, (int)n, (int2 *)x.dc); });
        printf(", \"zerocopy_read_24of56_payload_GBs\": %.2f, \"zc_read_ms\": %.3f", cbytes / t / GB, t * 1e3);
        t = timed(x, [&] { zc_read_bytes<<<148 * 8, 256, 0, x.s>>>((const uint4 *)hm, (uint4 *)x.dc, bytes / 16); });
        printf(", \"zerocopy_read_contig_GBs\": %.2f", bytes / t / GB);
        t = timed(x, [&] { CK(cudaMemcpyAsync(x.dv, x.hp, bytes, cudaMemcpyHostToDevice, x.s)); }, 2);
        printf(", \"h2d_pageable_GBs\": %.2f", bytes / t / GB);
        t = timed(x, [&] { CK(cudaMemcpyAsync(x.hp, x.dv, bytes, cudaMemcpyDeviceToHost, x.s)); }, 2);
        printf(", \"d2h_pageable_GBs\": %.2f", bytes / t / GB);
        t = timed(x, [&] { memcpy(x.h, x.hp, bytes); }, 2);
        printf(", \"host_memcpy_1thread_GBs\": %.2f", bytes / t / GB);
        // small round trips: 2.5k records up, tiny kernel, 2.5k x 24 B down, stream sync each
        for (int m : { 2500, 40000 }) {
            const int reps = 200;
            auto t0 = std::chrono::steady_clock::now();
            for (int r = 0; r < reps; ++r) {
                CK(cudaMemcpyAsync(x.dv, x.h, (size_t)m * 56, cudaMemcpyHostToDevice, x.s));
                tiny<<<1, 32, 0, x.s>>>((int *)x.dc);
                CK(cudaMemcpyAsync(x.h, x.dc, (size_t)m * 24, cudaMemcpyDeviceToHost, x.s));
                CK(cudaStreamSynchronize(x.s));
            }
            double us = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / reps * 1e6;
            printf(", \"roundtrip_%d_pairs_us\": %.1f", m, us);
        }
        {   // launch + sync latency of an empty kernel
            const int reps = 500;
            auto t0 = std::chrono::steady_clock::now();
            for (int r = 0; r < reps; ++r) { tiny<<<1, 32, 0, x.s>>>((int *)x.dc); CK(cudaStreamSynchronize(x.s)); }
            double us = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / reps * 1e6;
            printf(", \"launch_sync_us\": %.1f", us);
        }
        {   // cost of creating what one r01 handle creates: 116 streams + 8 pinned allocations
            auto t0 = std::chrono::steady_clock::now();
            std::vector<cudaStream_t> ss(116); std::vector<void *> pp(8);
            for (auto &s : ss) CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
            for (auto &p : pp) CK(cudaHostAlloc(&p, 256, cudaHostAllocDefault));
            double mk = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() * 1e6;
            t0 = std::chrono::steady_clock::now();
            for (auto &s : ss) CK(cudaStreamDestroy(s));
            for (auto &p : pp) CK(cudaFreeHost(p));
            double rm = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() * 1e6;
            printf(", \"r01_handle_resources_create_us\": %.0f, \"destroy_us\": %.0f", mk, rm);
        }
    }
    // ---- concurrency matrix: k GPUs copying at once (one host thread each), H2D alone, D2H alone, both
    printf(", \"matrix\": [");
    bool first = true;
    for (int k = 1; k <= ngpu; k *= 2) {
        for (int mode = 0; mode < 3; ++mode) {
            std::vector<double> secs(k, 0.0);
            std::atomic<int> ready(0);
            std::vector<std::thread> th;
            for (int d = 0; d < k; ++d)
                th.emplace_back([&, d] {
                    Dev &x = dv[d]; CK(cudaSetDevice(d));
                    ready++; while (ready.load() < k) { }
                    const int reps = 4;
                    auto t0 = std::chrono::steady_clock::now();
                    for (int r = 0; r < reps; ++r) {
                        if (mode != 1) CK(cudaMemcpyAsync(x.dv, x.h, bytes, cudaMemcpyHostToDevice, x.s));
                        if (mode != 0) CK(cudaMemcpyAsync(x.h, x.dc, bytes, cudaMemcpyDeviceToHost, x.s2));
                    }
                    CK(cudaDeviceSynchronize());
                    secs[d] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / reps;
                });
            for (auto &t : th) t.join();
            double worst = *std::max_element(secs.begin(), secs.end());
            printf("%s{\"gpus\": %d, \"mode\": \"%s\", \"per_gpu_GBs_each_dir\": %.2f, \"aggregate_GBs_each_dir\": %.2f}", first ? "" : ", ",
                   k, mode == 0 ? "h2d" : mode == 1 ? "d2h" : "duplex", bytes / worst / 1e9, k * (double)bytes / worst / 1e9);
            first = false;
        }
    }
    printf("]}\n");
    return 0;
}
