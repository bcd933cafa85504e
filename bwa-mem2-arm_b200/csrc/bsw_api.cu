/*
 * bsw_api.cu -- extern "C" boundary of the B200 BSW / kswv path (include/bsw_cuda.h).
 *
 * Host side of the drop-in replacement for BandedPairWiseSW::getScores8/16 and
 * kswv::getScores8/16 (bwa-mem2 src/bandedSWA.cpp, src/kswv.cpp; call sites in
 * mem_chain2aln_across_reads_V2 / mem_sam_pe_batch, SURVEY.md sections 3, 8b).
 * One handle per calling host thread (as the reference constructs one object
 * per worker); a handle owns, per device, four chunk slots (stream, grow-only
 * device scratch, pinned mailbox) and one priority stream per query-length bin.  No CPU fallback: any CUDA failure is fatal, as in the
 * reference (fprintf + exit).
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <string>
#include <algorithm>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <functional>
#include <cub/device/device_radix_sort.cuh>
#include "bsw_kernels.cuh"

using namespace bswk;

#define CK(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            fprintf(stderr, "[bsw-b200] CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_),  \
                    __FILE__, __LINE__, cudaGetErrorString(e_));                              \
            exit(EXIT_FAILURE);                                                               \
        }                                                                                     \
    } while (0)

static void die(const char *msg)
{
    fprintf(stderr, "[bsw-b200] fatal: %s\n", msg);
    exit(EXIT_FAILURE);
}

namespace {

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

/* grow-only pinned host buffer (staging of the gathered sequences) */
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

/* helper threads of one handle for the host-side gather (the reference's getScores has a numThreads argument too) */
struct Pool {
    std::vector<std::thread> th;
    std::mutex m;
    std::condition_variable cv, cv_done;
    std::function<void(int)> job;
    int gen = 0, pending = 0;
    bool stop = false;
    int size() const { return (int)th.size() + 1; }
    void start(int nthreads)
    {
        for (int t = 1; t < nthreads; ++t)
            th.emplace_back([this, t] {
                int seen = 0;
                std::unique_lock<std::mutex> lk(m);
                for (;;) {
                    cv.wait(lk, [&] { return stop || gen != seen; });
                    if (stop) return;
                    seen = gen;
                    lk.unlock(); job(t); lk.lock();
                    if (--pending == 0) cv_done.notify_one();
                }
            });
    }
    void run(const std::function<void(int)> &fn)     /* fn(t) for t = 0 .. size()-1; t = 0 on the calling thread */
    {
        if (th.empty()) { fn(0); return; }
        { std::lock_guard<std::mutex> lk(m); job = fn; pending = (int)th.size(); ++gen; }
        cv.notify_all();
        fn(0);
        std::unique_lock<std::mutex> lk(m);
        cv_done.wait(lk, [&] { return pending == 0; });
    }
    void shutdown()
    {
        { std::lock_guard<std::mutex> lk(m); stop = true; }
        cv.notify_all();
        for (auto &t : th) t.join();
        th.clear();
    }
};

#define BSW_NSLOT 4
#define BSW_NLEVELS 6
struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    DevBuf pairs, ref, qer, aln, keys_a, keys_b, idx_a, idx_b, cubtmp, plan, scratch, rowmax;
    DevBuf px_prev, px_list, px_sub, px_count;   /* device-side caller protocol (bsw_dev_extendChained) */
    int *h_count = nullptr;                       /* pinned mailbox of the retry count */
    PinBuf st_ref, st_qer, st_off;                /* gathered sequences + their offsets (host-pointer BSW calls) */
    DevBuf off;
    Plan *h_plan = nullptr;          /* pinned mailbox */
    bool busy = false;               /* a chunk is in flight on this slot */
    cudaEvent_t dev_done = nullptr;  /* end of the last device-resident call that used this slot's scratch */
    bool dev_used = false;
};

struct DevCtx {
    int dev = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    /* one stream per query-length bin: the bins' kernels run concurrently, so one bin's tail wave
     * overlaps the next bin's blocks (fork/join on events around them) */
    cudaStream_t bin_stream[BSW_NBINS + 1] = {};
    /* BSW: one stream per (priority level, bin).  The levels a call's non-empty bins get are decided per call from its
     * histogram (assign_levels): with fixed per-bin priorities a call used only three of the device's levels, and inside
     * one level the SHORTER bins win every freed shared-memory fragment, so the longest bin of a level ran last and its
     * long-lived warps made the call's tail (bin trace: NE=96 ending at 6.08 ms of a 6.08 ms call, NE=80 at 4.5 ms). */
    cudaStream_t lvl_stream[BSW_NLEVELS][BSW_NBINS + 1] = {};
    int nlevels = 1;
    cudaEvent_t fork_ev = nullptr, bin_ev[BSW_NBINS + 1] = {};
    int sm_count = 0, smem_optin = 0;
    DevBuf cells;
    int dev_rr = 0;                  /* next slot of the device-resident calls (see dev_slot) */
    /* in-flight chunk slots of the host-pointer path: while chunk c computes, chunk c+1 uploads and
     * chunk c-1 downloads (one stream, one set of scratch buffers and one pinned mailbox per slot) */
    Slot slot[BSW_NSLOT];
};

struct Core {
    Const c;
    std::vector<DevCtx> ctx;
    bsw_stats_t st;
    int timing = 0;
    int count_cells = 0;
    int chunk_pairs = 1 << 19;
    int ahead = 2;
    int mode = MODE_BSW;
    /* host-pointer BSW calls: gather the bases the chunk's pairs name into pinned staging with `gather` host threads and
     * upload those, instead of the [min, max) byte range of the caller's buffers (which holds the other tier's bases
     * too: bwa-mem2 passes ONE sequence arena to getScores8 and getScores16).  0 = range upload. */
    int gather = 0;
    Pool *pool = nullptr;
};

int env_int(const char *name, int dflt)
{
    const char *s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}

std::vector<int> pick_devices()
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) die("no CUDA device available (this library has no CPU fallback)");
    std::vector<int> out;
    const char *s = getenv("BSW_CUDA_DEVICES");
    if (s && *s) {
        std::string t(s); size_t p = 0;
        while (p < t.size()) {
            size_t q = t.find(',', p); if (q == std::string::npos) q = t.size();
            int d = atoi(t.substr(p, q - p).c_str());
            if (d < 0 || d >= n) die("BSW_CUDA_DEVICES names a device that does not exist");
            out.push_back(d); p = q + 1;
        }
    }
    if (out.empty()) {
        const char *lr = getenv("LOCAL_RANK");
        int d = (lr && *lr) ? atoi(lr) % n : 0;
        out.push_back(d);
    }
    return out;
}

void core_init(Core &k, int mode)
{
    k.mode = mode;
    memset(&k.st, 0, sizeof k.st);
    k.timing = env_int("BSW_CUDA_TIMING", 0);
    std::vector<int> devs = pick_devices();
    k.ctx.resize(devs.size());
    for (size_t i = 0; i < devs.size(); ++i) {
        DevCtx &x = k.ctx[i];
        x.dev = devs[i];
        CK(cudaSetDevice(x.dev));
        CK(cudaEventCreate(&x.ev0)); CK(cudaEventCreate(&x.ev1));
        {
            int lo = 0, hi = 0;                       /* numerically lowest = highest priority */
            CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
            CK(cudaEventCreateWithFlags(&x.fork_ev, cudaEventDisableTiming));
            for (int qb = 1; qb <= BSW_NBINS; ++qb) {
                int pr = hi + (BSW_NBINS - qb) * (lo - hi) / BSW_NBINS;   /* long queries first */
                CK(cudaStreamCreateWithPriority(&x.bin_stream[qb], cudaStreamNonBlocking, pr));
                CK(cudaEventCreateWithFlags(&x.bin_ev[qb], cudaEventDisableTiming));
            }
            x.nlevels = std::max(1, std::min(BSW_NLEVELS, lo - hi + 1));
            if (mode == MODE_BSW)
                for (int l = 0; l < x.nlevels; ++l)
                    for (int qb = 1; qb <= BSW_NBINS; ++qb)
                        CK(cudaStreamCreateWithPriority(&x.lvl_stream[l][qb], cudaStreamNonBlocking, hi + l));
        }
        CK(cudaDeviceGetAttribute(&x.sm_count, cudaDevAttrMultiProcessorCount, x.dev));
        CK(cudaDeviceGetAttribute(&x.smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, x.dev));
        for (Slot &sl : x.slot) {
            CK(cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&sl.dev_done, cudaEventDisableTiming));
            CK(cudaHostAlloc((void **)&sl.h_plan, sizeof(Plan), cudaHostAllocDefault));
            CK(cudaHostAlloc((void **)&sl.h_count, sizeof(int), cudaHostAllocDefault));
            sl.plan.reserve(sizeof(Plan));
        }
        x.cells.reserve(sizeof(unsigned long long));
        CK(cudaMemset(x.cells.p, 0, sizeof(unsigned long long)));
    }
    k.count_cells = env_int("BSW_CUDA_COUNT_CELLS", 0);
    k.chunk_pairs = env_int("BSW_CUDA_CHUNK_PAIRS", 1 << 19);
    k.ahead = env_int("BSW_CUDA_AHEAD", 2);
    k.gather = env_int("BSW_CUDA_GATHER_THREADS", 0);
    /* packed kernels: one warp per block, dynamic shared memory = NE x 192 B (NE <= 256) */
    const int max_smem = BSW_NBINS * 16 * BSW_WARP_ELEM_BYTES;
    for (DevCtx &x : k.ctx) {
        CK(cudaSetDevice(x.dev));
        if (max_smem > x.smem_optin) die("device shared memory too small for the packed kernels");
        const void *fns[] = { (const void *)bsw_packed_kernel<false, false, false>, (const void *)bsw_packed_kernel<false, true, false>,
                              (const void *)bsw_packed_kernel<false, false, true>, (const void *)bsw_packed_kernel<false, true, true>,
                              (const void *)bsw_packed_kernel<true, false, false>, (const void *)bsw_packed_kernel<true, true, false>,
                              (const void *)kswv_packed_kernel<false>, (const void *)kswv_packed_kernel<true> };
        for (const void *fn : fns) {
            CK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
            CK(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        }
    }
}

void core_destroy(Core &k)
{
    if (k.pool) { k.pool->shutdown(); delete k.pool; k.pool = nullptr; }
    for (DevCtx &x : k.ctx) {
        cudaSetDevice(x.dev);
        for (Slot &sl : x.slot) {
            if (sl.stream) cudaStreamSynchronize(sl.stream);
            sl.pairs.release(); sl.ref.release(); sl.qer.release(); sl.aln.release();
            sl.keys_a.release(); sl.keys_b.release(); sl.idx_a.release(); sl.idx_b.release();
            sl.cubtmp.release(); sl.plan.release(); sl.scratch.release(); sl.rowmax.release();
            sl.st_ref.release(); sl.st_qer.release(); sl.st_off.release(); sl.off.release();
            sl.px_prev.release(); sl.px_list.release(); sl.px_sub.release(); sl.px_count.release();
            if (sl.h_count) cudaFreeHost(sl.h_count);
            if (sl.h_plan) cudaFreeHost(sl.h_plan);
            if (sl.done) cudaEventDestroy(sl.done);
            if (sl.dev_done) cudaEventDestroy(sl.dev_done);
            if (sl.stream) cudaStreamDestroy(sl.stream);
        }
        x.cells.release();
        cudaEventDestroy(x.ev0); cudaEventDestroy(x.ev1);
        if (x.fork_ev) cudaEventDestroy(x.fork_ev);
        for (int qb = 1; qb <= BSW_NBINS; ++qb) { if (x.bin_stream[qb]) cudaStreamDestroy(x.bin_stream[qb]); if (x.bin_ev[qb]) cudaEventDestroy(x.bin_ev[qb]); }
        for (int l = 0; l < BSW_NLEVELS; ++l)
            for (int qb = 1; qb <= BSW_NBINS; ++qb) if (x.lvl_stream[l][qb]) cudaStreamDestroy(x.lvl_stream[l][qb]);
    }
    k.ctx.clear();
}

Const make_const(int o_del, int e_del, int o_ins, int e_ins, int zdrop, int end_bonus, int w_match, int w_mismatch)
{
    Const c;
    if (e_del <= 0 || e_ins <= 0) die("gap extension penalties must be positive");
    if (w_match <= 0) die("w_match must be positive");
    c.o_del = o_del; c.e_del = e_del; c.o_ins = o_ins; c.e_ins = e_ins;
    c.oe_del = o_del + e_del; c.oe_ins = o_ins + e_ins;
    c.zdrop = zdrop; c.end_bonus = end_bonus;
    c.a = w_match; c.b = w_mismatch < 0 ? -w_mismatch : w_mismatch;   /* upstream passes +b, the fork -b */
    c.ambig = -1;                                                      /* bandedSWA.h DEFAULT_AMBIG */
    c.maxsc = w_match;
    return c;
}

/* stages 1-3 + plan read-back; returns the host copy of the plan (valid until the next call) */
void plan_enqueue(Core &k, Slot &x, const bsw_seqpair_t *d_pairs, const uint8_t *d_qer, int n, cudaStream_t s, int tier8_check,
                  long long rlo = 0, long long rhi = (long long)1 << 62, long long qlo = 0, long long qhi = (long long)1 << 62)
{
    x.keys_a.reserve((size_t)n * 4); x.keys_b.reserve((size_t)n * 4);
    x.idx_a.reserve((size_t)n * 4);  x.idx_b.reserve((size_t)n * 4);
    size_t tmp_bytes = 0;
    CK(cub::DeviceRadixSort::SortPairsDescending(nullptr, tmp_bytes, (uint32_t *)nullptr, (uint32_t *)nullptr,
                                                 (uint32_t *)nullptr, (uint32_t *)nullptr, n, 0, BSW_KEY_BITS, s));
    x.cubtmp.reserve(tmp_bytes);
    Plan *d_plan = (Plan *)x.plan.p;
    CK(cudaMemsetAsync(d_plan, 0, sizeof(Plan), s));
    const int max_elems = BSW_NBINS * 16;
    prep_kernel<<<(n + 255) / 256, 256, 0, s>>>(d_pairs, d_qer, n, k.c, k.mode, max_elems, tier8_check, rlo, rhi, qlo, qhi,
                                                 (uint32_t *)x.keys_a.p, (uint32_t *)x.idx_a.p, d_plan);
    CK(cudaGetLastError());
    CK(cub::DeviceRadixSort::SortPairsDescending(x.cubtmp.p, tmp_bytes, (uint32_t *)x.keys_a.p, (uint32_t *)x.keys_b.p,
                                                 (uint32_t *)x.idx_a.p, (uint32_t *)x.idx_b.p, n, 0, BSW_KEY_BITS, s));
    CK(cudaMemcpyAsync(x.h_plan, d_plan, sizeof(Plan), cudaMemcpyDeviceToHost, s));
    k.st.kernel_launches += 1 + 8;    /* prep + the sort's internal launches (upper bound of cub's onesweep path) */
}

/* blocks until the histogram of plan_enqueue has landed in the slot's pinned mailbox */
const Plan *plan_wait(Slot &x, cudaStream_t s)
{
    CK(cudaStreamSynchronize(s));
    if (x.h_plan->bad_pairs) {
        fprintf(stderr, "[bsw-b200] fatal: %d pair(s) violate the call contract (negative length, h0 <= 0, or "
                        ">= %d bases in getScores8)\n", x.h_plan->bad_pairs, BSW_MAX_SEQ_LEN8);
        exit(EXIT_FAILURE);
    }
    return x.h_plan;
}

const Plan *run_plan(Core &k, Slot &x, const bsw_seqpair_t *d_pairs, const uint8_t *d_qer, int n, cudaStream_t s, int tier8_check)
{
    plan_enqueue(k, x, d_pairs, d_qer, n, s, tier8_check);
    return plan_wait(x, s);
}

void launch_bsw_dp(Core &k, DevCtx &x, Slot &sl, const Plan *pl, bsw_seqpair_t *d_pairs, const uint8_t *d_ref,
                   const uint8_t *d_qer, int n, int w, cudaStream_t s);

void launch_bsw(Core &k, DevCtx &x, Slot &sl, bsw_seqpair_t *d_pairs, const uint8_t *d_ref, const uint8_t *d_qer,
                int n, int w, cudaStream_t s, int tier8_check)
{
    if (n <= 0) return;
    const Plan *pl = run_plan(k, sl, d_pairs, d_qer, n, s, tier8_check);
    launch_bsw_dp(k, x, sl, pl, d_pairs, d_ref, d_qer, n, w, s);
}

/* Priority level (0 = highest) of every non-empty bin of a call: longest queries first, one level per bin while the
 * device has levels left; otherwise neighbouring bins at the two ends of the list share a level, the pair with the
 * smaller estimated work (pairs x columns^2) first. */
static void assign_levels(const Plan *pl, int nlev, int level_of[BSW_NBINS + 1])
{
    std::vector<int> last;              /* index (into `bins`) of the last bin of each group */
    std::vector<double> gw;
    std::vector<int> bins;
    for (int qb = BSW_NBINS; qb >= 1; --qb) {
        level_of[qb] = 0;
        const double cnt = (double)std::max(0, pl->hist[qb]) + (double)std::max(0, pl->hist_n[qb]);
        if (cnt > 0) { bins.push_back(qb); last.push_back((int)bins.size() - 1); gw.push_back(cnt * qb * qb); }
    }
    while ((int)gw.size() > nlev) {
        const size_t m = gw.size();
        if (gw[0] + gw[1] <= gw[m - 2] + gw[m - 1]) { gw[0] += gw[1]; last[0] = last[1]; gw.erase(gw.begin() + 1); last.erase(last.begin() + 1); }
        else { gw[m - 2] += gw[m - 1]; last[m - 2] = last[m - 1]; gw.pop_back(); last.pop_back(); }
    }
    for (size_t g = 0, b = 0; g < gw.size(); ++g)
        for (; (int)b <= last[g]; ++b) level_of[bins[b]] = (int)g;
}

void launch_bsw_dp(Core &k, DevCtx &x, Slot &sl, const Plan *pl, bsw_seqpair_t *d_pairs, const uint8_t *d_ref,
                   const uint8_t *d_qer, int n, int w, cudaStream_t s)
{
    int level_of[BSW_NBINS + 1];
    assign_levels(pl, x.nlevels, level_of);
    const uint32_t *sorted = (const uint32_t *)sl.idx_b.p;
    if (k.timing) CK(cudaEventRecord(x.ev0, s));
    int n_packed = 0;
    CK(cudaEventRecord(x.fork_ev, s));
    unsigned used = 0;
    /* sorted order: pairs whose query holds an N first (key bit 25), then the others; each longest query first */
    for (int pass = 0; pass < 2; ++pass)
        for (int qb = BSW_NBINS; qb >= 1; --qb) {
            const int cnt = pass == 0 ? pl->hist_n[qb] : pl->hist[qb];
            if (cnt <= 0) continue;
            const int NE = qb * 16, grid = (cnt + 63) / 64;
            cudaStream_t bs = x.lvl_stream[level_of[qb]][qb];
            if (!(used & (1u << qb))) CK(cudaStreamWaitEvent(bs, x.fork_ev, 0));
            used |= 1u << qb;
            /* template flags: exact cell counting (first-generation row loop); symmetric gap sums (one packed add less
             * per cell); general score table for queries with an N */
            const bool sym = k.c.oe_del == k.c.oe_ins;
            auto kern = k.count_cells ? (sym ? bsw_packed_kernel<true, true, false> : bsw_packed_kernel<true, false, false>)
                      : pass == 0     ? (sym ? bsw_packed_kernel<false, true, true> : bsw_packed_kernel<false, false, true>)
                                      : (sym ? bsw_packed_kernel<false, true, false> : bsw_packed_kernel<false, false, false>);
            kern<<<grid, 32, NE * BSW_WARP_ELEM_BYTES, bs>>>(d_pairs, d_ref, d_qer, sorted, n_packed, n_packed + cnt,
                                                            NE, k.c, w, (unsigned long long *)x.cells.p);
            CK(cudaGetLastError());
            k.st.kernel_launches++;
            n_packed += cnt;
        }
    cudaEvent_t trace_ev[BSW_NBINS + 1] = {};
    const bool trace = k.timing && env_int("BSW_CUDA_BIN_TRACE", 0);      /* debug: when does each bin's last block retire? */
    for (int qb = BSW_NBINS; qb >= 1; --qb)
        if (used & (1u << qb)) {
            cudaStream_t bs = x.lvl_stream[level_of[qb]][qb];
            if (trace) { CK(cudaEventCreate(&trace_ev[qb])); CK(cudaEventRecord(trace_ev[qb], bs)); }
            CK(cudaEventRecord(x.bin_ev[qb], bs));
            CK(cudaStreamWaitEvent(s, x.bin_ev[qb], 0));
        }
    if (pl->n_general > 0) {
        /* bounded scratch: at most ~1 GiB of eh rows per launch */
        const size_t per_pair = ((size_t)pl->max_len2_general + 2) * sizeof(int2);
        int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)pl->n_general, ((size_t)1 << 30) / per_pair));
        sl.scratch.reserve((size_t)chunk * per_pair);
        for (int first = 0; first < pl->n_general; first += chunk) {
            int cnt = std::min(chunk, pl->n_general - first);
            bsw_general_kernel<<<(cnt + 63) / 64, 64, 0, s>>>(d_pairs, d_ref, d_qer, sorted, n_packed + first, cnt,
                                                              k.c, w, (int2 *)sl.scratch.p, (unsigned long long *)x.cells.p);
            CK(cudaGetLastError());
            k.st.kernel_launches++;
        }
    }
    if (k.timing) {
        CK(cudaEventRecord(x.ev1, s)); CK(cudaEventSynchronize(x.ev1));
        float ms = 0; CK(cudaEventElapsedTime(&ms, x.ev0, x.ev1)); k.st.kernel_ms += ms;
        if (trace) {
            fprintf(stderr, "[bsw-b200] bin trace: call of %d pairs, DP window %.3f ms; last block of bin (columns: ms, pairs):", n, ms);
            for (int qb = BSW_NBINS; qb >= 1; --qb)
                if (trace_ev[qb]) {
                    float t = 0; CK(cudaEventElapsedTime(&t, x.ev0, trace_ev[qb]));
                    fprintf(stderr, " %d: %.3f, %d;", qb * 16, t, pl->hist[qb] + pl->hist_n[qb]);
                    cudaEventDestroy(trace_ev[qb]);
                }
            fprintf(stderr, "\n");
        }
    }
    k.st.pairs += n; k.st.pairs_packed += n_packed; k.st.pairs_general += pl->n_general;
}

void launch_kswv(Core &k, DevCtx &x, Slot &sl, const bsw_seqpair_t *d_pairs, const uint8_t *d_ref, const uint8_t *d_qer,
                 bsw_kswr_t *d_aln, int n, int is8, int phase, cudaStream_t s)
{
    if (n <= 0) return;
    const Plan *pl = run_plan(k, sl, d_pairs, d_qer, n, s, 0);
    const uint32_t *sorted = (const uint32_t *)sl.idx_b.p;
    if (k.timing) CK(cudaEventRecord(x.ev0, s));
    int n_packed = 0;
    for (int qb = 1; qb <= BSW_NBINS; ++qb) n_packed += pl->hist[qb];
    if (n_packed > 0) {
        const long long stride = n_packed;
        sl.rowmax.reserve((size_t)std::max(1, pl->max_len1) * (size_t)stride * sizeof(uint16_t));
        int first = 0;
        CK(cudaEventRecord(x.fork_ev, s));
        for (int qb = BSW_NBINS; qb >= 1; --qb) {
            const int cnt = pl->hist[qb];
            if (cnt <= 0) continue;
            const int NE = qb * 16, grid = (cnt + 63) / 64;
            cudaStream_t bs = x.bin_stream[qb];
            CK(cudaStreamWaitEvent(bs, x.fork_ev, 0));
            /* SYM=false on purpose: sharing h - oe puts a third op on kswv's f -> h -> f chain (measured 3 % slower) */
            auto kern = kswv_packed_kernel<false>;
            kern<<<grid, 32, NE * BSW_WARP_ELEM_BYTES, bs>>>(d_pairs, d_ref, d_qer, d_aln, sorted, first, first + cnt, NE,
                                                                          k.c, is8, phase, (uint16_t *)sl.rowmax.p, stride,
                                                                          (unsigned long long *)x.cells.p);
            CK(cudaGetLastError());
            CK(cudaEventRecord(x.bin_ev[qb], bs));
            CK(cudaStreamWaitEvent(s, x.bin_ev[qb], 0));
            k.st.kernel_launches++;
            first += cnt;
        }
    }
    if (pl->n_general > 0) {
        const size_t per_pair = ((size_t)pl->max_len2_general + 2) * sizeof(int2) + ((size_t)pl->max_len1 + 2) * sizeof(int);
        int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)pl->n_general, ((size_t)1 << 30) / per_pair));
        sl.scratch.reserve((size_t)chunk * per_pair);
        int2 *he = (int2 *)sl.scratch.p;
        int *rm = (int *)((char *)sl.scratch.p + (size_t)chunk * ((size_t)pl->max_len2_general + 2) * sizeof(int2));
        for (int first = 0; first < pl->n_general; first += chunk) {
            int cnt = std::min(chunk, pl->n_general - first);
            kswv_general_kernel<<<(cnt + 63) / 64, 64, 0, s>>>(d_pairs, d_ref, d_qer, d_aln, sorted, n_packed + first, cnt,
                                                               k.c, is8, phase, he, rm, (unsigned long long *)x.cells.p);
            CK(cudaGetLastError());
            k.st.kernel_launches++;
        }
    }
    if (k.timing) {
        CK(cudaEventRecord(x.ev1, s)); CK(cudaEventSynchronize(x.ev1));
        float ms = 0; CK(cudaEventElapsedTime(&ms, x.ev0, x.ev1)); k.st.kernel_ms += ms;
    }
    k.st.pairs += n; k.st.pairs_packed += n_packed; k.st.pairs_general += pl->n_general;
}

/* byte extents [lo, hi) of the sequence buffers touched by pairs[first, first+cnt) */
void extents(const bsw_seqpair_t *p, int first, int cnt, long long &rlo, long long &rhi, long long &qlo, long long &qhi)
{
    rlo = qlo = (long long)1 << 62; rhi = qhi = 0;
    for (int i = first; i < first + cnt; ++i) {
        long long a = p[i].idr, b = a + std::max(p[i].len1, 0), c = p[i].idq, d = c + std::max(p[i].len2, 0);
        rlo = std::min(rlo, a); rhi = std::max(rhi, b); qlo = std::min(qlo, c); qhi = std::max(qhi, d);
    }
    if (rhi <= rlo) { rlo = 0; rhi = 0; }
    if (qhi <= qlo) { qlo = 0; qhi = 0; }
}

/* Estimate of the same ranges from every 32nd pair plus the first and last one.  bwa-mem2 fills its sequence
 * arenas in pair order, so the estimate is normally exact; prep_kernel checks every pair against it on the
 * device (Plan::oob_pairs) and the chunk is redone with the exact scan if any pair falls outside.  This keeps
 * the calling host thread from reading all 56-byte records (0.56 GB per 10M pairs) before it can start a copy. */
void extents_sampled(const bsw_seqpair_t *p, int first, int cnt, long long &rlo, long long &rhi, long long &qlo, long long &qhi)
{
    rlo = qlo = (long long)1 << 62; rhi = qhi = 0;
    auto take = [&](int i) {
        long long a = p[i].idr, b = a + std::max(p[i].len1, 0), c = p[i].idq, d = c + std::max(p[i].len2, 0);
        rlo = std::min(rlo, a); rhi = std::max(rhi, b); qlo = std::min(qlo, c); qhi = std::max(qhi, d);
    };
    for (int i = first; i < first + cnt; i += 32) take(i);
    for (int i = std::max(first, first + cnt - 32); i < first + cnt; ++i) take(i);
    if (rhi <= rlo) { rlo = 0; rhi = 0; }
    if (qhi <= qlo) { qlo = 0; qhi = 0; }
}

/* ---- BSW host-pointer call: chunked, pipelined, sharded ---------------------------------------
 * The pair array is cut into chunks of BSW_CUDA_CHUNK_PAIRS pairs, dealt round-robin to the handle's
 * devices.  Each chunk uploads only the byte ranges of the sequence buffers its pairs touch, so the
 * H2D copy of chunk c+1 and the D2H copy of chunk c-1 overlap the kernels of chunk c (three slots per
 * device).  Results land directly in the caller's SeqPair array.  Synchronous on return. */
void slot_retire(Core &k, DevCtx &x, Slot &sl)
{
    if (sl.dev_used) { CK(cudaSetDevice(x.dev)); CK(cudaEventSynchronize(sl.dev_done)); sl.dev_used = false; }
    if (!sl.busy) return;
    CK(cudaSetDevice(x.dev));
    CK(cudaEventSynchronize(sl.done));
    sl.busy = false;
    (void)k;
}

void host_call_bsw(Core &k, bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, int n, int w, int tier8_check)
{
    k.st.calls++;
    if (n <= 0) return;
    if (!pairs || !ref || !qer) die("null buffer passed to getScores");
    const int nd = (int)k.ctx.size();
    const int chunk = std::max(1024, k.chunk_pairs);
    const int nchunks = (n + chunk - 1) / chunk;
    struct Job { int first, cnt; long long rlo, qlo; DevCtx *x; Slot *sl; };
    auto upload = [&](Job &j, bool exact) {
        long long rhi, qhi;
        if (exact) extents(pairs, j.first, j.cnt, j.rlo, rhi, j.qlo, qhi);
        else extents_sampled(pairs, j.first, j.cnt, j.rlo, rhi, j.qlo, qhi);
        Slot &sl = *j.sl;
        sl.ref.reserve((size_t)(rhi - j.rlo) + 64);
        sl.qer.reserve((size_t)(qhi - j.qlo) + 64);
        if (rhi > j.rlo) CK(cudaMemcpyAsync(sl.ref.p, ref + j.rlo, (size_t)(rhi - j.rlo), cudaMemcpyHostToDevice, sl.stream));
        if (qhi > j.qlo) CK(cudaMemcpyAsync(sl.qer.p, qer + j.qlo, (size_t)(qhi - j.qlo), cudaMemcpyHostToDevice, sl.stream));
        k.st.h2d_bytes += (rhi - j.rlo) + (qhi - j.qlo);
        plan_enqueue(k, sl, (const bsw_seqpair_t *)sl.pairs.p, (const uint8_t *)sl.qer.p - j.qlo, j.cnt, sl.stream, tier8_check, j.rlo, rhi, j.qlo, qhi);
    };
    /* gathered upload: (1) every helper thread sums the lengths of its share of the chunk's pairs, (2) exclusive prefix
     * over the shares, (3) every thread copies its pairs' bases to their place in the pinned staging buffers and notes
     * the new offsets; then pairs, offsets and the two dense buffers go up, and swap_offsets_kernel points the device
     * copy of the records at the dense buffers */
    auto upload_gathered = [&](Job &j) {
        Slot &sl = *j.sl;
        if (!k.pool) { k.pool = new Pool(); k.pool->start(std::max(1, k.gather)); }
        const int T = k.pool->size();
        std::vector<long long> sr(T + 1, 0), sq(T + 1, 0);
        auto share = [&](int t, int &b, int &e) { b = j.first + (int)((long long)j.cnt * t / T); e = j.first + (int)((long long)j.cnt * (t + 1) / T); };
        k.pool->run([&](int t) {
            int b, e; share(t, b, e);
            long long r = 0, q = 0;
            for (int i = b; i < e; ++i) { r += std::max(pairs[i].len1, 0); q += std::max(pairs[i].len2, 0); }
            sr[t + 1] = r; sq[t + 1] = q;
        });
        for (int t = 0; t < T; ++t) { sr[t + 1] += sr[t]; sq[t + 1] += sq[t]; }
        const long long tot_r = sr[T], tot_q = sq[T];
        if (tot_r >= ((long long)1 << 31) || tot_q >= ((long long)1 << 31)) die("chunk too large for 32-bit sequence offsets");
        sl.st_ref.reserve((size_t)tot_r + 64); sl.st_qer.reserve((size_t)tot_q + 64); sl.st_off.reserve((size_t)j.cnt * sizeof(int2));
        uint8_t *gr = (uint8_t *)sl.st_ref.p, *gq = (uint8_t *)sl.st_qer.p;
        int2 *go = (int2 *)sl.st_off.p;
        k.pool->run([&](int t) {
            int b, e; share(t, b, e);
            long long r = sr[t], q = sq[t];
            for (int i = b; i < e; ++i) {
                const int l1 = std::max(pairs[i].len1, 0), l2 = std::max(pairs[i].len2, 0);
                memcpy(gr + r, ref + pairs[i].idr, (size_t)l1);
                memcpy(gq + q, qer + pairs[i].idq, (size_t)l2);
                go[i - j.first] = make_int2((int)r, (int)q);
                r += l1; q += l2;
            }
        });
        sl.ref.reserve((size_t)tot_r + 64); sl.qer.reserve((size_t)tot_q + 64); sl.off.reserve((size_t)j.cnt * sizeof(int2));
        CK(cudaMemcpyAsync(sl.off.p, go, (size_t)j.cnt * sizeof(int2), cudaMemcpyHostToDevice, sl.stream));
        if (tot_r > 0) CK(cudaMemcpyAsync(sl.ref.p, gr, (size_t)tot_r, cudaMemcpyHostToDevice, sl.stream));
        if (tot_q > 0) CK(cudaMemcpyAsync(sl.qer.p, gq, (size_t)tot_q, cudaMemcpyHostToDevice, sl.stream));
        k.st.h2d_bytes += tot_r + tot_q + (long long)j.cnt * (long long)sizeof(int2);
        bswk::swap_offsets_kernel<<<(j.cnt + 255) / 256, 256, 0, sl.stream>>>((bsw_seqpair_t *)sl.pairs.p, (int2 *)sl.off.p, j.cnt);
        CK(cudaGetLastError());
        k.st.kernel_launches++;
        j.rlo = 0; j.qlo = 0;
        plan_enqueue(k, sl, (const bsw_seqpair_t *)sl.pairs.p, (const uint8_t *)sl.qer.p, j.cnt, sl.stream, tier8_check,
                     0, std::max(tot_r, 1LL), 0, std::max(tot_q, 1LL));
    };
    std::vector<Job> jobs(nchunks);
    /* stage A: extents, uploads, prep + sort + histogram read-back -- all asynchronous */
    auto stage_a = [&](int ci) {
        Job &j = jobs[ci];
        j.first = ci * chunk; j.cnt = std::min(chunk, n - j.first);
        j.x = &k.ctx[ci % nd]; j.sl = &j.x->slot[(ci / nd) % BSW_NSLOT];
        slot_retire(k, *j.x, *j.sl);              /* its previous chunk must have landed */
        Slot &sl = *j.sl;
        CK(cudaSetDevice(j.x->dev));
        sl.pairs.reserve((size_t)j.cnt * sizeof(bsw_seqpair_t));
        CK(cudaMemcpyAsync(sl.pairs.p, pairs + j.first, (size_t)j.cnt * sizeof(bsw_seqpair_t), cudaMemcpyHostToDevice, sl.stream));
        k.st.h2d_bytes += (int64_t)j.cnt * sizeof(bsw_seqpair_t);
        if (k.gather > 0) upload_gathered(j);
        else upload(j, false);
        sl.busy = true;
    };
    /* stage B: wait for the histogram, launch the DP kernels, download the results */
    auto stage_b = [&](int ci) {
        Job &j = jobs[ci];
        Slot &sl = *j.sl;
        CK(cudaSetDevice(j.x->dev));
        const Plan *pl = plan_wait(sl, sl.stream);
        if (pl->oob_pairs) {                      /* the sampled ranges missed a pair: redo with the exact scan */
            if (k.gather > 0) die("internal: gathered offsets out of range");
            upload(j, true);
            pl = plan_wait(sl, sl.stream);
            if (pl->oob_pairs) die("internal: exact extents still miss a pair");
        }
        /* the kernels index the sequence buffers with the caller's idr/idq: rebase the pointers */
        launch_bsw_dp(k, *j.x, sl, pl, (bsw_seqpair_t *)sl.pairs.p, (const uint8_t *)sl.ref.p - j.rlo,
                      (const uint8_t *)sl.qer.p - j.qlo, j.cnt, w, sl.stream);
        if (k.gather > 0) {                       /* the caller's idr / idq go back into the records before the download */
            bswk::swap_offsets_kernel<<<(j.cnt + 255) / 256, 256, 0, sl.stream>>>((bsw_seqpair_t *)sl.pairs.p, (int2 *)sl.off.p, j.cnt);
            CK(cudaGetLastError());
            k.st.kernel_launches++;
        }
        CK(cudaMemcpyAsync(pairs + j.first, sl.pairs.p, (size_t)j.cnt * sizeof(bsw_seqpair_t), cudaMemcpyDeviceToHost, sl.stream));
        CK(cudaEventRecord(sl.done, sl.stream));
        k.st.d2h_bytes += (int64_t)j.cnt * sizeof(bsw_seqpair_t);
    };
    /* chunk c+nd is uploading while chunk c waits for its histogram: the copy engine never idles */
    const int ahead = nd * std::max(1, std::min(BSW_NSLOT - 2, k.ahead));   /* chunks uploading ahead of the one waiting */
    for (int ci = 0; ci < std::min(ahead, nchunks); ++ci) stage_a(ci);
    for (int ci = 0; ci < nchunks; ++ci) {
        if (ci + ahead < nchunks) stage_a(ci + ahead);
        stage_b(ci);
    }
    for (DevCtx &x : k.ctx)
        for (Slot &sl : x.slot) slot_retire(k, x, sl);
}

/* ---- kswv host-pointer call: one slice per device (slot 0), synchronous -------------------------- */
void host_call_kswv_part(Core &k, bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, bsw_kswr_t *aln,
                         int n, int is8, int phase);

/* bounded device scratch: the row-maximum scratch is (max len1) x (pairs) x 2 B, so large calls go in parts of
 * BSW_CUDA_KSWV_PART pairs (phase 0 indexes aln by position, phase 1 by regid) */
void host_call_kswv(Core &k, bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, bsw_kswr_t *aln,
                    int n, int is8, int phase)
{
    k.st.calls++;
    const int part = std::max(1024, env_int("BSW_CUDA_KSWV_PART", 1 << 21));
    for (int off = 0; off < n; off += part)
        host_call_kswv_part(k, pairs + off, ref, qer, phase == 0 ? aln + off : aln, std::min(part, n - off), is8, phase);
}

void host_call_kswv_part(Core &k, bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, bsw_kswr_t *aln,
                         int n, int is8, int phase)
{
    if (n <= 0) return;
    if (!pairs || !ref || !qer || !aln) die("null buffer passed to kswv getScores");
    const int nd = (int)k.ctx.size();
    struct Slice { int first, cnt; long long rlo, rhi, qlo, qhi; int alo, ahi; };
    std::vector<Slice> sl(nd);
    for (int d = 0; d < nd; ++d) {
        Slice &s = sl[d];
        s.first = (int)((long long)n * d / nd);
        s.cnt = (int)((long long)n * (d + 1) / nd) - s.first;
        if (s.cnt <= 0) continue;
        extents(pairs, s.first, s.cnt, s.rlo, s.rhi, s.qlo, s.qhi);
        DevCtx &x = k.ctx[d];
        Slot &t = x.slot[0];
        CK(cudaSetDevice(x.dev));
        t.pairs.reserve((size_t)s.cnt * sizeof(bsw_seqpair_t));
        t.ref.reserve((size_t)(s.rhi - s.rlo) + 64);
        t.qer.reserve((size_t)(s.qhi - s.qlo) + 64);
        CK(cudaMemcpyAsync(t.pairs.p, pairs + s.first, (size_t)s.cnt * sizeof(bsw_seqpair_t), cudaMemcpyHostToDevice, t.stream));
        if (s.rhi > s.rlo) CK(cudaMemcpyAsync(t.ref.p, ref + s.rlo, (size_t)(s.rhi - s.rlo), cudaMemcpyHostToDevice, t.stream));
        if (s.qhi > s.qlo) CK(cudaMemcpyAsync(t.qer.p, qer + s.qlo, (size_t)(s.qhi - s.qlo), cudaMemcpyHostToDevice, t.stream));
        k.st.h2d_bytes += (int64_t)s.cnt * sizeof(bsw_seqpair_t) + (s.rhi - s.rlo) + (s.qhi - s.qlo);
        if (phase == 0) { s.alo = s.first; s.ahi = s.first + s.cnt; t.aln.reserve((size_t)s.cnt * sizeof(bsw_kswr_t)); }
        else {
            /* phase 1 refines aln[regid]: send the span [min regid, max regid] this slice names */
            int lo = 0x7fffffff, hi = -1;
            for (int i = s.first; i < s.first + s.cnt; ++i) { lo = std::min(lo, pairs[i].regid); hi = std::max(hi, pairs[i].regid); }
            if (lo < 0) die("kswv phase 1: negative regid");
            s.alo = lo; s.ahi = hi + 1;
            t.aln.reserve((size_t)(hi + 1 - lo) * sizeof(bsw_kswr_t));
            CK(cudaMemcpyAsync(t.aln.p, aln + lo, (size_t)(hi + 1 - lo) * sizeof(bsw_kswr_t), cudaMemcpyHostToDevice, t.stream));
            k.st.h2d_bytes += (int64_t)(hi + 1 - lo) * sizeof(bsw_kswr_t);
        }
    }
    for (int d = 0; d < nd; ++d) {
        const Slice &s = sl[d];
        if (s.cnt <= 0) continue;
        DevCtx &x = k.ctx[d];
        Slot &t = x.slot[0];
        CK(cudaSetDevice(x.dev));
        const uint8_t *d_ref = (const uint8_t *)t.ref.p - s.rlo;
        const uint8_t *d_qer = (const uint8_t *)t.qer.p - s.qlo;
        /* phase 0 indexes aln by slice-local pair index; phase 1 by regid, rebased to the span */
        bsw_kswr_t *d_aln = (bsw_kswr_t *)t.aln.p - (phase == 0 ? 0 : s.alo);
        launch_kswv(k, x, t, (const bsw_seqpair_t *)t.pairs.p, d_ref, d_qer, d_aln, s.cnt, is8, phase, t.stream);
        const size_t cnt = (size_t)(s.ahi - s.alo);
        if (phase == 0) {
            CK(cudaMemcpyAsync(aln + s.alo, t.aln.p, cnt * sizeof(bsw_kswr_t), cudaMemcpyDeviceToHost, t.stream));
        } else {
            /* only tb/qb of the named entries change; copy back through a bounce so that entries of
             * other devices inside the span are not clobbered */
            std::vector<bsw_kswr_t> tmp(cnt);
            CK(cudaMemcpyAsync(tmp.data(), t.aln.p, cnt * sizeof(bsw_kswr_t), cudaMemcpyDeviceToHost, t.stream));
            CK(cudaStreamSynchronize(t.stream));
            for (int i = s.first; i < s.first + s.cnt; ++i) {
                const int r = pairs[i].regid;
                aln[r].tb = tmp[r - s.alo].tb; aln[r].qb = tmp[r - s.alo].qb;
            }
        }
        k.st.d2h_bytes += (int64_t)cnt * sizeof(bsw_kswr_t);
    }
    for (int d = 0; d < nd; ++d) {
        if (sl[d].cnt <= 0) continue;
        CK(cudaSetDevice(k.ctx[d].dev));
        CK(cudaStreamSynchronize(k.ctx[d].slot[0].stream));
    }
}

} /* namespace */

struct bsw_handle  { Core k; };
struct kswv_handle { Core k; int32_t maxRefLen, maxQerLen; };

extern "C" {

const char *bsw_version(void) { return "bsw-b200 0.1 (sm_100a, packed s16x2 DPX)"; }

int bsw_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

bsw_t *bsw_create(int o_del, int e_del, int o_ins, int e_ins, int zdrop, int end_bonus,
                  const int8_t *mat, int8_t w_match, int8_t w_mismatch, int numThreads)
{
    (void)mat; (void)numThreads;
    bsw_t *h = new bsw_handle();
    h->k.c = make_const(o_del, e_del, o_ins, e_ins, zdrop, end_bonus, w_match, w_mismatch);
    core_init(h->k, MODE_BSW);
    return h;
}

void bsw_destroy(bsw_t *h) { if (!h) return; core_destroy(h->k); delete h; }

void bsw_getScores8(bsw_t *h, bsw_seqpair_t *pairArray, const uint8_t *seqBufRef, const uint8_t *seqBufQer,
                    int32_t numPairs, uint16_t numThreads, int32_t w)
{
    (void)numThreads;
    host_call_bsw(h->k, pairArray, seqBufRef, seqBufQer, numPairs, w, 1);
}

void bsw_getScores16(bsw_t *h, bsw_seqpair_t *pairArray, const uint8_t *seqBufRef, const uint8_t *seqBufQer,
                     int32_t numPairs, uint16_t numThreads, int32_t w)
{
    (void)numThreads;
    host_call_bsw(h->k, pairArray, seqBufRef, seqBufQer, numPairs, w, 0);
}

void bsw_scalarBandedSWAWrapper(bsw_t *h, bsw_seqpair_t *pairArray, const uint8_t *seqBufRef, const uint8_t *seqBufQer,
                                int32_t numPairs, int nthreads, int32_t w)
{
    (void)nthreads;
    host_call_bsw(h->k, pairArray, seqBufRef, seqBufQer, numPairs, w, 0);
}

/* Device-resident calls rotate over the slots (sort keys, permutation, histogram mailbox, general-path scratch), so
 * that calls issued on DIFFERENT caller streams overlap: the next call's prep / sort / DP blocks fill the SMs while the
 * previous call's last warps drain (bench step of four 2.5M-pair calls: 18.54 ms on one stream, 18.16 ms on four).
 * Calls on one stream stay ordered by that stream.  A slot is reused only after the call that last used it has finished (stream-side wait). */
static Slot &dev_slot(DevCtx &x, cudaStream_t s, bool rotate)
{
    /* kswv keeps to one slot: its row-maxima scratch is ~1 GB per million pairs, and a slot's buffers grow on first use */
    Slot &sl = x.slot[rotate ? x.dev_rr : 0];
    if (rotate) x.dev_rr = (x.dev_rr + 1) % BSW_NSLOT;
    if (sl.dev_used) CK(cudaStreamWaitEvent(s, sl.dev_done, 0));
    sl.dev_used = true;
    return sl;
}

void bsw_dev_getScores(bsw_t *h, bsw_seqpair_t *d_pairArray, const uint8_t *d_seqBufRef, const uint8_t *d_seqBufQer,
                       int32_t numPairs, int32_t w, void *stream)
{
    DevCtx &x = h->k.ctx[0];
    CK(cudaSetDevice(x.dev));
    h->k.st.calls++;
    Slot &sl = dev_slot(x, (cudaStream_t)stream, true);
    launch_bsw(h->k, x, sl, d_pairArray, d_seqBufRef, d_seqBufQer, numPairs, w, (cudaStream_t)stream, 0);
    CK(cudaEventRecord(sl.dev_done, (cudaStream_t)stream));
}

/* One side (left or right extensions) of the caller protocol on resident pairs: attempt 0 with band w on all pairs,
 * then up to maxBandTry-1 retries with the band doubled, each on the pairs the upstream rule selects (score changed by
 * the attempt and max_off >= 3/4 of the attempt's band), moved to a dense sub-array and back.  One 4-byte read-back per
 * retry (the count) on top of the histogram read-back of every BSW launch. */
static void dev_extend_side(Core &k, DevCtx &x, Slot &sl, bsw_seqpair_t *d_pairs, const uint8_t *d_ref, const uint8_t *d_qer,
                            int n, int w, int max_try, cudaStream_t s)
{
    if (n <= 0) return;
    const int T = 256, G = (n + T - 1) / T;
    sl.px_prev.reserve((size_t)n * sizeof(int));
    sl.px_list.reserve((size_t)n * sizeof(int));
    sl.px_count.reserve(sizeof(int));
    bswk::proto_save_score_kernel<<<G, T, 0, s>>>(d_pairs, n, (int *)sl.px_prev.p);
    launch_bsw(k, x, sl, d_pairs, d_ref, d_qer, n, w, s, 0);
    for (int attempt = 1; attempt < max_try; ++attempt) {
        const int w_prev = w << (attempt - 1), w_now = w << attempt;
        CK(cudaMemsetAsync(sl.px_count.p, 0, sizeof(int), s));
        bswk::proto_select_retry_kernel<<<G, T, 0, s>>>(d_pairs, (const int *)sl.px_prev.p, n, (w_prev >> 1) + (w_prev >> 2),
                                                        (int *)sl.px_list.p, (int *)sl.px_count.p);
        CK(cudaMemcpyAsync(sl.h_count, sl.px_count.p, sizeof(int), cudaMemcpyDeviceToHost, s));
        if (attempt + 1 < max_try) bswk::proto_save_score_kernel<<<G, T, 0, s>>>(d_pairs, n, (int *)sl.px_prev.p);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(s));
        const int m = *sl.h_count;
        k.st.kernel_launches += 2;
        if (m <= 0) break;
        sl.px_sub.reserve((size_t)m * sizeof(bsw_seqpair_t));
        const int Gm = (m + T - 1) / T;
        bswk::proto_gather_kernel<<<Gm, T, 0, s>>>(d_pairs, (const int *)sl.px_list.p, m, (bsw_seqpair_t *)sl.px_sub.p);
        launch_bsw(k, x, sl, (bsw_seqpair_t *)sl.px_sub.p, d_ref, d_qer, m, w_now, s, 0);
        bswk::proto_scatter_kernel<<<Gm, T, 0, s>>>((const bsw_seqpair_t *)sl.px_sub.p, (const int *)sl.px_list.p, m, d_pairs);
        CK(cudaGetLastError());
        k.st.kernel_launches += 2;
    }
}

void bsw_dev_extendChained(bsw_t *h, bsw_seqpair_t *d_left, const uint8_t *d_leftRef, const uint8_t *d_leftQer, int32_t numLeft,
                           bsw_seqpair_t *d_right, const uint8_t *d_rightRef, const uint8_t *d_rightQer, int32_t numRight,
                           int32_t w, int32_t maxBandTry, void *stream)
{
    DevCtx &x = h->k.ctx[0];
    CK(cudaSetDevice(x.dev));
    cudaStream_t s = (cudaStream_t)stream;
    h->k.st.calls++;
    Slot &sl = dev_slot(x, s, true);
    const int tries = maxBandTry < 1 ? 1 : maxBandTry;
    dev_extend_side(h->k, x, sl, d_left, d_leftRef, d_leftQer, numLeft, w, tries, s);
    if (numRight > 0) {
        bswk::proto_chain_h0_kernel<<<(numRight + 255) / 256, 256, 0, s>>>(d_left, numLeft, d_right, numRight);
        CK(cudaGetLastError());
        h->k.st.kernel_launches++;
        dev_extend_side(h->k, x, sl, d_right, d_rightRef, d_rightQer, numRight, w, tries, s);
    }
    CK(cudaEventRecord(sl.dev_done, s));
}

kswv_t *kswv_create(int o_del, int e_del, int o_ins, int e_ins, int8_t w_match, int8_t w_mismatch,
                    int numThreads, int32_t maxRefLen, int32_t maxQerLen)
{
    (void)numThreads;
    kswv_t *h = new kswv_handle();
    h->k.c = make_const(o_del, e_del, o_ins, e_ins, 0, 0, w_match, w_mismatch);
    h->maxRefLen = maxRefLen; h->maxQerLen = maxQerLen;
    core_init(h->k, MODE_KSWV);
    return h;
}

void kswv_destroy(kswv_t *h) { if (!h) return; core_destroy(h->k); delete h; }

void kswv_getScores8(kswv_t *h, bsw_seqpair_t *pairArray, const uint8_t *seqBufRef, const uint8_t *seqBufQer,
                     bsw_kswr_t *aln, int32_t numPairs, uint16_t numThreads, int phase)
{
    (void)numThreads;
    host_call_kswv(h->k, pairArray, seqBufRef, seqBufQer, aln, numPairs, 1, phase);
}

void kswv_getScores16(kswv_t *h, bsw_seqpair_t *pairArray, const uint8_t *seqBufRef, const uint8_t *seqBufQer,
                      bsw_kswr_t *aln, int32_t numPairs, uint16_t numThreads, int phase)
{
    (void)numThreads;
    host_call_kswv(h->k, pairArray, seqBufRef, seqBufQer, aln, numPairs, 0, phase);
}

void kswv_dev_getScores(kswv_t *h, const bsw_seqpair_t *d_pairArray, const uint8_t *d_seqBufRef, const uint8_t *d_seqBufQer,
                        bsw_kswr_t *d_aln, int32_t numPairs, int is8, int phase, void *stream)
{
    DevCtx &x = h->k.ctx[0];
    CK(cudaSetDevice(x.dev));
    h->k.st.calls++;
    Slot &sl = dev_slot(x, (cudaStream_t)stream, false);
    launch_kswv(h->k, x, sl, d_pairArray, d_seqBufRef, d_seqBufQer, d_aln, numPairs, is8, phase, (cudaStream_t)stream);
    CK(cudaEventRecord(sl.dev_done, (cudaStream_t)stream));
}

void *bsw_host_alloc(size_t bytes)
{
    void *p = nullptr;
    CK(cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable));
    return p;
}
void bsw_host_free(void *p) { if (p) CK(cudaFreeHost(p)); }

static void stats_with_pending(Core &k, bsw_stats_t *out)
{
    long long cells = 0;
    for (DevCtx &x : k.ctx) {
        CK(cudaSetDevice(x.dev));
        CK(cudaDeviceSynchronize());
        unsigned long long c = 0;
        CK(cudaMemcpy(&c, x.cells.p, sizeof c, cudaMemcpyDeviceToHost));
        cells += (long long)c;
    }
    k.st.cells = cells;
    *out = k.st;
}
static void stats_reset(Core &k)
{
    memset(&k.st, 0, sizeof k.st);
    for (DevCtx &x : k.ctx) {
        CK(cudaSetDevice(x.dev));
        CK(cudaDeviceSynchronize());
        CK(cudaMemset(x.cells.p, 0, sizeof(unsigned long long)));
    }
}

void bsw_get_stats(bsw_t *h, bsw_stats_t *out) { stats_with_pending(h->k, out); }
void bsw_reset_stats(bsw_t *h) { stats_reset(h->k); }
void bsw_set_timing(bsw_t *h, int on) { h->k.timing = on; }
void bsw_set_cell_counting(bsw_t *h, int on) { h->k.count_cells = on; }
void kswv_get_stats(kswv_t *h, bsw_stats_t *out) { stats_with_pending(h->k, out); }
void kswv_reset_stats(kswv_t *h) { stats_reset(h->k); }
void kswv_set_timing(kswv_t *h, int on) { h->k.timing = on; }

double bsw_measure_dpx_peak(int device, int which)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return 0.0;
    CK(cudaSetDevice(device));
    int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const int threads = 1024, blocks = sms * 2, iters = 4096;
    uint32_t *out = nullptr; CK(cudaMalloc(&out, (size_t)threads * blocks * 4));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        CK(cudaEventRecord(e0));
        switch (which) {
        case 0: dpx_peak_kernel<0><<<blocks, threads>>>(out, iters, 12345u + rep); break;
        case 1: dpx_peak_kernel<1><<<blocks, threads>>>(out, iters, 12345u + rep); break;
        case 2: dpx_peak_kernel<2><<<blocks, threads>>>(out, iters, 12345u + rep); break;
        case 3: dpx_peak_kernel<3><<<blocks, threads>>>(out, iters, 12345u + rep); break;
        default: dpx_peak_kernel<4><<<blocks, threads>>>(out, iters, 12345u + rep); break;
        }
        CK(cudaGetLastError());
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
        /* thread-level instruction count: 64 per iteration (8 chains x 8 unrolled); WHICH==3 issues 2 per step */
        double instr = (double)threads * blocks * iters * 64.0 * (which == 3 ? 2.0 : 1.0);
        if (rep > 0) best = std::max(best, instr / (ms * 1e-3));
    }
    CK(cudaEventDestroy(e0)); CK(cudaEventDestroy(e1)); CK(cudaFree(out));
    return best;
}

} /* extern "C" */
