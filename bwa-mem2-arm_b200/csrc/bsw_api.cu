/*
 * bsw_api.cu -- extern "C" boundary of the B200 BSW / kswv path (include/bsw_cuda.h).
 *
 * Host side of the drop-in replacement for BandedPairWiseSW::getScores8/16 and
 * kswv::getScores8/16 (bwa-mem2 src/bandedSWA.cpp, src/kswv.cpp; call sites in
 * mem_chain2aln_across_reads_V2 / mem_sam_pe_batch, SURVEY.md sections 3, 8b).
 *
 * A handle is cheap: scoring constants plus pointers.  Everything expensive is process-wide per device and
 * built once (DevRt): the priority streams of the per-bin launches, a pool of per-call contexts (chunk slots with
 * their streams, events, pinned mailboxes and grow-only scratch -- handed to a handle on its first large call and
 * returned to the pool by bsw_destroy) and the staged-call engine (engine.cuh).  The reference constructs its
 * BandedPairWiseSW / kswv objects per task and per worker thread; a handle that built
 * its own 116 streams and 8 pinned allocations cost 0.57 s to create (profiles/r02_pcie_probe_1gpu.json).
 *
 * Three ways in:
 *   host-pointer calls below BSW_CUDA_DIRECT_MIN pairs -> staged engine (merged with the other threads' calls);
 *   larger host-pointer calls -> "direct": chunked DMA from the caller's buffers, pipelined over chunk slots;
 *   bsw_dev_* / kswv_dev_* -> buffers already resident in HBM.
 * No CPU fallback: any CUDA failure is fatal, as in the reference (fprintf + exit).
 */
#include <vector>
#include <thread>
#include <string>
#include <algorithm>
#include <mutex>
#include <memory>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <thrust/iterator/transform_iterator.h>
#include <thrust/iterator/counting_iterator.h>
#include "host_common.cuh"
#include "bsw_kernels.cuh"
#include "engine.cuh"

using namespace bswk;
using namespace bswh;

namespace {

#define BSW_NSLOT 4
#define BSW_NLEVELS 6

#define BSW_ARENA_PIECES 32

struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    DevBuf pairs, ref, qer, aln, keys_a, keys_b, idx_a, idx_b, cubtmp, plan, scratch;
    DevBuf px_prev, px_list, px_sub, px_count;   /* device-side caller protocol (bsw_dev_extendChained) */
    DevBuf kx_off, kx_pairs, kx_ref, kx_qer;     /* device-side kswv hand-off (kswv_dev_align2) */
    DevBuf gx_eh, gx_z, gx_zoff, gx_coff, gx_jobs, gx_cigar;   /* CIGAR generation (bsw_global2) */
    DevBuf rpos, toff, reftmp;                    /* reference-resident calls: positions, target offsets, expanded targets */
    PinBuf h_res;                                 /* kswv phase 1: pinned landing area of a chunk's pass results */
    int *h_count = nullptr;                       /* pinned mailbox of the retry count */
    long long *h_tot = nullptr;                   /* pinned mailbox of 64-bit totals */
    Plan *h_plan = nullptr;                       /* pinned mailbox of the histogram */
    bool busy = false;                            /* a chunk is in flight on this slot */
    cudaEvent_t dev_done = nullptr;               /* end of the last device-resident call that used this slot's scratch */
    bool dev_used = false;
};

/* what one handle needs on one device while it makes large or resident calls; pooled per device */
struct DevCtx {
    int dev = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t fork_ev = nullptr, bin_ev[BSW_NBINS + 1] = {};
    DevBuf cells;
    int dev_rr = 0;                  /* next slot of the device-resident calls (see dev_slot) */
    Slot slot[BSW_NSLOT];
    /* sequence arenas declared with bsw_arena_begin: device copies + the event that marks them uploaded */
    DevBuf arena_ref, arena_qer;
    cudaStream_t arena_stream = nullptr;
    cudaEvent_t arena_ev[BSW_ARENA_PIECES] = {};     /* piece p of both arenas is on the device */
    int arena_enq = 0;                               /* pieces whose copies have been enqueued */
};

/* resident 2-bit reference (bsw_ref_upload), one copy per device */
struct RefCopy { int dev; uint8_t *d_pac; };

} /* namespace */

struct bsw_ref {
    int64_t l_pac = 0;
    std::vector<RefCopy> copies;
    const uint8_t *on(int dev) const
    {
        for (const RefCopy &c : copies) if (c.dev == dev) return c.d_pac;
        return nullptr;
    }
};

namespace {

/* process-wide, per device */
struct DevRt {
    int dev = 0;
    int sm_count = 0, smem_optin = 0, nlevels = 1;
    /* one stream per query-length bin (kswv) and per (priority level, bin) (BSW): the bins' kernels run concurrently,
     * so one bin's tail wave overlaps the next bin's blocks (fork/join on events around them).  The levels a call's
     * non-empty bins get are decided per call from its histogram (assign_levels). */
    cudaStream_t bin_stream[BSW_NBINS + 1] = {};
    cudaStream_t lvl_stream[BSW_NLEVELS][BSW_NBINS + 1] = {};
    std::mutex mu;
    std::vector<DevCtx *> free_ctx;
    std::unique_ptr<Engine> engine;
    std::once_flag engine_once;

    Engine &eng()
    {
        std::call_once(engine_once, [this] { engine.reset(new Engine(dev)); });
        return *engine;
    }
};

std::mutex g_rt_mu;
std::vector<DevRt *> g_rt;            /* index = CUDA device ordinal; entries live until process exit */

void stop_engines()
{
    std::lock_guard<std::mutex> lk(g_rt_mu);
    for (DevRt *r : g_rt) if (r && r->engine) r->engine->shutdown();
}

DevRt &devrt(int dev)
{
    std::lock_guard<std::mutex> lk(g_rt_mu);
    if ((int)g_rt.size() <= dev) g_rt.resize(dev + 1, nullptr);
    if (g_rt[dev]) return *g_rt[dev];
    static bool hooked = false;
    if (!hooked) { atexit(stop_engines); hooked = true; }
    DevRt *r = new DevRt();
    r->dev = dev;
    CK(cudaSetDevice(dev));
    CK(cudaDeviceGetAttribute(&r->sm_count, cudaDevAttrMultiProcessorCount, dev));
    CK(cudaDeviceGetAttribute(&r->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    int lo = 0, hi = 0;                       /* numerically lowest = highest priority */
    CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    r->nlevels = std::max(1, std::min(BSW_NLEVELS, lo - hi + 1));
    for (int qb = 1; qb <= BSW_NBINS; ++qb) {
        const int pr = hi + (BSW_NBINS - qb) * (lo - hi) / BSW_NBINS;   /* long queries first */
        CK(cudaStreamCreateWithPriority(&r->bin_stream[qb], cudaStreamNonBlocking, pr));
        for (int l = 0; l < r->nlevels; ++l)
            CK(cudaStreamCreateWithPriority(&r->lvl_stream[l][qb], cudaStreamNonBlocking, hi + l));
    }
    /* packed kernels: one warp per block, dynamic shared memory = NE x 192 B (NE <= 256) */
    const int max_smem = BSW_NBINS * 16 * BSW_WARP_ELEM_BYTES;
    if (max_smem > r->smem_optin) die("device shared memory too small for the packed kernels");
    const void *fns[] = { (const void *)bsw_packed_kernel<false, false, false>, (const void *)bsw_packed_kernel<false, true, false>,
                          (const void *)bsw_packed_kernel<false, false, true>, (const void *)bsw_packed_kernel<false, true, true>,
                          (const void *)bsw_packed_kernel<true, false, true>, (const void *)bsw_packed_kernel<true, true, true> };
    for (const void *fn : fns) {
        CK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
        CK(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    }
    g_rt[dev] = r;
    return *r;
}

DevCtx *ctx_acquire(DevRt &rt)
{
    {
        std::lock_guard<std::mutex> lk(rt.mu);
        if (!rt.free_ctx.empty()) { DevCtx *x = rt.free_ctx.back(); rt.free_ctx.pop_back(); return x; }
    }
    DevCtx *x = new DevCtx();
    x->dev = rt.dev;
    CK(cudaSetDevice(rt.dev));
    CK(cudaEventCreate(&x->ev0)); CK(cudaEventCreate(&x->ev1));
    CK(cudaEventCreateWithFlags(&x->fork_ev, cudaEventDisableTiming));
    for (int qb = 1; qb <= BSW_NBINS; ++qb) CK(cudaEventCreateWithFlags(&x->bin_ev[qb], cudaEventDisableTiming));
    for (Slot &sl : x->slot) {
        CK(cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&sl.dev_done, cudaEventDisableTiming));
        CK(cudaHostAlloc((void **)&sl.h_plan, sizeof(Plan), cudaHostAllocDefault));
        CK(cudaHostAlloc((void **)&sl.h_count, 4 * sizeof(int), cudaHostAllocDefault));
        CK(cudaHostAlloc((void **)&sl.h_tot, 4 * sizeof(long long), cudaHostAllocDefault));
        sl.plan.reserve(sizeof(Plan));
    }
    CK(cudaStreamCreateWithFlags(&x->arena_stream, cudaStreamNonBlocking));
    for (cudaEvent_t &e : x->arena_ev) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    x->cells.reserve(sizeof(unsigned long long));
    CK(cudaMemset(x->cells.p, 0, sizeof(unsigned long long)));
    return x;
}

/* back to the pool (buffers, streams and mailboxes are kept for the next handle); all of its work must be done */
void ctx_release(DevRt &rt, DevCtx *x)
{
    CK(cudaSetDevice(rt.dev));
    for (Slot &sl : x->slot) { CK(cudaStreamSynchronize(sl.stream)); sl.busy = false; sl.dev_used = false; }
    CK(cudaStreamSynchronize(x->arena_stream));
    CK(cudaMemset(x->cells.p, 0, sizeof(unsigned long long)));
    x->dev_rr = 0;
    std::lock_guard<std::mutex> lk(rt.mu);
    rt.free_ctx.push_back(x);
}

/* one host-pointer BSW call on the direct route (direct_call_bsw_multi) */
struct DirectSeg { bsw_seqpair_t *pairs; const uint8_t *ref, *qer; int n, w, tier8_check; };

struct Core {
    Const c;
    std::vector<int> devs;
    std::vector<DevCtx *> ctx;       /* per device of the handle; created on first large / resident call */
    bsw_stats_t st;
    int timing = 0;
    int count_cells = 0;
    int trace = 0;                   /* BSW_CUDA_TRACE: host-side timing of every direct call on stderr */
    int chunk_pairs = 1 << 19;
    bool chunk_ramp = true;                       /* direct calls: chunk sizes ramp up / down at the ends (chunk_starts) */
    int kswv_no_w = 0;                            /* BSW_CUDA_KSWV_NO_W=1: resident kswv calls keep the 16-bit shapes off (probe) */
    int kswv_lanes = 0;                           /* kswv_sys.cuh: lanes per group (8 / 16 / 32); 0 = by batch size */
    int ahead = 2;
    int direct_min = 1 << 17;        /* host-pointer calls with fewer pairs go through the staged engine */
    long long direct_bytes = 8ll << 20;  /* ... unless they carry at least this many bases (call_is_heavy) */
    int wide_max = 32768;            /* calls / batches up to this many pairs run on the lanes-per-pair kernel */
    int wide_lat = 32768;             /* ... and up to this many on its shapes with more lanes per pair (latency) */
    int wide_split = 0;              /* staged batches in between: bins up to this index stay on the thread kernel */
    int mode = MODE_BSW;
    /* arenas declared with bsw_arena_begin */
    const uint8_t *arena_ref_h = nullptr, *arena_qer_h = nullptr;
    size_t arena_ref_n = 0, arena_qer_n = 0;
    size_t arena_ref_piece = 0, arena_qer_piece = 0;     /* bytes per upload piece */
    bool arena_on = false;
    bool arena_defer = false;                     /* bsw_arena_defer: large calls of a bracket complete at bsw_arena_end */
    std::vector<DirectSeg> pending;               /* ... and these are the ones waiting for it */

    DevRt &rt(int i) { return devrt(devs[i]); }
    DevCtx &cx(int i)
    {
        if (!ctx[i]) ctx[i] = ctx_acquire(rt(i));
        return *ctx[i];
    }
};

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
    k.devs = pick_devices();
    k.ctx.assign(k.devs.size(), nullptr);
    for (size_t i = 0; i < k.devs.size(); ++i) (void)devrt(k.devs[i]);       /* first handle on a device builds its runtime */
    k.count_cells = env_int("BSW_CUDA_COUNT_CELLS", 0);
    k.chunk_pairs = env_int("BSW_CUDA_CHUNK_PAIRS", 1 << 19);
    k.kswv_lanes = env_int("BSW_CUDA_KSWV_LANES", 0);
    k.kswv_no_w = env_int("BSW_CUDA_KSWV_NO_W", 0) ? 1 : 0;
    k.chunk_ramp = env_int("BSW_CUDA_CHUNK_RAMP", 1) != 0;
    k.ahead = env_int("BSW_CUDA_AHEAD", 2);
    k.trace = env_int("BSW_CUDA_TRACE", 0);
    k.direct_min = env_int("BSW_CUDA_DIRECT_MIN", 1 << 17);
    k.direct_bytes = (long long)std::max(1, env_int("BSW_CUDA_DIRECT_MB", 8)) << 20;
    k.wide_max = env_int("BSW_CUDA_WIDE_MAX", 32768);
    k.wide_lat = env_int("BSW_CUDA_WIDE_LAT", 32768);
    k.wide_split = env_int("BSW_CUDA_WIDE_SPLIT", 0);
}

void core_destroy(Core &k)
{
    for (size_t i = 0; i < k.ctx.size(); ++i)
        if (k.ctx[i]) { ctx_release(k.rt((int)i), k.ctx[i]); k.ctx[i] = nullptr; }
}

Const make_const(int o_del, int e_del, int o_ins, int e_ins, int zdrop, int end_bonus, int w_match, int w_mismatch)
{
    Const c;
    memset(&c, 0, sizeof c);
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

[[noreturn]] void die_contract(int bad)
{
    fprintf(stderr, "[bsw-b200] fatal: %d pair(s) violate the call contract (negative length, h0 <= 0, or "
                    ">= %d bases in getScores8)\n", bad, BSW_MAX_SEQ_LEN8);
    exit(EXIT_FAILURE);
}

/* ================================================================ large / resident calls: prep, sort, per-bin launches */

/* stages 1-3 + plan read-back (asynchronous) */
void plan_enqueue(Core &k, Slot &x, const bsw_seqpair_t *d_pairs, const uint8_t *d_qer, int n, cudaStream_t s, int tier8_check,
                  long long rlo = 0, long long rhi = (long long)1 << 62, long long qlo = 0, long long qhi = (long long)1 << 62,
                  bool clear_plan = true)
{
    x.keys_a.reserve((size_t)n * 4); x.keys_b.reserve((size_t)n * 4);
    x.idx_a.reserve((size_t)n * 4);  x.idx_b.reserve((size_t)n * 4);
    size_t tmp_bytes = 0;
    CK(cub::DeviceRadixSort::SortPairsDescending(nullptr, tmp_bytes, (uint32_t *)nullptr, (uint32_t *)nullptr,
                                                 (uint32_t *)nullptr, (uint32_t *)nullptr, n, 0, BSW_KEY_BITS, s));
    x.cubtmp.reserve(tmp_bytes);
    Plan *d_plan = (Plan *)x.plan.p;
    if (clear_plan) CK(cudaMemsetAsync(d_plan, 0, sizeof(Plan), s));     /* else: cleared before ref_expand_kernel reported into it */
    const int max_elems = BSW_NBINS * 16;
    prep_kernel<<<(n + 255) / 256, 256, 0, s>>>(d_pairs, d_qer, n, k.c, k.mode, max_elems, tier8_check, k.count_cells ? 0 : 1, rlo, rhi, qlo, qhi,
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
    if (x.h_plan->bad_pairs) die_contract(x.h_plan->bad_pairs);
    return x.h_plan;
}

const Plan *run_plan(Core &k, Slot &x, const bsw_seqpair_t *d_pairs, const uint8_t *d_qer, int n, cudaStream_t s, int tier8_check)
{
    plan_enqueue(k, x, d_pairs, d_qer, n, s, tier8_check);
    return plan_wait(x, s);
}

/* Priority level (0 = highest) of every non-empty bin of a call: longest queries first, one level per bin while the
 * device has levels left; otherwise neighbouring bins at the two ends of the list share a level, the pair with the
 * smaller estimated work (pairs x columns^2) first. */
void assign_levels(const Plan *pl, int nlev, int level_of[BSW_NBINS + 1])
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

void launch_bsw_dp(Core &k, DevRt &rt, DevCtx &x, Slot &sl, const Plan *pl, bsw_seqpair_t *d_pairs, const uint8_t *d_ref,
                   const uint8_t *d_qer, int n, int w, cudaStream_t s)
{
    int level_of[BSW_NBINS + 1];
    assign_levels(pl, rt.nlevels, level_of);
    const uint32_t *sorted = (const uint32_t *)sl.idx_b.p;
    const PairIO io = pairio_records(d_pairs);
    if (k.timing) CK(cudaEventRecord(x.ev0, s));
    int n_packed = 0;
    CK(cudaEventRecord(x.fork_ev, s));
    unsigned used = 0;
    /* sorted order: pairs whose query holds an N first (key bit 25), then the others; each longest query first.
     * Small calls (n <= BSW_CUDA_WIDE_MAX) are latency-bound in the thread-owns-two-pairs kernel (its longest pair
     * sets the call's duration): their bins go to the lanes-per-pair kernel instead, like the wide-only pairs below. */
    const bool sym = k.c.oe_del == k.c.oe_ins;
    const bool small = n <= k.wide_max && !k.count_cells && wide_eligible(k.c, 0, 0, 1);
    std::vector<WideSeg> segs;             /* everything the lanes-per-pair kernel runs: one launch (bsw_kernels.cuh) */
    for (int pass = 0; pass < 2; ++pass)
        for (int qb = BSW_NBINS; qb >= 1; --qb) {
            const int cnt = pass == 0 ? pl->hist_n[qb] : pl->hist[qb];
            if (cnt <= 0) continue;
            if (small) {
                const int cf = wide_cfg_of_bin(qb);
                if (!segs.empty() && segs.back().cfg == cf && segs.back().first + segs.back().count == n_packed) segs.back().count += cnt;
                else segs.push_back(WideSeg{ cf, n_packed, cnt });
                n_packed += cnt;
                continue;
            }
            const int NE = qb * 16, grid = (cnt + 63) / 64;
            cudaStream_t bs = rt.lvl_stream[level_of[qb]][qb];
            if (!(used & (1u << qb))) CK(cudaStreamWaitEvent(bs, x.fork_ev, 0));
            used |= 1u << qb;
            /* template flags: exact cell counting; symmetric gap sums (one packed add less per cell); general score
             * table for queries with an N (the counting variant always uses it) */
            auto kern = k.count_cells ? (sym ? bsw_packed_kernel<true, true, true> : bsw_packed_kernel<true, false, true>)
                      : pass == 0     ? (sym ? bsw_packed_kernel<false, true, true> : bsw_packed_kernel<false, false, true>)
                                      : (sym ? bsw_packed_kernel<false, true, false> : bsw_packed_kernel<false, false, false>);
            kern<<<grid, 32, NE * BSW_WARP_ELEM_BYTES, bs>>>(io, d_ref, d_qer, sorted, bintable_single(n_packed, cnt, NE),
                                                            k.c, w, (unsigned long long *)x.cells.p);
            CK(cudaGetLastError());
            k.st.kernel_launches++;
            n_packed += cnt;
        }
    cudaEvent_t trace_ev[BSW_NBINS + 1] = {};
    const bool trace = k.timing && env_int("BSW_CUDA_BIN_TRACE", 0);      /* debug: when does each bin's last block retire? */
    for (int qb = BSW_NBINS; qb >= 1; --qb)
        if (used & (1u << qb)) {
            cudaStream_t bs = rt.lvl_stream[level_of[qb]][qb];
            if (trace) { CK(cudaEventCreate(&trace_ev[qb])); CK(cudaEventRecord(trace_ev[qb], bs)); }
            CK(cudaEventRecord(x.bin_ev[qb], bs));
            CK(cudaStreamWaitEvent(s, x.bin_ev[qb], 0));
        }
    /* pairs only the 16-bit lanes-per-pair kernel admits: sorted after the packed ones, largest configuration first */
    int n_wide = 0;
    for (int cfg = BSW_WIDE_NCFG - 1; cfg >= 0; --cfg) {
        const int cnt = pl->hist_w[cfg];
        if (cnt <= 0) continue;
        segs.push_back(WideSeg{ cfg, n_packed + n_wide, cnt });
        n_wide += cnt;
    }
    if (!segs.empty()) {
        k.st.kernel_launches += launch_wide_segs(segs.data(), (int)segs.size(), n <= k.wide_lat, sym, io, d_ref, d_qer, sorted, k.c, w, s);
        CK(cudaGetLastError());
    }
    if (pl->n_general > 0) {
        /* bounded scratch: at most ~1 GiB of eh rows per launch */
        const size_t per_pair = ((size_t)pl->max_len2_general + 2) * sizeof(int2);
        int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)pl->n_general, ((size_t)1 << 30) / per_pair));
        sl.scratch.reserve((size_t)chunk * per_pair);
        for (int first = 0; first < pl->n_general; first += chunk) {
            int cnt = std::min(chunk, pl->n_general - first);
            bsw_general_kernel<<<(cnt + 63) / 64, 64, 0, s>>>(io, d_ref, d_qer, sorted, n_packed + n_wide + first, cnt,
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
    k.st.pairs += n; k.st.pairs_general += pl->n_general;
    if (small) k.st.pairs_wide += n_packed + n_wide; else { k.st.pairs_packed += n_packed; k.st.pairs_wide += n_wide; }
}

void launch_bsw(Core &k, DevRt &rt, DevCtx &x, Slot &sl, bsw_seqpair_t *d_pairs, const uint8_t *d_ref, const uint8_t *d_qer,
                int n, int w, cudaStream_t s, int tier8_check)
{
    if (n <= 0) return;
    const Plan *pl = run_plan(k, sl, d_pairs, d_qer, n, s, tier8_check);
    launch_bsw_dp(k, rt, x, sl, pl, d_pairs, d_ref, d_qer, n, w, s);
}

/* kswv DP launches of a planned call: one registers-and-shuffles launch (kswv_sys.cuh) per non-empty query-length bin on
 * the bin's stream -- queries with an N first (key bit 25; the general score table), each longest query first -- then
 * the general 32-bit kernel for what prep did not admit. */
void launch_kswv_dp(Core &k, DevRt &rt, DevCtx &x, Slot &sl, const Plan *pl, bsw_seqpair_t *d_pairs, const uint8_t *d_ref,
                    const uint8_t *d_qer, bsw_kswr_t *d_aln, int n, int is8, int phase, cudaStream_t s)
{
    const uint32_t *sorted = (const uint32_t *)sl.idx_b.p;
    const PairIO io = pairio_records(d_pairs);
    if (k.timing) CK(cudaEventRecord(x.ev0, s));
    int n_packed = 0;
    const bool sym = k.c.oe_del == k.c.oe_ins;
    int lanes = k.kswv_lanes ? k.kswv_lanes : kswv_sys_lanes_for(n);
    if (!sym) lanes = 8;
    if (lanes > 8) {
        /* a small call: ONE launch over all packed pairs on the shape with more lanes per group (see engine.cuh) */
        int any_n = 0, ml2 = 1;
        for (int qb = 1; qb <= BSW_NBINS; ++qb) {
            n_packed += pl->hist_n[qb] + pl->hist[qb]; any_n += pl->hist_n[qb];
            if (pl->hist_n[qb] + pl->hist[qb] > 0) ml2 = std::max(ml2, std::min(qb * 16, std::max(1, pl->max_len2_bin[qb])));
        }
        if (n_packed > 0) {
            launch_kswv_sys(ml2, sym, any_n > 0, lanes, 0, n_packed, std::max(1, pl->max_len1_packed), io, d_ref, d_qer, d_aln, sorted,
                            k.c, is8, phase, (unsigned long long *)x.cells.p, s);
            CK(cudaGetLastError());
            k.st.kernel_launches++;
        }
    } else {
        unsigned used = 0;
        CK(cudaEventRecord(x.fork_ev, s));
        for (int pass = 0; pass < 2; ++pass)
            for (int qb = BSW_NBINS; qb >= 1; --qb) {
                const int cnt = pass == 0 ? pl->hist_n[qb] : pl->hist[qb];
                if (cnt <= 0) continue;
                cudaStream_t bs = rt.bin_stream[qb];
                if (!(used & (1u << qb))) CK(cudaStreamWaitEvent(bs, x.fork_ev, 0));
                used |= 1u << qb;
                launch_kswv_sys(std::min(qb * 16, std::max(1, pl->max_len2_bin[qb])), sym, pass == 0, 8, n_packed, cnt,
                                std::max(1, pl->max_len1_packed), io, d_ref, d_qer, d_aln, sorted,
                                k.c, is8, phase, (unsigned long long *)x.cells.p, bs);
                CK(cudaGetLastError());
                k.st.kernel_launches++;
                n_packed += cnt;
            }
        for (int qb = BSW_NBINS; qb >= 1; --qb)
            if (used & (1u << qb)) {
                CK(cudaEventRecord(x.bin_ev[qb], rt.bin_stream[qb]));
                CK(cudaStreamWaitEvent(s, x.bin_ev[qb], 0));
            }
    }
    /* the 16-bit tier: sorted after the byte-key pairs, largest configuration first */
    int n_w = 0;
    for (int cfg = KSWV_SYSW_NCFG - 1; cfg >= 0; --cfg) {
        const int cnt = pl->hist_w[cfg];
        if (cnt <= 0) continue;
        launch_kswv_sysw(cfg, sym, n_packed + n_w, cnt, std::max(1, pl->max_len1_packed), io, d_ref, d_qer, d_aln, sorted, k.c, is8, phase,
                         (unsigned long long *)x.cells.p, s);
        CK(cudaGetLastError());
        k.st.kernel_launches++;
        n_w += cnt;
    }
    if (pl->n_general > 0) {
        const size_t per_pair = ((size_t)pl->max_len2_general + 2) * sizeof(int2) + ((size_t)pl->max_len1 + 2) * sizeof(int);
        int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)pl->n_general, ((size_t)1 << 30) / per_pair));
        sl.scratch.reserve((size_t)chunk * per_pair);
        int2 *he = (int2 *)sl.scratch.p;
        int *rm = (int *)((char *)sl.scratch.p + (size_t)chunk * ((size_t)pl->max_len2_general + 2) * sizeof(int2));
        for (int first = 0; first < pl->n_general; first += chunk) {
            int cnt = std::min(chunk, pl->n_general - first);
            kswv_general_kernel<<<(cnt + 63) / 64, 64, 0, s>>>(io, d_ref, d_qer, d_aln, sorted, n_packed + n_w + first, cnt,
                                                               k.c, is8, phase, he, rm, (unsigned long long *)x.cells.p);
            CK(cudaGetLastError());
            k.st.kernel_launches++;
        }
    }
    if (k.timing) {
        CK(cudaEventRecord(x.ev1, s)); CK(cudaEventSynchronize(x.ev1));
        float ms = 0; CK(cudaEventElapsedTime(&ms, x.ev0, x.ev1)); k.st.kernel_ms += ms;
    }
    k.st.pairs += n; k.st.pairs_packed += n_packed; k.st.pairs_wide += n_w; k.st.pairs_general += pl->n_general;
}

/* returns false (nothing launched) when a pair names bases outside [rlo, rhi) / [qlo, qhi): the caller uploaded ranges
 * estimated from a sample of the pairs and has to redo them exactly */
bool launch_kswv(Core &k, DevRt &rt, DevCtx &x, Slot &sl, bsw_seqpair_t *d_pairs, const uint8_t *d_ref, const uint8_t *d_qer,
                 bsw_kswr_t *d_aln, int n, int is8, int phase, cudaStream_t s,
                 long long rlo = 0, long long rhi = (long long)1 << 62, long long qlo = 0, long long qhi = (long long)1 << 62)
{
    if (n <= 0) return true;
    /* kswv: the plan's tier flag = is8 (admission of the 16-bit shapes) */
    plan_enqueue(k, sl, d_pairs, d_qer, n, s, is8 | k.kswv_no_w, rlo, rhi, qlo, qhi);
    const Plan *pl = plan_wait(sl, s);
    if (pl->oob_pairs) return false;
    launch_kswv_dp(k, rt, x, sl, pl, d_pairs, d_ref, d_qer, d_aln, n, is8, phase, s);
    return true;
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
void extents_sampled(const bsw_seqpair_t *p, int first, int cnt, long long &rlo, long long &rhi, long long &qlo, long long &qhi,
                     long long *len1_sum32 = nullptr)
{
    rlo = qlo = (long long)1 << 62; rhi = qhi = 0;
    long long l1 = 0;
    auto take = [&](int i) {
        long long a = p[i].idr, b = a + std::max(p[i].len1, 0), c = p[i].idq, d = c + std::max(p[i].len2, 0);
        rlo = std::min(rlo, a); rhi = std::max(rhi, b); qlo = std::min(qlo, c); qhi = std::max(qhi, d);
    };
    for (int i = first; i < first + cnt; i += 32) { take(i); l1 += std::max(p[i].len1, 0) + 4; }
    for (int i = std::max(first, first + cnt - 32); i < first + cnt; ++i) take(i);
    if (rhi <= rlo) { rlo = 0; rhi = 0; }
    if (qhi <= qlo) { qlo = 0; qhi = 0; }
    if (len1_sum32) *len1_sum32 = l1;
}

void slot_retire(DevCtx &x, Slot &sl)
{
    if (sl.dev_used) { CK(cudaSetDevice(x.dev)); CK(cudaEventSynchronize(sl.dev_done)); sl.dev_used = false; }
    if (!sl.busy) return;
    CK(cudaSetDevice(x.dev));
    CK(cudaEventSynchronize(sl.done));
    sl.busy = false;
}

/* enqueues the copies of the arena pieces [x.arena_enq, piece] on the device's arena stream, each followed by its event */
void arena_pieces_upto(Core &k, DevCtx &x, int piece)
{
    for (; x.arena_enq <= piece; ++x.arena_enq) {
        const int pc = x.arena_enq;
        const size_t r0 = std::min(k.arena_ref_n, k.arena_ref_piece * pc), r1 = pc == BSW_ARENA_PIECES - 1 ? k.arena_ref_n : std::min(k.arena_ref_n, k.arena_ref_piece * (pc + 1));
        const size_t q0 = std::min(k.arena_qer_n, k.arena_qer_piece * pc), q1 = pc == BSW_ARENA_PIECES - 1 ? k.arena_qer_n : std::min(k.arena_qer_n, k.arena_qer_piece * (pc + 1));
        if (r1 > r0) CK(cudaMemcpyAsync((char *)x.arena_ref.p + r0, k.arena_ref_h + r0, r1 - r0, cudaMemcpyHostToDevice, x.arena_stream));
        if (q1 > q0) CK(cudaMemcpyAsync((char *)x.arena_qer.p + q0, k.arena_qer_h + q0, q1 - q0, cudaMemcpyHostToDevice, x.arena_stream));
        CK(cudaEventRecord(x.arena_ev[pc], x.arena_stream));
        k.st.h2d_bytes += (long long)(r1 - r0) + (long long)(q1 - q0);
    }
}

/* Chunk boundaries of a direct call.  A call cannot start computing before its first chunk is up, sorted and planned, and
 * cannot return before its last chunk is computed and home: with equal chunks that is ~2 x (one chunk's upload + DP) of
 * unoverlapped time per call.  So the chunks ramp up from chunk / 8 (at least 16 k pairs) by doubling and ramp down the
 * same way at the end, each size once per device (BSW_CUDA_CHUNK_RAMP=0: equal chunks).  Used by the kswv calls, whose
 * kernel is quick on small batches: C4 e2e step 40.5 -> 37.6 ms. */
std::vector<int> chunk_starts(int n, int chunk, int nd, bool ramp)
{
    std::vector<int> st;
    if (!ramp || n <= chunk) { for (int f = 0; f < n; f += chunk) st.push_back(f); st.push_back(n); return st; }
    const int lo = std::max(16384, chunk >> 3);
    std::vector<int> head, tail;
    long long used = 0;
    for (int sz = lo; sz < chunk; sz <<= 1)
        for (int d = 0; d < nd; ++d) { head.push_back(sz); tail.push_back(sz); used += 2ll * sz; }
    if (used >= n) { for (int f = 0; f < n; f += chunk) st.push_back(f); st.push_back(n); return st; }
    int f = 0;
    for (int sz : head) { st.push_back(f); f += sz; }
    long long tail_sum = used / 2;
    while (n - f - tail_sum > 0) { st.push_back(f); f += (int)std::min<long long>(chunk, n - f - tail_sum); }
    for (size_t i = tail.size(); i-- > 0;) { st.push_back(f); f += tail[i]; }
    st.push_back(n);
    return st;
}

/* ---- direct BSW call: chunked, pipelined, sharded ---------------------------------------------
 * The pair array is cut into chunks of BSW_CUDA_CHUNK_PAIRS pairs, dealt round-robin to the handle's
 * devices.  Each chunk uploads its 56-byte records and only the byte ranges of the sequence buffers its pairs
 * touch (nothing when the buffers lie in a declared arena, bsw_arena_begin; no target bases at all when the
 * targets come from the resident reference, `rpos` != nullptr), so the H2D copy of chunk c+1 and the D2H copy of
 * chunk c-1 overlap the kernels of chunk c.  Results land directly in the caller's SeqPair array (whole-record
 * DMA: a 24-of-56-byte strided copy reaches 1.5 GB/s up and 15 GB/s down against 55 GB/s for the contiguous
 * records, profiles/r02_pcie_probe_1gpu.json).  Synchronous on return. */
/* One or more direct calls in ONE chunk pipeline.  Several segments = the deferred calls of an arena bracket
 * (bsw_arena_defer): their chunks are interleaved in proportion, so that a call that mostly waits for PCIe (the first
 * tier call of a side: it brings the arenas up) and one that mostly waits for the GPU (the second: only records go up)
 * fill each other's gaps. */
void direct_call_bsw_multi(Core &k, const std::vector<DirectSeg> &segs, const bsw_ref *rref = nullptr, const int64_t *rpos = nullptr)
{
    const int nd = (int)k.devs.size();
    const int chunk = std::max(1024, k.chunk_pairs);
    /* equal chunks: a warp of the thread-owns-two-pairs kernel lives 0.4-0.9 ms whatever the chunk holds, so small ramp chunks only
     * add sorts and tails (measured on c2: 53.1 ms ramped vs 49.6 ms) */
    struct Job { int seg, first, cnt, di; long long rlo, qlo; const uint8_t *d_ref, *d_qer; DevCtx *x; Slot *sl; };
    std::vector<Job> jobs;
    {
        /* chunks of every segment, merged by their relative position inside their segment */
        std::vector<std::pair<double, Job>> all;
        for (size_t sg = 0; sg < segs.size(); ++sg) {
            const std::vector<int> cs = chunk_starts(segs[sg].n, chunk, nd, k.chunk_ramp && env_int("BSW_CUDA_CHUNK_RAMP_BSW", 0));
            const int nc = (int)cs.size() - 1;
            for (int ci = 0; ci < nc; ++ci) {
                Job j{}; j.seg = (int)sg; j.first = cs[ci]; j.cnt = cs[ci + 1] - cs[ci];
                all.push_back({ (ci + 0.5) / nc, j });
            }
        }
        std::stable_sort(all.begin(), all.end(), [](const std::pair<double, Job> &a, const std::pair<double, Job> &b) { return a.first < b.first; });
        for (auto &e : all) jobs.push_back(e.second);
    }
    const int nchunks = (int)jobs.size();
    long long n = 0;
    for (const DirectSeg &sg : segs) n += sg.n;
    auto upload = [&](Job &j, bool exact) {
        bsw_seqpair_t *pairs = segs[j.seg].pairs;
        const uint8_t *ref = segs[j.seg].ref, *qer = segs[j.seg].qer;
        const int tier8_check = segs[j.seg].tier8_check;
        const bool in_arena_r = k.arena_on && !rref && ref >= k.arena_ref_h && ref < k.arena_ref_h + k.arena_ref_n;
        const bool in_arena_q = k.arena_on && qer >= k.arena_qer_h && qer < k.arena_qer_h + k.arena_qer_n;
        long long rlo, rhi, qlo, qhi, l1s = 0;
        if (exact) { extents(pairs, j.first, j.cnt, rlo, rhi, qlo, qhi); for (int i = j.first; i < j.first + j.cnt; ++i) l1s += std::max(pairs[i].len1, 0) + 4; }
        else extents_sampled(pairs, j.first, j.cnt, rlo, rhi, qlo, qhi, &l1s);
        Slot &sl = *j.sl;
        DevCtx &x = *j.x;
        const bsw_seqpair_t *d_pairs = (const bsw_seqpair_t *)sl.pairs.p;
        /* query bases */
        /* Arenas go up in BSW_ARENA_PIECES pieces behind the calls (bsw_arena_begin): the chunk waits for the piece that
         * holds the last byte its (sampled) extents name, and tells prep_kernel that only the pieces up to there are valid,
         * so a pair beyond them redoes the chunk with the exact extents instead of reading bytes still in flight. */
        int piece = -1;
        if (in_arena_q) {
            const long long hi = (qer - k.arena_qer_h) + std::max(qhi, 1ll) - 1;          /* arena offset of the last byte */
            piece = std::max(piece, (int)std::min<long long>(BSW_ARENA_PIECES - 1, hi / (long long)k.arena_qer_piece));
        }
        if (in_arena_r) {
            const long long hi = (ref - k.arena_ref_h) + std::max(rhi, 1ll) - 1;
            piece = std::max(piece, (int)std::min<long long>(BSW_ARENA_PIECES - 1, hi / (long long)k.arena_ref_piece));
        }
        if (piece >= 0) { arena_pieces_upto(k, x, piece); CK(cudaStreamWaitEvent(sl.stream, x.arena_ev[piece], 0)); }
        auto arena_valid = [&](size_t piece_bytes, size_t total) {
            return (long long)std::min<unsigned long long>((unsigned long long)total, (unsigned long long)piece_bytes * (unsigned long long)(piece + 1));
        };
        if (in_arena_q) {
            j.d_qer = (const uint8_t *)x.arena_qer.p + (qer - k.arena_qer_h);
            qlo = k.arena_qer_h - qer; qhi = qlo + (piece == BSW_ARENA_PIECES - 1 ? (long long)k.arena_qer_n : arena_valid(k.arena_qer_piece, k.arena_qer_n));
        } else {
            sl.qer.reserve((size_t)(qhi - qlo) + 64);
            if (qhi > qlo) CK(cudaMemcpyAsync(sl.qer.p, qer + qlo, (size_t)(qhi - qlo), cudaMemcpyHostToDevice, sl.stream));
            k.st.h2d_bytes += qhi - qlo;
            j.d_qer = (const uint8_t *)sl.qer.p - qlo;
        }
        /* target bases */
        if (rref) {
            /* expand the targets from the resident 2-bit reference into a dense scratch; the records' idr are rewritten on
             * the device to point into it.  Its size is estimated from the sampled lengths (x 1.25) unless `exact`;
             * the kernel reports an overflow through Plan::oob_pairs, like a pair outside the sampled byte ranges. */
            const long long cap = exact ? l1s + 64 : (l1s * 32 * 5) / 4 + 65536;
            CK(cudaMemsetAsync(sl.plan.p, 0, sizeof(Plan), sl.stream));
            sl.reftmp.reserve((size_t)cap + 64);
            sl.rpos.reserve((size_t)j.cnt * 8); sl.toff.reserve((size_t)j.cnt * 4 + 4);
            CK(cudaMemcpyAsync(sl.rpos.p, rpos + j.first, (size_t)j.cnt * 8, cudaMemcpyHostToDevice, sl.stream));
            k.st.h2d_bytes += (long long)j.cnt * 8;
            size_t tmp_bytes = 0;
            auto it = thrust::make_transform_iterator(thrust::make_counting_iterator(0), Len1Rounded{ d_pairs });
            CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, it, (int *)sl.toff.p, j.cnt, sl.stream));
            sl.cubtmp.reserve(tmp_bytes);
            CK(cub::DeviceScan::ExclusiveSum(sl.cubtmp.p, tmp_bytes, it, (int *)sl.toff.p, j.cnt, sl.stream));
            const uint8_t *d_pac = rref->on(x.dev);
            if (!d_pac) die("the reference was not uploaded to a device this handle uses");
            ref_expand_kernel<<<(j.cnt + 31) / 32, 256, 0, sl.stream>>>((bsw_seqpair_t *)sl.pairs.p, j.cnt, (const long long *)sl.rpos.p,
                                                                       (int *)sl.toff.p, d_pac, rref->l_pac, (uint8_t *)sl.reftmp.p,
                                                                       cap, (Plan *)sl.plan.p);
            CK(cudaGetLastError());
            k.st.kernel_launches += 3;
            j.d_ref = (const uint8_t *)sl.reftmp.p;
            rlo = 0; rhi = cap;
        } else if (in_arena_r) {
            j.d_ref = (const uint8_t *)x.arena_ref.p + (ref - k.arena_ref_h);
            rlo = k.arena_ref_h - ref; rhi = rlo + (piece == BSW_ARENA_PIECES - 1 ? (long long)k.arena_ref_n : arena_valid(k.arena_ref_piece, k.arena_ref_n));
        } else {
            sl.ref.reserve((size_t)(rhi - rlo) + 64);
            if (rhi > rlo) CK(cudaMemcpyAsync(sl.ref.p, ref + rlo, (size_t)(rhi - rlo), cudaMemcpyHostToDevice, sl.stream));
            k.st.h2d_bytes += rhi - rlo;
            j.d_ref = (const uint8_t *)sl.ref.p - rlo;
        }
        plan_enqueue(k, sl, d_pairs, j.d_qer, j.cnt, sl.stream, tier8_check, rlo, rhi, qlo, qhi, !rref);
    };
    /* stage A: extents, uploads, prep + sort + histogram read-back -- all asynchronous */
    auto stage_a = [&](int ci) {
        Job &j = jobs[ci];
        bsw_seqpair_t *pairs = segs[j.seg].pairs;
        j.di = ci % nd;
        j.x = &k.cx(j.di); j.sl = &j.x->slot[(ci / nd) % BSW_NSLOT];
        slot_retire(*j.x, *j.sl);              /* its previous chunk must have landed */
        Slot &sl = *j.sl;
        CK(cudaSetDevice(j.x->dev));
        sl.pairs.reserve((size_t)j.cnt * sizeof(bsw_seqpair_t));
        CK(cudaMemcpyAsync(sl.pairs.p, pairs + j.first, (size_t)j.cnt * sizeof(bsw_seqpair_t), cudaMemcpyHostToDevice, sl.stream));
        k.st.h2d_bytes += (int64_t)j.cnt * sizeof(bsw_seqpair_t);
        upload(j, false);
        sl.busy = true;
    };
    /* stage B: wait for the histogram, launch the DP kernels, download the results */
    auto stage_b = [&](int ci) {
        Job &j = jobs[ci];
        Slot &sl = *j.sl;
        bsw_seqpair_t *pairs = segs[j.seg].pairs;
        const int w = segs[j.seg].w;
        CK(cudaSetDevice(j.x->dev));
        const Plan *pl = plan_wait(sl, sl.stream);
        if (pl->oob_pairs) {                      /* the sampled ranges missed a pair: redo with the exact scan */
            if (rref) CK(cudaMemcpyAsync(sl.pairs.p, pairs + j.first, (size_t)j.cnt * sizeof(bsw_seqpair_t), cudaMemcpyHostToDevice, sl.stream));
            upload(j, true);
            pl = plan_wait(sl, sl.stream);
            if (pl->oob_pairs) die("a pair names bases outside its sequence buffer / declared arena");
        }
        /* the kernels index the sequence buffers with the caller's idr/idq: the pointers were rebased in upload() */
        launch_bsw_dp(k, k.rt(j.di), *j.x, sl, pl, (bsw_seqpair_t *)sl.pairs.p, j.d_ref, j.d_qer, j.cnt, w, sl.stream);
        if (rref) {                                /* the caller's idr come back untouched */
            restore_idr_kernel<<<(j.cnt + 255) / 256, 256, 0, sl.stream>>>((bsw_seqpair_t *)sl.pairs.p, j.cnt, (const int *)sl.toff.p);
            k.st.kernel_launches++;
        }
        CK(cudaMemcpyAsync(pairs + j.first, sl.pairs.p, (size_t)j.cnt * sizeof(bsw_seqpair_t), cudaMemcpyDeviceToHost, sl.stream));
        CK(cudaEventRecord(sl.done, sl.stream));
        k.st.d2h_bytes += (int64_t)j.cnt * sizeof(bsw_seqpair_t);
    };
    /* chunk c+nd is uploading while chunk c waits for its histogram: the copy engine never idles */
    const int ahead = nd * std::max(1, std::min(BSW_NSLOT - 2, k.ahead));   /* chunks uploading ahead of the one waiting */
    const bool trace = k.trace != 0;
    long long t_a = 0, t_b = 0, t0 = trace ? now_ns() : 0;
    for (int ci = 0; ci < std::min(ahead, nchunks); ++ci) stage_a(ci);
    if (trace) t_a += now_ns() - t0;
    for (int ci = 0; ci < nchunks; ++ci) {
        long long t1 = trace ? now_ns() : 0;
        if (ci + ahead < nchunks) stage_a(ci + ahead);
        long long t2 = trace ? now_ns() : 0;
        stage_b(ci);
        if (trace) { t_a += t2 - t1; t_b += now_ns() - t2; }
    }
    long long t3 = trace ? now_ns() : 0;
    for (size_t i = 0; i < k.ctx.size(); ++i)
        if (k.ctx[i]) for (Slot &sl : k.ctx[i]->slot) slot_retire(*k.ctx[i], sl);
    if (trace)
        fprintf(stderr, "[bsw-b200] direct call: %d pairs, %d chunks, total %.3f ms: stage A (retire slot + enqueue uploads, prep, sort) %.3f, "
                        "stage B (wait histogram + launch DP + enqueue download) %.3f, final drain %.3f\n",
                (int)n, nchunks, (now_ns() - t0) * 1e-6, t_a * 1e-6, t_b * 1e-6, (now_ns() - t3) * 1e-6);
}

void direct_call_bsw(Core &k, bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, int n, int w, int tier8_check,
                     const bsw_ref *rref = nullptr, const int64_t *rpos = nullptr)
{
    direct_call_bsw_multi(k, std::vector<DirectSeg>{ DirectSeg{ pairs, ref, qer, n, w, tier8_check } }, rref, rpos);
}

/* ---- direct kswv call: chunks of BSW_CUDA_KSWV_CHUNK pairs, pipelined over the slots like BSW -----
 * phase 0 writes aln[i]; phase 1 refines aln[regid]: the device computes the reverse pass of every pair into a dense
 * per-chunk result array and the host applies the rule (score equal -> tb, qb), so no aln span goes up. */
void direct_call_kswv(Core &k, bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, bsw_kswr_t *aln,
                      int n, int is8, int phase)
{
    const int nd = (int)k.devs.size();
    const int chunk = std::max(1024, env_int("BSW_CUDA_KSWV_CHUNK", 1 << 18));
    const std::vector<int> cstart = chunk_starts(n, chunk, nd, k.chunk_ramp);
    const int nchunks = (int)cstart.size() - 1;
    struct Job { int first, cnt, di; long long rlo, qlo; DevCtx *x; Slot *sl; bsw_kswr_t *h_res; bool landed; };
    std::vector<Job> jobs(nchunks);
    auto finish = [&](Job &j) {              /* chunk landed: phase 1 applies the refinement rule on the host */
        if (j.landed) return;
        slot_retire(*j.x, *j.sl);
        if (phase == 1)
            for (int i = 0; i < j.cnt; ++i) {
                const bsw_kswr_t &rr = j.h_res[i];
                bsw_kswr_t *a = &aln[pairs[j.first + i].regid];
                if (a->score == rr.score) { a->tb = a->te - rr.te; a->qb = a->qe - rr.qe; }
            }
        j.landed = true;
    };
    auto upload = [&](Job &j, bool exact) {
        Slot &sl = *j.sl;
        long long rlo, rhi, qlo, qhi;
        if (exact) extents(pairs, j.first, j.cnt, rlo, rhi, qlo, qhi);
        else extents_sampled(pairs, j.first, j.cnt, rlo, rhi, qlo, qhi);
        sl.pairs.reserve((size_t)j.cnt * sizeof(bsw_seqpair_t));
        sl.ref.reserve((size_t)(rhi - rlo) + 64); sl.qer.reserve((size_t)(qhi - qlo) + 64);
        sl.aln.reserve((size_t)j.cnt * sizeof(bsw_kswr_t));
        CK(cudaMemcpyAsync(sl.pairs.p, pairs + j.first, (size_t)j.cnt * sizeof(bsw_seqpair_t), cudaMemcpyHostToDevice, sl.stream));
        if (rhi > rlo) CK(cudaMemcpyAsync(sl.ref.p, ref + rlo, (size_t)(rhi - rlo), cudaMemcpyHostToDevice, sl.stream));
        if (qhi > qlo) CK(cudaMemcpyAsync(sl.qer.p, qer + qlo, (size_t)(qhi - qlo), cudaMemcpyHostToDevice, sl.stream));
        k.st.h2d_bytes += (int64_t)j.cnt * sizeof(bsw_seqpair_t) + (rhi - rlo) + (qhi - qlo);
        j.rlo = rlo; j.qlo = qlo;
        plan_enqueue(k, sl, (const bsw_seqpair_t *)sl.pairs.p, (const uint8_t *)sl.qer.p - qlo, j.cnt, sl.stream, is8, rlo, rhi, qlo, qhi);
    };
    /* stage A: the slot's previous chunk lands, then uploads + prep + sort + histogram read-back are enqueued */
    auto stage_a = [&](int ci) {
        Job &j = jobs[ci];
        j.first = cstart[ci]; j.cnt = cstart[ci + 1] - cstart[ci];
        j.di = ci % nd; j.landed = false;
        j.x = &k.cx(j.di); j.sl = &j.x->slot[(ci / nd) % BSW_NSLOT];
        if (ci >= nd * BSW_NSLOT) finish(jobs[ci - nd * BSW_NSLOT]);
        CK(cudaSetDevice(j.x->dev));
        upload(j, false);
        j.sl->busy = true;
    };
    /* stage B: histogram in, DP kernels, results on their way home */
    auto stage_b = [&](int ci) {
        Job &j = jobs[ci];
        Slot &sl = *j.sl;
        CK(cudaSetDevice(j.x->dev));
        const Plan *pl = plan_wait(sl, sl.stream);
        if (pl->oob_pairs) {
            upload(j, true);
            pl = plan_wait(sl, sl.stream);
            if (pl->oob_pairs) die("a pair names bases outside its sequence buffer");
        }
        /* the device always stores the raw pass result of pair i at chunk-local index i (phase-0 store rule) */
        launch_kswv_dp(k, k.rt(j.di), *j.x, sl, pl, (bsw_seqpair_t *)sl.pairs.p, (const uint8_t *)sl.ref.p - j.rlo,
                       (const uint8_t *)sl.qer.p - j.qlo, (bsw_kswr_t *)sl.aln.p, j.cnt, is8, 0, sl.stream);
        if (phase == 0) j.h_res = aln + j.first;
        else { sl.h_res.reserve((size_t)chunk * sizeof(bsw_kswr_t)); j.h_res = (bsw_kswr_t *)sl.h_res.p; }
        CK(cudaMemcpyAsync(j.h_res, sl.aln.p, (size_t)j.cnt * sizeof(bsw_kswr_t), cudaMemcpyDeviceToHost, sl.stream));
        CK(cudaEventRecord(sl.done, sl.stream));
        k.st.d2h_bytes += (int64_t)j.cnt * sizeof(bsw_kswr_t);
    };
    /* like the BSW path: chunk c + ahead is uploading while chunk c waits for its histogram */
    const int ahead = nd * std::max(1, std::min(BSW_NSLOT - 2, k.ahead));
    for (int ci = 0; ci < std::min(ahead, nchunks); ++ci) stage_a(ci);
    for (int ci = 0; ci < nchunks; ++ci) {
        if (ci + ahead < nchunks) stage_a(ci + ahead);
        stage_b(ci);
    }
    for (int ci = std::max(0, nchunks - nd * BSW_NSLOT); ci < nchunks; ++ci) finish(jobs[ci]);
}

/* ================================================================ staged calls (engine.cuh) */

/* the device with the fewest pairs outstanding in its engine */
int pick_engine(Core &k)
{
    int best = 0; long long load = -1;
    for (int i = 0; i < (int)k.devs.size(); ++i) {
        const long long l = k.rt(i).eng().outstanding.load(std::memory_order_relaxed);
        if (load < 0 || l < load) { load = l; best = i; }
    }
    return best;
}

inline bool bytes_have_n(const uint8_t *p, int len)
{
    uint8_t acc = 0;
    for (int j = 0; j < len; ++j) acc |= p[j];
    return (acc & 0xfc) != 0;
}

/* class of one pair: the same rules prep_kernel applies on the device (bsw_kernels.cuh) */
inline int classify(const Core &k, const bsw_seqpair_t &p, const uint8_t *qer, int is8)
{
    const bool ok = k.mode == MODE_BSW ? packed_eligible(k.c, p.len1, p.len2, p.h0) : kswv_sys_eligible(k.c, p.len1, p.len2);
    int qb = (p.len2 + (k.mode == MODE_BSW ? BSW_SWEEP_SLACK : 0) + 15) >> 4;
    qb = qb < 1 ? 1 : qb;
    if (!ok || qb > BSW_NBINS) {
        if (k.mode == MODE_BSW && wide_eligible(k.c, p.len1, p.len2, p.h0)) return CLS_WIDE0 + wide_cfg_for(p.len2);
        if (k.mode == MODE_KSWV && kswv_sysw_eligible(k.c, is8, p.len1, p.len2)) return CLS_WIDE0 + kswv_sysw_cfg_for(p.len2);
        return CLS_GENERAL;
    }
    if (p.len2 > 0 && bytes_have_n(qer + p.idq, p.len2)) return BSW_NBINS + qb;
    return qb;
}

/* one host-pointer call through the engine, in parts that fit a staging slot */
void staged_call(Core &k, bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, bsw_kswr_t *aln, int n, int w,
                 int tier8_check, int is8, int phase)
{
    Engine &eng = k.rt(pick_engine(k)).eng();
    const BatchKey key = make_key(k.mode, k.c, w, is8, k.wide_max, k.wide_lat, k.wide_split);
    std::vector<uint8_t> cls;
    int first = 0;
    while (first < n) {
        /* pass 1 over the part: contract, classes, byte sums */
        const long long t0 = now_ns();
        int cnt[BSW_NCLASS]; memset(cnt, 0, sizeof cnt);
        long long bytes = 0; int m = 0, bad = 0, ml1p = 0, ml2g = 0, ml1g = 0;
        cls.clear();
        for (int i = first; i < n && m < eng.cap_pairs; ++i, ++m) {
            const bsw_seqpair_t &p = pairs[i];
            if (p.len1 < 0 || p.len2 < 0 || (k.mode == MODE_BSW && p.h0 <= 0) ||
                (tier8_check && (p.len1 >= BSW_MAX_SEQ_LEN8 || p.len2 >= BSW_MAX_SEQ_LEN8))) { ++bad; cls.push_back(0); continue; }
            const long long b = (long long)p.len1 + p.len2;
            if (bytes + b + 64 > eng.cap_bytes) {
                if (m == 0) die("one pair is larger than a staging slot (raise BSW_CUDA_STAGE_MB)");
                break;
            }
            bytes += b;
            const int c = classify(k, p, qer, is8);
            cls.push_back((uint8_t)c); cnt[c]++;
            if (c == CLS_GENERAL) { ml2g = std::max(ml2g, p.len2); ml1g = std::max(ml1g, p.len1); }
            else if (c < CLS_GENERAL || k.mode == MODE_KSWV) ml1p = std::max(ml1p, p.len1);     /* kswv: the 16-bit classes size the `b` lists too */
        }
        if (bad) die_contract(bad);
        const long long t1 = now_ns();
        Reservation r = eng.reserve(key, m, bytes, cnt, ml1p, ml2g, ml1g);
        StagedSlot &s = *r.slot;
        const long long t2 = now_ns();
        /* pass 2: pack */
        {
            int *rec = reinterpret_cast<int *>(s.h_rec + (size_t)r.pair_base * 24);
            uint8_t *seq = reinterpret_cast<uint8_t *>(s.h_seq);
            long long off = r.byte_base;
            int pos[BSW_NCLASS]; memcpy(pos, r.cls_pos, sizeof pos);
            for (int t = 0; t < m; ++t, rec += 6) {
                const bsw_seqpair_t &p = pairs[first + t];
                const int c = cls[t];
                rec[0] = (int)off; rec[1] = (int)(off + p.len1);
                rec[2] = (int)(((uint32_t)c << 24) | (uint32_t)pos[c]++);
                rec[3] = p.len1; rec[4] = p.len2; rec[5] = p.h0;
                memcpy(seq + off, ref + p.idr, (size_t)p.len1);
                memcpy(seq + off + p.len1, qer + p.idq, (size_t)p.len2);
                off += (long long)p.len1 + p.len2;
            }
        }
        eng.packed(r, m);
        const long long t3 = now_ns();
        eng.wait_done(r);
        const long long t4 = now_ns();
        /* results */
        if (k.mode == MODE_BSW) {
            const char *res = s.h_res + (size_t)r.pair_base * 24;
            for (int t = 0; t < m; ++t) memcpy(&pairs[first + t].score, res + (size_t)t * 24, 24);
            k.st.d2h_bytes += (long long)m * 24;
        } else {
            const bsw_kswr_t *res = reinterpret_cast<const bsw_kswr_t *>(s.h_res) + r.pair_base;
            if (phase == 0) memcpy(aln + first, res, (size_t)m * sizeof(bsw_kswr_t));
            else
                for (int t = 0; t < m; ++t) {
                    bsw_kswr_t *a = &aln[pairs[first + t].regid];
                    if (a->score == res[t].score) { a->tb = a->te - res[t].te; a->qb = a->qe - res[t].qe; }
                }
            k.st.d2h_bytes += (long long)m * (long long)sizeof(bsw_kswr_t);
        }
        eng.release(r, m);
        {
            const long long t5 = now_ns();
            EngineProf &pf = eng.prof;
            pf.scan_ns += t1 - t0; pf.reserve_ns += t2 - t1; pf.pack_ns += t3 - t2; pf.wait_ns += t4 - t3; pf.unpack_ns += t5 - t4;
        }
        k.st.h2d_bytes += (long long)m * 24 + bytes;
        k.st.pairs += m; k.st.pairs_general += cnt[CLS_GENERAL];
        {   /* which kernel served the packed-eligible classes is the engine's choice per batch (engine stats); the handle
             * books them as packed and the wide-only classes as wide */
            int nw = 0; for (int cf = 0; cf < BSW_WIDE_NCFG; ++cf) nw += cnt[CLS_WIDE0 + cf];
            k.st.pairs_wide += nw; k.st.pairs_packed += m - cnt[CLS_GENERAL] - nw;
        }
        first += m;
    }
}

/* The staged route exists to MERGE the small calls of many host threads; its callers pack every base themselves.  A single
 * call that already carries a device batch worth of work -- at least BSW_CUDA_DIRECT_MB of bases (default 8), estimated
 * from 16 evenly spaced pairs -- takes the direct route however few pairs it has (50 k mate-rescue pairs of 250 x 550 bp:
 * 31 ms staged, 3.4 ms direct). */
bool call_is_heavy(const Core &k, const bsw_seqpair_t *pairs, int n)
{
    if (n < 2048) return false;
    long long sum = 0;
    for (int t = 0; t < 16; ++t) { const bsw_seqpair_t &p = pairs[(long long)t * (n - 1) / 15]; sum += std::max(p.len1, 0) + std::max(p.len2, 0); }
    return sum * (long long)n / 16 >= k.direct_bytes;
}

void host_call_bsw(Core &k, bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, int n, int w, int tier8_check)
{
    k.st.calls++;
    if (n <= 0) return;
    if (!pairs || !ref || !qer) die("null buffer passed to getScores");
    if (n >= k.direct_min || k.count_cells || k.timing || call_is_heavy(k, pairs, n)) {
        if (k.arena_on && k.arena_defer && !k.count_cells && !k.timing) {      /* runs with the other calls of the bracket, at bsw_arena_end */
            k.pending.push_back(DirectSeg{ pairs, ref, qer, n, w, tier8_check });
            return;
        }
        direct_call_bsw(k, pairs, ref, qer, n, w, tier8_check);
    }
    else staged_call(k, pairs, ref, qer, nullptr, n, w, tier8_check, 0, 0);
}

void host_call_kswv(Core &k, bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, bsw_kswr_t *aln,
                    int n, int is8, int phase)
{
    k.st.calls++;
    if (n <= 0) return;
    if (!pairs || !ref || !qer || !aln) die("null buffer passed to kswv getScores");
    if (phase == 1)
        for (int i = 0; i < n; ++i) if (pairs[i].regid < 0) die("kswv phase 1: negative regid");
    if (n >= k.direct_min || k.timing || call_is_heavy(k, pairs, n)) direct_call_kswv(k, pairs, ref, qer, aln, n, is8, phase);
    else staged_call(k, pairs, ref, qer, aln, n, 0, 0, is8, phase);
}

} /* namespace */

struct bsw_handle  { Core k; };
struct kswv_handle { Core k; int32_t maxRefLen, maxQerLen; };

extern "C" {

const char *bsw_version(void) { return "bsw-b200 0.2 (sm_100a, packed s16x2 DPX)"; }

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

/* ---- sequence arenas ------------------------------------------------------------------------- */
void bsw_arena_begin(bsw_t *h, const uint8_t *seqBufRef, size_t refBytes, const uint8_t *seqBufQer, size_t qerBytes)
{
    Core &k = h->k;
    if (k.arena_on) die("bsw_arena_begin: an arena is already declared on this handle");
    k.arena_ref_h = seqBufRef; k.arena_ref_n = seqBufRef ? refBytes : 0;
    k.arena_qer_h = seqBufQer; k.arena_qer_n = seqBufQer ? qerBytes : 0;
    /* pieces of at least 4 MB (a smaller copy does not reach the link's bandwidth), at most BSW_ARENA_PIECES of them */
    k.arena_ref_piece = std::max<size_t>((size_t)4 << 20, ((k.arena_ref_n + BSW_ARENA_PIECES - 1) / BSW_ARENA_PIECES + 255) & ~(size_t)255);
    k.arena_qer_piece = std::max<size_t>((size_t)4 << 20, ((k.arena_qer_n + BSW_ARENA_PIECES - 1) / BSW_ARENA_PIECES + 255) & ~(size_t)255);
    for (int i = 0; i < (int)k.devs.size(); ++i) {
        DevCtx &x = k.cx(i);
        CK(cudaSetDevice(x.dev));
        /* the slots may still read the previous arena's copies */
        for (Slot &sl : x.slot) slot_retire(x, sl);
        x.arena_ref.reserve(k.arena_ref_n + 64); x.arena_qer.reserve(k.arena_qer_n + 64);
        /* the pieces go up lazily, each just before the first chunk that needs it (arena_pieces_upto): copies are served in
         * submission order, so enqueueing the whole arena here would put every chunk's records behind it */
        x.arena_enq = 0;
    }
    k.arena_on = true;
}

void bsw_arena_end(bsw_t *h)
{
    Core &k = h->k;
    if (!k.arena_on) return;
    if (!k.pending.empty()) {                      /* the deferred calls of the bracket: one interleaved chunk pipeline */
        const std::vector<DirectSeg> segs = std::move(k.pending);
        k.pending.clear();
        direct_call_bsw_multi(k, segs);
    }
    for (size_t i = 0; i < k.ctx.size(); ++i)
        if (k.ctx[i]) { CK(cudaSetDevice(k.ctx[i]->dev)); CK(cudaStreamSynchronize(k.ctx[i]->arena_stream)); }
    k.arena_on = false;
    k.arena_ref_h = k.arena_qer_h = nullptr; k.arena_ref_n = k.arena_qer_n = 0;
}

void bsw_arena_defer(bsw_t *h, int on)
{
    Core &k = h->k;
    k.arena_defer = on != 0;
    if (!on && !k.pending.empty()) {               /* switched off inside a bracket: what is pending runs now */
        const std::vector<DirectSeg> segs = std::move(k.pending);
        k.pending.clear();
        direct_call_bsw_multi(k, segs);
    }
}

/* ---- resident reference --------------------------------------------------- */
bsw_ref_t *bsw_ref_upload(const uint8_t *pac, int64_t l_pac)
{
    if (!pac || l_pac <= 0) die("bsw_ref_upload: empty reference");
    bsw_ref_t *r = new bsw_ref();
    r->l_pac = l_pac;
    const size_t bytes = (size_t)((l_pac + 3) >> 2);
    for (int d : pick_devices()) {
        (void)devrt(d);
        CK(cudaSetDevice(d));
        RefCopy c; c.dev = d; c.d_pac = nullptr;
        CK(cudaMalloc((void **)&c.d_pac, bytes + 16));
        CK(cudaMemset(c.d_pac + bytes, 0, 16));
        CK(cudaMemcpy(c.d_pac, pac, bytes, cudaMemcpyHostToDevice));
        r->copies.push_back(c);
    }
    return r;
}

void bsw_ref_free(bsw_ref_t *r)
{
    if (!r) return;
    for (RefCopy &c : r->copies) { cudaSetDevice(c.dev); cudaFree(c.d_pac); }
    delete r;
}

void bsw_getScoresRef(bsw_t *h, bsw_seqpair_t *pairArray, const bsw_ref_t *ref, const int64_t *rpos,
                      const uint8_t *seqBufQer, int32_t numPairs, int32_t w)
{
    Core &k = h->k;
    k.st.calls++;
    if (numPairs <= 0) return;
    if (!pairArray || !ref || !rpos || !seqBufQer) die("null buffer passed to bsw_getScoresRef");
    direct_call_bsw(k, pairArray, nullptr, seqBufQer, numPairs, w, 0, ref, rpos);
}

/* Device-resident calls rotate over the slots (sort keys, permutation, histogram mailbox, general-path scratch), so
 * that calls issued on DIFFERENT caller streams overlap: the next call's prep / sort / DP blocks fill the SMs while the
 * previous call's last warps drain (bench step of four 2.5M-pair calls: 18.54 ms on one stream, 18.16 ms on four).
 * Calls on one stream stay ordered by that stream.  A slot is reused only after the call that last used it has finished (stream-side wait). */
static Slot &dev_slot(DevCtx &x, cudaStream_t s, bool rotate)
{
    /* rotate = false pins a caller to slot 0 (kswv_dev_align2: its hand-off scratch grows on first use) */
    Slot &sl = x.slot[rotate ? x.dev_rr : 0];
    if (rotate) x.dev_rr = (x.dev_rr + 1) % BSW_NSLOT;
    if (sl.dev_used) CK(cudaStreamWaitEvent(s, sl.dev_done, 0));
    sl.dev_used = true;
    return sl;
}

void bsw_dev_getScores(bsw_t *h, bsw_seqpair_t *d_pairArray, const uint8_t *d_seqBufRef, const uint8_t *d_seqBufQer,
                       int32_t numPairs, int32_t w, void *stream)
{
    DevCtx &x = h->k.cx(0);
    CK(cudaSetDevice(x.dev));
    h->k.st.calls++;
    Slot &sl = dev_slot(x, (cudaStream_t)stream, true);
    launch_bsw(h->k, h->k.rt(0), x, sl, d_pairArray, d_seqBufRef, d_seqBufQer, numPairs, w, (cudaStream_t)stream, 0);
    CK(cudaEventRecord(sl.dev_done, (cudaStream_t)stream));
}

/* One side (left or right extensions) of the caller protocol on resident pairs: attempt 0 with band w on all pairs,
 * then up to maxBandTry-1 retries with the band doubled, each on the pairs the upstream rule selects (score changed by
 * the attempt and max_off >= 3/4 of the attempt's band), moved to a dense sub-array and back.  One 4-byte read-back per
 * retry (the count) on top of the histogram read-back of every BSW launch. */
static void dev_extend_side(Core &k, DevRt &rt, DevCtx &x, Slot &sl, bsw_seqpair_t *d_pairs, const uint8_t *d_ref, const uint8_t *d_qer,
                            int n, int w, int max_try, cudaStream_t s)
{
    if (n <= 0) return;
    const int T = 256, G = (n + T - 1) / T;
    sl.px_prev.reserve((size_t)n * sizeof(int));
    sl.px_list.reserve((size_t)n * sizeof(int));
    sl.px_count.reserve(sizeof(int));
    bswk::proto_save_score_kernel<<<G, T, 0, s>>>(d_pairs, n, (int *)sl.px_prev.p);
    launch_bsw(k, rt, x, sl, d_pairs, d_ref, d_qer, n, w, s, 0);
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
        launch_bsw(k, rt, x, sl, (bsw_seqpair_t *)sl.px_sub.p, d_ref, d_qer, m, w_now, s, 0);
        bswk::proto_scatter_kernel<<<Gm, T, 0, s>>>((const bsw_seqpair_t *)sl.px_sub.p, (const int *)sl.px_list.p, m, d_pairs);
        CK(cudaGetLastError());
        k.st.kernel_launches += 2;
    }
}

void bsw_dev_extendChained(bsw_t *h, bsw_seqpair_t *d_left, const uint8_t *d_leftRef, const uint8_t *d_leftQer, int32_t numLeft,
                           bsw_seqpair_t *d_right, const uint8_t *d_rightRef, const uint8_t *d_rightQer, int32_t numRight,
                           int32_t w, int32_t maxBandTry, void *stream)
{
    DevCtx &x = h->k.cx(0);
    DevRt &rt = h->k.rt(0);
    CK(cudaSetDevice(x.dev));
    cudaStream_t s = (cudaStream_t)stream;
    h->k.st.calls++;
    Slot &sl = dev_slot(x, s, true);
    const int tries = maxBandTry < 1 ? 1 : maxBandTry;
    dev_extend_side(h->k, rt, x, sl, d_left, d_leftRef, d_leftQer, numLeft, w, tries, s);
    if (numRight > 0) {
        bswk::proto_chain_h0_kernel<<<(numRight + 255) / 256, 256, 0, s>>>(d_left, numLeft, d_right, numRight);
        CK(cudaGetLastError());
        h->k.st.kernel_launches++;
        dev_extend_side(h->k, rt, x, sl, d_right, d_rightRef, d_rightQer, numRight, w, tries, s);
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
    DevCtx &x = h->k.cx(0);
    CK(cudaSetDevice(x.dev));
    h->k.st.calls++;
    Slot &sl = dev_slot(x, (cudaStream_t)stream, true);
    launch_kswv(h->k, h->k.rt(0), x, sl, const_cast<bsw_seqpair_t *>(d_pairArray), d_seqBufRef, d_seqBufQer, d_aln, numPairs, is8, phase,
                (cudaStream_t)stream);
    CK(cudaEventRecord(sl.dev_done, (cudaStream_t)stream));
}

/* ksw_align2 for a resident batch in one submission: forward pass, hand-off on the device (which hits
 * get a second pass, their reversed prefixes, the phase-1 records), reverse pass refining d_aln in place.  The pairs and
 * the sequence buffers are not modified.  Two small read-backs (the hand-off totals, the second pass's histogram) on
 * top of the first pass's histogram. */
/* ksw_align2 on resident buffers: forward pass, hand-off, reverse pass.  sl = scratch slot.  Returns
 * false when the forward pass found a pair outside the given byte ranges (nothing was computed). */
static bool dev_align2_core(Core &k, DevRt &rt, DevCtx &x, Slot &sl, bsw_seqpair_t *d_pairs, const uint8_t *d_seqBufRef, const uint8_t *d_seqBufQer,
                            bsw_kswr_t *d_aln, int n, int is8, cudaStream_t s,
                            long long rlo = 0, long long rhi = (long long)1 << 62, long long qlo = 0, long long qhi = (long long)1 << 62)
{
    if (!launch_kswv(k, rt, x, sl, d_pairs, d_seqBufRef, d_seqBufQer, d_aln, n, is8, 0, s, rlo, rhi, qlo, qhi)) return false;
    /* hand-off: three exclusive scans (slot, target bytes, query bytes), totals, fill */
    sl.kx_off.reserve((size_t)n * 3 * sizeof(int) + 4 * sizeof(int));
    int *off_k = (int *)sl.kx_off.p, *off_r = off_k + n, *off_q = off_r + n, *totals = off_q + n;
    for (int which = 0; which < 3; ++which) {
        auto it = thrust::make_transform_iterator(thrust::make_counting_iterator(0), HandoffLens{ d_pairs, d_aln, which });
        size_t tmp_bytes = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, it, off_k + (size_t)which * n, n, s));
        sl.cubtmp.reserve(tmp_bytes);
        CK(cub::DeviceScan::ExclusiveSum(sl.cubtmp.p, tmp_bytes, it, off_k + (size_t)which * n, n, s));
    }
    /* the fill needs the scratch sizes: last offsets + last lengths */
    kswv_handoff_totals_kernel<<<1, 32, 0, s>>>(d_pairs, d_aln, n, off_k, off_r, off_q, totals);
    CK(cudaMemcpyAsync(sl.h_count, totals, 3 * sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(s));
    const int n1 = sl.h_count[0], tot_r = sl.h_count[1], tot_q = sl.h_count[2];
    k.st.kernel_launches += 7;
    if (n1 < 0 || tot_r < 0 || tot_q < 0) die("kswv align2: hand-off sizes overflow 31 bits (split the batch)");
    if (n1 > 0) {
        sl.kx_pairs.reserve((size_t)n1 * sizeof(bsw_seqpair_t));
        sl.kx_ref.reserve((size_t)tot_r + 64); sl.kx_qer.reserve((size_t)tot_q + 64);
        kswv_handoff_fill_kernel<<<(int)(((long long)n * 8 + 255) / 256), 256, 0, s>>>(d_pairs, d_aln, n, d_seqBufRef, d_seqBufQer, off_k, off_r, off_q,
                                                                                  (bsw_seqpair_t *)sl.kx_pairs.p, (uint8_t *)sl.kx_ref.p,
                                                                                  (uint8_t *)sl.kx_qer.p, totals);
        CK(cudaGetLastError());
        k.st.kernel_launches++;
        launch_kswv(k, rt, x, sl, (bsw_seqpair_t *)sl.kx_pairs.p, (const uint8_t *)sl.kx_ref.p, (const uint8_t *)sl.kx_qer.p, d_aln, n1, is8, 1, s);
    }
    return true;
}

void kswv_dev_align2(kswv_t *h, const bsw_seqpair_t *d_pairArray, const uint8_t *d_seqBufRef, const uint8_t *d_seqBufQer,
                     bsw_kswr_t *d_aln, int32_t numPairs, int is8, void *stream)
{
    Core &k = h->k;
    DevCtx &x = k.cx(0);
    CK(cudaSetDevice(x.dev));
    cudaStream_t s = (cudaStream_t)stream;
    k.st.calls++;
    if (numPairs <= 0) return;
    Slot &sl = dev_slot(x, s, false);
    dev_align2_core(k, k.rt(0), x, sl, const_cast<bsw_seqpair_t *>(d_pairArray), d_seqBufRef, d_seqBufQer, d_aln, numPairs, is8, s);
    CK(cudaEventRecord(sl.dev_done, s));
}

/* ksw_align2 on HOST buffers in one call: what mem_sam_pe_batch does with getScores(phase 0), a host loop that reverses the
 * prefixes of the qualifying hits, and getScores(phase 1) -- here records and bases go up once, the hand-off and both
 * passes stay on the device, the finished kswr_t come down.  Chunks of BSW_CUDA_KSWV_CHUNK pairs (ramped at the ends):
 * chunk c + 1 uploads on its staging slot's stream while chunk c computes on slot 0's. */
/* kswv_align2 / kswv_align2Ref: rref != nullptr -> the targets are expanded on the device from the resident 2-bit reference
 * (rpos[i] as in bsw_getScoresRef) into the staging slot's scratch, the device copies of the records point into it, and no
 * target base crosses PCIe */
static void align2_host(Core &k, const bsw_seqpair_t *pairArray, const uint8_t *seqBufRef, const uint8_t *seqBufQer, bsw_kswr_t *aln,
                        int32_t numPairs, int is8, const bsw_ref *rref, const int64_t *rpos)
{
    k.st.calls++;
    const int n = numPairs;
    if (n <= 0) return;
    if (!pairArray || (!seqBufRef && !rref) || (rref && !rpos) || !seqBufQer || !aln) die("null buffer passed to kswv_align2 / kswv_align2Ref");
    const int nd = (int)k.devs.size();
    for (int i = 0; i < nd; ++i) { DevCtx &xi = k.cx(i); CK(cudaSetDevice(xi.dev)); for (Slot &sl : xi.slot) slot_retire(xi, sl); }
    /* windows from the resident reference: little to upload, so larger chunks (each costs ~0.7 ms of device under-use) */
    const int chunk = std::max(1024, env_int("BSW_CUDA_KSWV_CHUNK", rref ? 1 << 19 : 1 << 18));
    const std::vector<int> cstart = chunk_starts(n, chunk, nd, k.chunk_ramp);
    const int nchunks = (int)cstart.size() - 1;
    /* A chunk's three phases each end in a small read-back the host waits for (two plans, the hand-off sizes), and the
     * device drains meanwhile.  Two workers per device -- the calling thread and helpers, each on a context of its own --
     * take the chunks in turn, so one chunk's kernels fill the other's gaps; worker w serves device w mod (number of
     * devices of the handle) (BSW_CUDA_ALIGN2_WORKERS = workers per device, default 2). */
    const int per_dev = std::max(1, std::min(4, env_int("BSW_CUDA_ALIGN2_WORKERS", 2)));
    const int nworkers = std::max(1, std::min(nd * per_dev, nchunks >= 4 ? nchunks / 2 : 1));
    auto run = [&](Core &kk, DevCtx &xx, int worker) {
        DevRt &rt = k.rt(worker % nd);
        CK(cudaSetDevice(xx.dev));
        Slot &cs = xx.slot[0];                     /* scratch (plan, sort, hand-off) and the compute stream */
        struct Up { long long rlo, rhi, qlo, qhi; } up[2];
        std::vector<int> mine;
        for (int ci = worker; ci < nchunks; ci += nworkers) mine.push_back(ci);
        auto upload = [&](int m, bool exact) {
            Slot &st = xx.slot[1 + (m & 1)];
            slot_retire(xx, st);                   /* the chunk that used this staging slot two chunks ago is home */
            const int first = cstart[mine[m]], cnt = cstart[mine[m] + 1] - first;
            Up &u = up[m & 1];
            long long l1s = 0;
            if (exact) {
                extents(pairArray, first, cnt, u.rlo, u.rhi, u.qlo, u.qhi);
                if (rref) for (int i = first; i < first + cnt; ++i) l1s += std::max(pairArray[i].len1, 0) + 4;
            } else extents_sampled(pairArray, first, cnt, u.rlo, u.rhi, u.qlo, u.qhi, &l1s);
            st.pairs.reserve((size_t)cnt * sizeof(bsw_seqpair_t));
            st.qer.reserve((size_t)(u.qhi - u.qlo) + 64);
            st.aln.reserve((size_t)cnt * sizeof(bsw_kswr_t));
            CK(cudaMemcpyAsync(st.pairs.p, pairArray + first, (size_t)cnt * sizeof(bsw_seqpair_t), cudaMemcpyHostToDevice, st.stream));
            if (u.qhi > u.qlo) CK(cudaMemcpyAsync(st.qer.p, seqBufQer + u.qlo, (size_t)(u.qhi - u.qlo), cudaMemcpyHostToDevice, st.stream));
            kk.st.h2d_bytes += (int64_t)cnt * sizeof(bsw_seqpair_t) + (u.qhi - u.qlo);
            if (rref) {
                /* targets: dense scratch filled from the resident reference; its size is estimated from the sampled lengths
                 * (x 1.25) unless `exact` -- the expansion counts the pairs that do not fit (Plan::oob_pairs, read back with
                 * the upload) and the chunk is redone with the exact size, like a chunk with a pair outside the sampled ranges */
                const long long cap = exact ? l1s + 64 : (l1s * 32 * 5) / 4 + 65536;
                st.reftmp.reserve((size_t)cap + 64); st.rpos.reserve((size_t)cnt * 8); st.toff.reserve((size_t)cnt * 4 + 4);
                st.plan.reserve(sizeof(Plan));
                CK(cudaMemsetAsync(st.plan.p, 0, sizeof(Plan), st.stream));
                CK(cudaMemcpyAsync(st.rpos.p, rpos + first, (size_t)cnt * 8, cudaMemcpyHostToDevice, st.stream));
                size_t tmp_bytes = 0;
                auto it = thrust::make_transform_iterator(thrust::make_counting_iterator(0), Len1Rounded{ (const bsw_seqpair_t *)st.pairs.p });
                CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, it, (int *)st.toff.p, cnt, st.stream));
                st.cubtmp.reserve(tmp_bytes);
                CK(cub::DeviceScan::ExclusiveSum(st.cubtmp.p, tmp_bytes, it, (int *)st.toff.p, cnt, st.stream));
                const uint8_t *d_pac = rref->on(xx.dev);
                if (!d_pac) die("the reference was not uploaded to a device this handle uses");
                ref_expand_kernel<<<(cnt + 31) / 32, 256, 0, st.stream>>>((bsw_seqpair_t *)st.pairs.p, cnt, (const long long *)st.rpos.p, (int *)st.toff.p,
                                                                         d_pac, rref->l_pac, (uint8_t *)st.reftmp.p, cap, (Plan *)st.plan.p);
                CK(cudaGetLastError());
                CK(cudaMemcpyAsync(st.h_plan, st.plan.p, sizeof(Plan), cudaMemcpyDeviceToHost, st.stream));
                kk.st.h2d_bytes += (int64_t)cnt * 8; kk.st.d2h_bytes += sizeof(Plan); kk.st.kernel_launches += 3;
                u.rlo = 0; u.rhi = cap;
            } else {
                st.ref.reserve((size_t)(u.rhi - u.rlo) + 64);
                if (u.rhi > u.rlo) CK(cudaMemcpyAsync(st.ref.p, seqBufRef + u.rlo, (size_t)(u.rhi - u.rlo), cudaMemcpyHostToDevice, st.stream));
                kk.st.h2d_bytes += u.rhi - u.rlo;
            }
            CK(cudaEventRecord(st.dev_done, st.stream));          /* "uploaded" */
        };
        if (!mine.empty()) upload(0, false);
        for (int m = 0; m < (int)mine.size(); ++m) {
            if (m + 1 < (int)mine.size()) upload(m + 1, false);
            Slot &st = xx.slot[1 + (m & 1)];
            const int first = cstart[mine[m]], cnt = cstart[mine[m] + 1] - first;
            for (int attempt = 0; ; ++attempt) {
                const Up &u = up[m & 1];
                CK(cudaStreamWaitEvent(cs.stream, st.dev_done, 0));
                const uint8_t *d_ref = rref ? (const uint8_t *)st.reftmp.p : (const uint8_t *)st.ref.p - u.rlo;
                const bool ok = dev_align2_core(kk, rt, xx, cs, (bsw_seqpair_t *)st.pairs.p, d_ref, (const uint8_t *)st.qer.p - u.qlo,
                                                (bsw_kswr_t *)st.aln.p, cnt, is8, cs.stream, u.rlo, u.rhi, u.qlo, u.qhi);
                /* (the core waited for its stream, which had waited for the upload: the expansion's counters are home) */
                if (rref && st.h_plan->bad_pairs) die("kswv_align2Ref: a pair's window (rpos, len1) leaves the reference");
                if (ok && !(rref && st.h_plan->oob_pairs)) break;      /* oob here: the estimated target scratch was too small */
                if (attempt) die("a pair names bases outside its sequence buffer");
                CK(cudaStreamSynchronize(cs.stream));
                upload(m, true);                   /* the sampled ranges missed a pair: exact scan */
            }
            CK(cudaMemcpyAsync(aln + first, st.aln.p, (size_t)cnt * sizeof(bsw_kswr_t), cudaMemcpyDeviceToHost, cs.stream));
            CK(cudaEventRecord(st.done, cs.stream));
            st.busy = true;
            kk.st.d2h_bytes += (int64_t)cnt * sizeof(bsw_kswr_t);
        }
        for (Slot &sl : xx.slot) slot_retire(xx, sl);
        CK(cudaStreamSynchronize(cs.stream));
    };
    DevCtx &x0 = k.cx(0);
    if (nworkers == 1) { run(k, x0, 0); return; }
    /* every helper gets its own view of the handle (own counters) and its own context: the handle's context of its device
     * for the first worker of a device, one from the pool for the others */
    std::vector<Core> kb((size_t)nworkers - 1, k);
    std::vector<DevCtx *> xb((size_t)nworkers - 1, nullptr);
    std::vector<char> pooled((size_t)nworkers - 1, 0);
    std::vector<std::thread> helpers;
    for (int w = 1; w < nworkers; ++w) {
        const int di = w % nd;
        if (w < nd) xb[w - 1] = &k.cx(di);
        else { xb[w - 1] = ctx_acquire(k.rt(di)); pooled[w - 1] = 1; }
        kb[w - 1].ctx[0] = xb[w - 1];
        memset(&kb[w - 1].st, 0, sizeof(bsw_stats_t));
    }
    for (int w = 1; w < nworkers; ++w) helpers.emplace_back([&, w] { run(kb[w - 1], *xb[w - 1], w); });
    run(k, x0, 0);
    for (std::thread &t : helpers) t.join();
    for (int w = 1; w < nworkers; ++w) {
        const bsw_stats_t &b = kb[w - 1].st;
        k.st.pairs += b.pairs; k.st.pairs_packed += b.pairs_packed; k.st.pairs_general += b.pairs_general; k.st.pairs_wide += b.pairs_wide;
        k.st.kernel_launches += b.kernel_launches; k.st.h2d_bytes += b.h2d_bytes; k.st.d2h_bytes += b.d2h_bytes; k.st.kernel_ms += b.kernel_ms;
        if (!pooled[w - 1]) continue;
        /* a pooled context's cell counter joins the handle's context on the same device */
        DevCtx &home = k.cx(w % nd);
        CK(cudaSetDevice(home.dev));
        unsigned long long ca = 0, cb = 0;
        CK(cudaMemcpy(&ca, home.cells.p, sizeof ca, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(&cb, xb[w - 1]->cells.p, sizeof cb, cudaMemcpyDeviceToHost));
        ca += cb;
        CK(cudaMemcpy(home.cells.p, &ca, sizeof ca, cudaMemcpyHostToDevice));
        ctx_release(k.rt(w % nd), xb[w - 1]);
    }
    CK(cudaSetDevice(x0.dev));
}

void kswv_align2(kswv_t *h, const bsw_seqpair_t *pairArray, const uint8_t *seqBufRef, const uint8_t *seqBufQer, bsw_kswr_t *aln,
                 int32_t numPairs, int is8)
{
    align2_host(h->k, pairArray, seqBufRef, seqBufQer, aln, numPairs, is8, nullptr, nullptr);
}

void kswv_align2Ref(kswv_t *h, const bsw_seqpair_t *pairArray, const bsw_ref_t *ref, const int64_t *rpos, const uint8_t *seqBufQer,
                    bsw_kswr_t *aln, int32_t numPairs, int is8)
{
    if (numPairs > 0 && !ref) die("null reference passed to kswv_align2Ref");
    align2_host(h->k, pairArray, nullptr, seqBufQer, aln, numPairs, is8, ref, rpos);
}

/* ---- CIGAR generation ------------------------------------------------------ */
__global__ void global2_last_kernel(const bswk::GJob *jobs, int count, const long long *zoff, long long *out)
{
    if (threadIdx.x || blockIdx.x) return;
    const bswk::GJob g = jobs[count - 1];
    out[0] = zoff[count - 1] + bswk::global2_zwords(g.tlen, g.qlen, g.w);
}
__global__ void global2_last_cigar_kernel(const bswk::GJob *jobs, int count, const int *coff, long long *out)
{
    if (threadIdx.x || blockIdx.x) return;
    long long bad = 0;
    out[1] = (long long)coff[count - 1] + (jobs[count - 1].n_cigar > 0 ? jobs[count - 1].n_cigar : 0);
    (void)bad;
}
__global__ void global2_count_bad_kernel(const bswk::GJob *jobs, int count, int *bad)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < count && jobs[t].n_cigar < 0) atomicAdd(bad, 1);
}

/* resident jobs and sequences; ops to d_cigar (dense, job order).  Returns the total number of ops; the ops of the jobs
 * that fit below cigarCap are written, the others only counted. */
static long long dev_global2(Core &k, DevCtx &x, Slot &sl, bswk::GJob *d_jobs, int n, const uint8_t *d_ref, const uint8_t *d_qer,
                             uint32_t *d_cigar, long long cap, int max_w, cudaStream_t s)
{
    const int M = 2 * std::max(0, max_w) + 2;
    const int chunk_max = std::max(1024, env_int("BSW_CUDA_GLOBAL_CHUNK", 1 << 18));
    long long total = 0;
    for (int first = 0; first < n; first += chunk_max) {
        const int cnt = std::min(chunk_max, n - first);
        bswk::GJob *jb = d_jobs + first;
        sl.gx_zoff.reserve((size_t)cnt * sizeof(long long));
        sl.gx_coff.reserve((size_t)cnt * sizeof(int) + 16);
        sl.gx_eh.reserve((size_t)cnt * (size_t)M * sizeof(int2));
        sl.px_count.reserve(sizeof(int));
        {
            auto it = thrust::make_transform_iterator(thrust::make_counting_iterator(0), GlobalZWords{ jb });
            size_t tmp = 0;
            CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, it, (long long *)sl.gx_zoff.p, cnt, s));
            sl.cubtmp.reserve(tmp);
            CK(cub::DeviceScan::ExclusiveSum(sl.cubtmp.p, tmp, it, (long long *)sl.gx_zoff.p, cnt, s));
        }
        global2_last_kernel<<<1, 32, 0, s>>>(jb, cnt, (const long long *)sl.gx_zoff.p, (long long *)sl.cubtmp.p);
        CK(cudaMemcpyAsync(sl.h_tot, sl.cubtmp.p, sizeof(long long), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        const long long zwords = sl.h_tot[0];
        sl.gx_z.reserve((size_t)(zwords + 4) * sizeof(uint32_t));
        global2_dp_kernel<<<(cnt + 63) / 64, 64, 0, s>>>(jb, cnt, d_ref, d_qer, k.c, (int2 *)sl.gx_eh.p, M, (uint32_t *)sl.gx_z.p,
                                                         (const long long *)sl.gx_zoff.p, (unsigned long long *)x.cells.p);
        CK(cudaGetLastError());
        {
            auto it = thrust::make_transform_iterator(thrust::make_counting_iterator(0), GlobalNCigar{ jb });
            size_t tmp = 0;
            CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, it, (int *)sl.gx_coff.p, cnt, s));
            sl.cubtmp.reserve(tmp + 64);
            CK(cub::DeviceScan::ExclusiveSum(sl.cubtmp.p, tmp, it, (int *)sl.gx_coff.p, cnt, s));
        }
        CK(cudaMemsetAsync(sl.px_count.p, 0, sizeof(int), s));
        global2_count_bad_kernel<<<(cnt + 255) / 256, 256, 0, s>>>(jb, cnt, (int *)sl.px_count.p);
        global2_emit_kernel<<<(cnt + 63) / 64, 64, 0, s>>>(jb, cnt, (const uint32_t *)sl.gx_z.p, (const long long *)sl.gx_zoff.p,
                                                           (const int *)sl.gx_coff.p, total, d_cigar, cap);
        global2_last_cigar_kernel<<<1, 32, 0, s>>>(jb, cnt, (const int *)sl.gx_coff.p, (long long *)sl.gx_zoff.p);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(sl.h_tot, (const long long *)sl.gx_zoff.p + 1, sizeof(long long), cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(sl.h_count, sl.px_count.p, sizeof(int), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        if (*sl.h_count) die("bsw_global2: a job has a negative length or band, or a band above maxBand");
        total += sl.h_tot[0];
        if (total > 0x7fffffffll) die("bsw_global2: more than 2^31 CIGAR ops in one call (split the batch)");
        k.st.kernel_launches += 9;
        k.st.pairs += cnt;
    }
    return total;
}

int64_t bsw_dev_global2(bsw_t *h, bsw_global_job_t *d_jobs, int32_t numJobs, const uint8_t *d_seqBufRef, const uint8_t *d_seqBufQer,
                        uint32_t *d_cigar, int64_t cigarCap, int32_t maxBand, void *stream)
{
    Core &k = h->k;
    DevCtx &x = k.cx(0);
    CK(cudaSetDevice(x.dev));
    k.st.calls++;
    if (numJobs <= 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    Slot &sl = dev_slot(x, s, false);
    const long long tot = dev_global2(k, x, sl, (bswk::GJob *)d_jobs, numJobs, d_seqBufRef, d_seqBufQer, d_cigar, cigarCap, maxBand, s);
    CK(cudaEventRecord(sl.dev_done, s));
    return tot;
}

int64_t bsw_global2(bsw_t *h, bsw_global_job_t *jobs, int32_t numJobs, const uint8_t *seqBufRef, const uint8_t *seqBufQer,
                    uint32_t *cigar, int64_t cigarCap)
{
    Core &k = h->k;
    k.st.calls++;
    if (numJobs <= 0) return 0;
    if (!jobs || !seqBufRef || !seqBufQer || (!cigar && cigarCap > 0)) die("null buffer passed to bsw_global2");
    DevCtx &x = k.cx(0);
    CK(cudaSetDevice(x.dev));
    Slot &sl = x.slot[0];
    slot_retire(x, sl);
    cudaStream_t s = sl.stream;
    long long rlo = (long long)1 << 62, rhi = 0, qlo = (long long)1 << 62, qhi = 0;
    int max_w = 0, bad = 0;
    for (int i = 0; i < numJobs; ++i) {
        const bsw_global_job_t &g = jobs[i];
        if (g.tlen < 0 || g.qlen < 0 || g.w < 0) { ++bad; continue; }
        rlo = std::min<long long>(rlo, g.idr); rhi = std::max<long long>(rhi, (long long)g.idr + g.tlen);
        qlo = std::min<long long>(qlo, g.idq); qhi = std::max<long long>(qhi, (long long)g.idq + g.qlen);
        max_w = std::max(max_w, g.w);
    }
    if (bad) die("bsw_global2: a job has a negative length or band");
    if (rhi <= rlo) { rlo = 0; rhi = 0; }
    if (qhi <= qlo) { qlo = 0; qhi = 0; }
    sl.gx_jobs.reserve((size_t)numJobs * sizeof(bsw_global_job_t));
    sl.ref.reserve((size_t)(rhi - rlo) + 64); sl.qer.reserve((size_t)(qhi - qlo) + 64);
    sl.gx_cigar.reserve((size_t)std::max<long long>(cigarCap, 1) * sizeof(uint32_t));
    CK(cudaMemcpyAsync(sl.gx_jobs.p, jobs, (size_t)numJobs * sizeof(bsw_global_job_t), cudaMemcpyHostToDevice, s));
    if (rhi > rlo) CK(cudaMemcpyAsync(sl.ref.p, seqBufRef + rlo, (size_t)(rhi - rlo), cudaMemcpyHostToDevice, s));
    if (qhi > qlo) CK(cudaMemcpyAsync(sl.qer.p, seqBufQer + qlo, (size_t)(qhi - qlo), cudaMemcpyHostToDevice, s));
    k.st.h2d_bytes += (long long)numJobs * sizeof(bsw_global_job_t) + (rhi - rlo) + (qhi - qlo);
    const long long tot = dev_global2(k, x, sl, (bswk::GJob *)sl.gx_jobs.p, numJobs, (const uint8_t *)sl.ref.p - rlo, (const uint8_t *)sl.qer.p - qlo,
                                      (uint32_t *)sl.gx_cigar.p, cigarCap, max_w, s);
    CK(cudaMemcpyAsync(jobs, sl.gx_jobs.p, (size_t)numJobs * sizeof(bsw_global_job_t), cudaMemcpyDeviceToHost, s));
    const long long fit = std::min<long long>(tot, cigarCap);
    if (fit > 0) CK(cudaMemcpyAsync(cigar, sl.gx_cigar.p, (size_t)fit * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    k.st.d2h_bytes += (long long)numJobs * sizeof(bsw_global_job_t) + fit * 4;
    return tot <= cigarCap ? tot : -tot;
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
    for (DevCtx *x : k.ctx) {
        if (!x) continue;
        CK(cudaSetDevice(x->dev));
        CK(cudaDeviceSynchronize());
        unsigned long long c = 0;
        CK(cudaMemcpy(&c, x->cells.p, sizeof c, cudaMemcpyDeviceToHost));
        cells += (long long)c;
    }
    k.st.cells = cells;
    *out = k.st;
}
static void stats_reset(Core &k)
{
    memset(&k.st, 0, sizeof k.st);
    for (DevCtx *x : k.ctx) {
        if (!x) continue;
        CK(cudaSetDevice(x->dev));
        CK(cudaDeviceSynchronize());
        CK(cudaMemset(x->cells.p, 0, sizeof(unsigned long long)));
    }
}

void bsw_get_stats(bsw_t *h, bsw_stats_t *out) { stats_with_pending(h->k, out); }
void bsw_reset_stats(bsw_t *h) { stats_reset(h->k); }
void bsw_set_timing(bsw_t *h, int on) { h->k.timing = on; }
void bsw_set_cell_counting(bsw_t *h, int on) { h->k.count_cells = on; }
void bsw_set_direct_min(bsw_t *h, int pairs) { h->k.direct_min = pairs; }
void bsw_set_wide_max(bsw_t *h, int pairs) { h->k.wide_max = pairs; }
void bsw_set_wide_policy(bsw_t *h, int wide_max, int wide_lat, int wide_split) { h->k.wide_max = wide_max; h->k.wide_lat = wide_lat; h->k.wide_split = wide_split; }
void kswv_get_stats(kswv_t *h, bsw_stats_t *out) { stats_with_pending(h->k, out); }
void kswv_reset_stats(kswv_t *h) { stats_reset(h->k); }
void kswv_set_timing(kswv_t *h, int on) { h->k.timing = on; }
void kswv_set_direct_min(kswv_t *h, int pairs) { h->k.direct_min = pairs; }

void bsw_engine_get_stats(int device, bsw_engine_stats_t *out)
{
    memset(out, 0, sizeof *out);
    DevRt *r = nullptr;
    { std::lock_guard<std::mutex> lk(g_rt_mu); if (device >= 0 && device < (int)g_rt.size()) r = g_rt[device]; }
    if (!r || !r->engine) return;
    const EngineStats s = r->engine->stats();
    out->batches = s.batches; out->pairs = s.pairs; out->calls = s.calls; out->kernel_launches = s.launches;
    out->h2d_bytes = s.h2d_bytes; out->d2h_bytes = s.d2h_bytes; out->max_batch_pairs = s.max_batch_pairs; out->wide_batches = s.wide_batches;
    out->scan_ns = s.scan_ns; out->reserve_ns = s.reserve_ns; out->pack_ns = s.pack_ns; out->wait_ns = s.wait_ns; out->unpack_ns = s.unpack_ns;
    out->packers_ns = s.packers_ns; out->submit_ns = s.submit_ns; out->flight_ns = s.flight_ns;
}
void bsw_engine_reset_stats(int device)
{
    DevRt *r = nullptr;
    { std::lock_guard<std::mutex> lk(g_rt_mu); if (device >= 0 && device < (int)g_rt.size()) r = g_rt[device]; }
    if (r && r->engine) r->engine->reset_stats();
}

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

/* ------------------------------------------------------------------ bsw_dev_getScores_static
 * The allocation-free variant of bsw_dev_getScores (include/bsw_cuda.h; no reference counterpart).
 * The general kernel's DP rows live in the file-scope array below instead of Slot::scratch.
 * bsw_static_scratch is a HOST-side stand-in in host code, not a device address: the pointer the kernel gets comes from
 * cudaGetSymbolAddress, once per device, with that device current. */
static __device__ __align__(256) unsigned char bsw_static_scratch[BSW_STATIC_SCRATCH_BYTES];
/* layout: [0, 4) status word = pairs refused by the kernel (cleared by every call); [256, ...) int2 rows, entry j of
 * pair g at (j * numPairs + g) -- the layout of bsw_general_kernel */

namespace {

__global__ void __launch_bounds__(64)
bsw_static_kernel(bsw_seqpair_t *__restrict__ pairs, const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                  int n, int max_qlen, Const c, int w, unsigned char *scratch)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    const PairIO io = { (const char *)pairs, (char *)pairs + offsetof(bsw_seqpair_t, score), (int)sizeof(bsw_seqpair_t), (int)sizeof(bsw_seqpair_t) };
    const PairIn P = load_pair(io, g);
    PairOut o;
    if (P.len1 < 0 || P.len2 < 0 || P.len2 > max_qlen) {
        o.score = o.tle = o.gtle = o.qle = o.gscore = o.max_off = -1;
        atomicAdd(reinterpret_cast<int *>(bsw_static_scratch), 1);      /* device code names the array directly */
    } else {
        bsw_general_thread<long long>(c, w, P, ref, qer, reinterpret_cast<int2 *>(scratch + 256) + g, (long long)n, o);
    }
    store_bsw(io, g, o);
}

struct StaticScratch { std::once_flag once; unsigned char *ptr = nullptr; cudaError_t err = cudaSuccess; };
StaticScratch g_static_scratch[64];

int static_fail(int code, const char *what, cudaError_t e = cudaSuccess)
{
    if (e != cudaSuccess) fprintf(stderr, "[bsw-b200] bsw_dev_getScores_static: %s: %s (%s)\n", what, cudaGetErrorName(e), cudaGetErrorString(e));
    else fprintf(stderr, "[bsw-b200] bsw_dev_getScores_static: %s\n", what);
    return code;
}

} /* namespace */

extern "C" int bsw_dev_getScores_static(bsw_t *h, bsw_seqpair_t *d_pairArray, const uint8_t *d_seqBufRef, const uint8_t *d_seqBufQer,
                                        int32_t numPairs, int32_t maxQueryLen, int32_t w, void *stream)
{
    if (!h || numPairs < 0 || maxQueryLen < 0) return static_fail(-1, "invalid argument (null handle, negative numPairs or maxQueryLen)");
    if (numPairs == 0) return 0;
    if (!d_pairArray || !d_seqBufRef || !d_seqBufQer) return static_fail(-1, "invalid argument (null device pointer)");
    const long long entries = (long long)numPairs * ((long long)maxQueryLen + 1);
    if (entries > (long long)BSW_STATIC_MAX_ENTRIES) {
        fprintf(stderr, "[bsw-b200] bsw_dev_getScores_static: numPairs * (maxQueryLen + 1) = %d * %d = %lld exceeds the static scratch of "
                        "%d entries (BSW_STATIC_MAX_ENTRIES); split the batch or use bsw_dev_getScores\n",
                numPairs, maxQueryLen + 1, entries, (int)BSW_STATIC_MAX_ENTRIES);
        return -2;
    }
    const int dev = h->k.devs[0];
    if (dev < 0 || dev >= 64) return static_fail(-1, "device index beyond the per-device table (64)");
    cudaError_t e = cudaSetDevice(dev);
    if (e != cudaSuccess) return static_fail(-3, "cudaSetDevice", e);
    StaticScratch &ss = g_static_scratch[dev];
    std::call_once(ss.once, [&ss]() {           /* set-up: once per device, that device current */
        void *p = nullptr;
        ss.err = cudaGetSymbolAddress(&p, bsw_static_scratch);
        ss.ptr = (unsigned char *)p;
    });
    if (ss.err != cudaSuccess || !ss.ptr) return static_fail(-3, "cudaGetSymbolAddress(bsw_static_scratch)", ss.err);
    cudaStream_t s = (cudaStream_t)stream;
    e = cudaMemsetAsync(ss.ptr, 0, 256, s);      /* the status word must start at zero; the rows are rewritten by the kernel */
    if (e != cudaSuccess) return static_fail(-3, "cudaMemsetAsync", e);
    bsw_static_kernel<<<(numPairs + 63) / 64, 64, 0, s>>>(d_pairArray, d_seqBufRef, d_seqBufQer, numPairs, maxQueryLen, h->k.c, w, ss.ptr);
    e = cudaGetLastError();
    if (e != cudaSuccess) return static_fail(-3, "kernel launch", e);
    /* a capturing stream cannot be waited for: the launch is recorded into the caller's graph and the status word stays
     * on the device (pairs the kernel refuses carry -1 in their result words) */
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cap) == cudaSuccess && cap != cudaStreamCaptureStatusNone) return 0;
    int refused = 0;
    e = cudaMemcpyAsync(&refused, ss.ptr, sizeof(int), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return static_fail(-3, "status read-back", e);
    if (refused) {
        fprintf(stderr, "[bsw-b200] bsw_dev_getScores_static: %d pair(s) with len2 > maxQueryLen = %d or a negative length (result words set to -1)\n",
                refused, maxQueryLen);
        return -4;
    }
    return 0;
}
