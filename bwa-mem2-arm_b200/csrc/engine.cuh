/*
 * engine.cuh -- the staged-call engine: how the SMALL boundary calls of an unchanged bwa-mem2 reach the GPU.
 *
 * The reference calls BandedPairWiseSW::getScores8/16 and kswv::getScores8/16 from every kt_for worker thread
 * with ~1-3 k pairs per call (R:docs-archive/WEEK1_BENCHMARK_RESULTS.md:88-95: 958,581 pairs in 391 batches;
 * R:ARM_OPTIMIZATION_IMPLEMENTATION.md:46-59), each worker on its own object and its own pageable arenas
 * (SURVEY.md 8b "Threading").  A GPU wants >= ~50 k pairs per submission.  This engine closes that gap without
 * changing a signature or a result:
 *
 *   caller thread (any number, no CUDA call at all)          engine thread (one per device, process-wide)
 *   -----------------------------------------------          --------------------------------------------
 *   1 scan its records: contract checks, class of each       a) closes an open slot as soon as the device has
 *     pair (query-length bin / has-N / general), byte sums      room for another batch (<= BSW_CUDA_INFLIGHT in
 *   2 reserve a range of the OPEN slot with the same            flight) or the slot is full; waits for its packers
 *     (mode, scoring, band) key -- one short mutex hold      b) submits the batch on the slot's stream: 2 H2D
 *   3 pack, outside the lock: 24-byte compact records           copies, class scatter, ONE multi-bin DP launch
 *     (offsets rebased into the slot) + exactly the bases       (+ one for N queries, + general kernel if any),
 *     its pairs name, into the slot's pinned staging            1 D2H copy, 1 event
 *   4 sleep until the slot's batch is done                   c) polls the events, wakes the callers
 *   5 copy its 24-byte (BSW) / 28-byte (kswv) results
 *     into its own SeqPair / kswr_t array
 *
 * So concurrent small calls merge into one device batch ("group commit"): the batch is whatever arrived while the
 * previous batches were in flight.  The host never waits for a device-side histogram: the callers compute the
 * classes while they pack, so the bin ranges are known when the batch closes.  Packing is done by the worker threads
 * themselves, in parallel, and it is the only pass over the caller's (pageable) memory: only the 24 input bytes of
 * each 56-byte record and the named bases cross PCIe, and only the result bytes come back.
 *
 * Results are independent of batching: every pair is computed by the same exact kernels whoever shares its batch,
 * warp or thread (tests/test_gpu_parity.py::test_staged_*).
 */
#pragma once
#include <atomic>
#include <mutex>
#include <condition_variable>
#include <thread>
#include <vector>
#include <chrono>
#include <algorithm>
#include <memory>
#include "host_common.cuh"
#include "bsw_kernels.cuh"

namespace bswh {

using namespace bswk;

/* what must be equal for two calls to share a batch */
struct BatchKey {
    int mode;            /* MODE_BSW / MODE_KSWV */
    Const c;
    int w;               /* BSW band */
    int is8;             /* kswv precision rule */
    int wide_max;        /* BSW: batches up to this many pairs run on the lanes-per-pair kernel (handle settings) */
    int wide_lat;        /*      ... up to this many on its shapes with more lanes per pair */
    int wide_split;      /*      ... bins up to this index stay on the thread kernel in between (0: none) */
    bool operator==(const BatchKey &o) const { return memcmp(this, &o, sizeof *this) == 0; }
};
static inline BatchKey make_key(int mode, const Const &c, int w, int is8, int wide_max, int wide_lat, int wide_split)
{
    BatchKey k; memset(&k, 0, sizeof k);
    k.mode = mode; k.c = c; k.w = mode == MODE_BSW ? w : 0; k.is8 = mode == MODE_KSWV ? is8 : 0;
    k.wide_max = mode == MODE_BSW ? wide_max : 0;
    k.wide_lat = mode == MODE_BSW ? wide_lat : 0;
    k.wide_split = mode == MODE_BSW ? wide_split : 0;
    return k;
}

/* classes 1..16: packed, bin = class; 17..32: packed with an N in the query; 33: general 32-bit kernel;
 * 34..38: only the 16-bit lanes-per-pair kernel admits them (bsw_wide.cuh), by configuration */
enum { CLS_GENERAL = 33, CLS_WIDE0 = 34 };
static_assert(CLS_WIDE0 + BSW_WIDE_NCFG <= BSW_NCLASS, "class table too small");

struct EngineStats {
    long long batches, pairs, calls, launches, h2d_bytes, d2h_bytes, max_batch_pairs, wide_batches;
    /* where the time goes, ns summed over calls (caller side) and batches (engine side) */
    long long scan_ns, reserve_ns, pack_ns, wait_ns, unpack_ns;      /* callers: steps 1..5 of the header comment */
    long long packers_ns, submit_ns, flight_ns;                     /* engine: wait for packers, enqueue, submit -> done seen */
};
struct EngineProf {
    std::atomic<long long> scan_ns{0}, reserve_ns{0}, pack_ns{0}, wait_ns{0}, unpack_ns{0}, packers_ns{0}, submit_ns{0}, flight_ns{0};
    void clear() { scan_ns = 0; reserve_ns = 0; pack_ns = 0; wait_ns = 0; unpack_ns = 0; packers_ns = 0; submit_ns = 0; flight_ns = 0; }
};
static inline long long now_ns()
{
    return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct Reservation {
    struct StagedSlot *slot = nullptr;
    int gen = 0;
    int pair_base = 0;
    long long byte_base = 0;
    int cls_pos[BSW_NCLASS];
};

struct StagedSlot {
    enum State { FREE, OPEN, CLOSED, INFLIGHT, DONE };
    State state = FREE;
    int gen = 0;
    BatchKey key;
    /* fill level (under the engine mutex) */
    int npairs = 0, ncallers = 0;
    long long nbytes = 0;
    int cls_cnt[BSW_NCLASS];
    int max_len1_packed = 0, max_len2_general = 0, max_len1_general = 0;
    bool want_close = false;
    std::atomic<int> packed{0};
    /* pinned staging + device mirrors */
    char *h_rec = nullptr, *h_seq = nullptr, *h_res = nullptr;
    char *d_rec = nullptr, *d_seq = nullptr, *d_res = nullptr;
    uint32_t *d_sorted = nullptr;
    DevBuf scratch;
    cudaStream_t stream = nullptr, stream2 = nullptr;
    cudaEvent_t done_ev = nullptr, fork_ev = nullptr, join_ev = nullptr;
    long long t_submit = 0;
    std::mutex done_mu;
    std::condition_variable done_cv;
};

class Engine {
public:
    int dev = 0;
    int cap_pairs = 1 << 16;
    long long cap_bytes = 32ll << 20;
    int nslots = 6, max_inflight = 3;
    int spin_iters = 400;                         /* wait_done: pause iterations before blocking */
    int smem_optin = 0;
    std::atomic<long long> outstanding{0};        /* pairs reserved and not yet handed back: load measure for device choice */

    explicit Engine(int device) : dev(device)
    {
        cap_pairs = std::max(1024, env_int("BSW_CUDA_STAGE_PAIRS", cap_pairs));
        cap_bytes = (long long)std::max(1, env_int("BSW_CUDA_STAGE_MB", 32)) << 20;
        nslots = std::max(2, std::min(16, env_int("BSW_CUDA_STAGE_SLOTS", nslots)));
        max_inflight = std::max(1, std::min(nslots - 1, env_int("BSW_CUDA_INFLIGHT", max_inflight)));
        spin_iters = std::max(0, env_int("BSW_CUDA_SPIN", spin_iters));
        memset(&st, 0, sizeof st);
        slots.reset(new StagedSlot[nslots]);
        done_gens.reset(new std::atomic<int>[nslots]);
        for (int i = 0; i < nslots; ++i) done_gens[i].store(0);
        std::unique_lock<std::mutex> lk(mu);
        th = std::thread([this] { run(); });
        cv_ready.wait(lk, [this] { return ready; });
    }
    ~Engine() { shutdown(); }

    void shutdown()
    {
        { std::lock_guard<std::mutex> lk(mu); if (stop) return; stop = true; }
        cv_engine.notify_all();
        if (th.joinable()) th.join();
    }

    /* step 2 of the header comment.  `cnt` = per-class pair counts of the call (or of the part of it that fits one
     * slot), `bytes` = bases it will pack.  Blocks while no slot can take it. */
    Reservation reserve(const BatchKey &key, int n, long long bytes, const int *cnt, int max_len1_packed,
                        int max_len2_general, int max_len1_general)
    {
        if (n > cap_pairs || bytes > cap_bytes) die("internal: staged reservation larger than a slot");
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            StagedSlot *s = nullptr;
            for (StagedSlot &x : all()) if (x.state == StagedSlot::OPEN && x.key == key) { s = &x; break; }
            if (s && (s->npairs + n > cap_pairs || s->nbytes + bytes + 64 > cap_bytes)) {
                s->want_close = true;                 /* full: the engine submits it; take a fresh one */
                wake_engine();
                s = nullptr;
                /* fall through to a FREE slot; two OPEN slots with one key never coexist because the full one is
                 * closed by the engine before it looks at anything else (want_close is checked first) */
                cv_free.wait(lk, [&] {
                    for (StagedSlot &x : all()) if (x.state == StagedSlot::OPEN && x.key == key && x.want_close) return false;
                    return true; });
                continue;
            }
            if (!s) {
                for (StagedSlot &x : all()) if (x.state == StagedSlot::FREE) { s = &x; break; }
                if (!s) { cv_free.wait(lk); continue; }
                s->state = StagedSlot::OPEN; s->gen++; s->key = key;
                s->npairs = 0; s->nbytes = 0; s->ncallers = 0; s->want_close = false;
                s->max_len1_packed = s->max_len2_general = s->max_len1_general = 0;
                memset(s->cls_cnt, 0, sizeof s->cls_cnt);
                s->packed.store(0, std::memory_order_relaxed);
            }
            Reservation r;
            r.slot = s; r.gen = s->gen; r.pair_base = s->npairs; r.byte_base = s->nbytes;
            for (int c = 0; c < BSW_NCLASS; ++c) { r.cls_pos[c] = s->cls_cnt[c]; s->cls_cnt[c] += cnt[c]; }
            s->npairs += n; s->nbytes += (bytes + 3) & ~3ll; s->ncallers++;
            s->max_len1_packed = std::max(s->max_len1_packed, max_len1_packed);
            s->max_len2_general = std::max(s->max_len2_general, max_len2_general);
            s->max_len1_general = std::max(s->max_len1_general, max_len1_general);
            st.calls++;
            outstanding.fetch_add(n, std::memory_order_relaxed);
            return r;
        }
    }

    /* step 3 done: the caller's records and bases are in the slot */
    void packed(const Reservation &r, int n)
    {
        r.slot->packed.fetch_add(n, std::memory_order_release);
        wake_engine();
    }

    /* step 4 */
    void wait_done(const Reservation &r)
    {
        StagedSlot *s = r.slot;
        for (int spin = 0; spin < spin_iters; ++spin) {    /* short spin: a wake-up through the futex costs more than a small batch */
            if (done_gen(s) >= r.gen) return;
            cpu_relax();
            if ((spin & 63) == 63) std::this_thread::yield();
        }
        std::unique_lock<std::mutex> lk(s->done_mu);       /* per slot: a landing batch wakes its own callers only */
        s->done_cv.wait(lk, [&] { return done_gen(s) >= r.gen; });
    }

    /* step 5 done */
    void release(const Reservation &r, int n)
    {
        outstanding.fetch_sub(n, std::memory_order_relaxed);
        std::lock_guard<std::mutex> lk(mu);
        if (--r.slot->ncallers == 0) { r.slot->state = StagedSlot::FREE; cv_free.notify_all(); }
    }

    EngineProf prof;
    EngineStats stats()
    {
        std::lock_guard<std::mutex> lk(mu);
        EngineStats r = st;
        r.scan_ns = prof.scan_ns; r.reserve_ns = prof.reserve_ns; r.pack_ns = prof.pack_ns; r.wait_ns = prof.wait_ns;
        r.unpack_ns = prof.unpack_ns; r.packers_ns = prof.packers_ns; r.submit_ns = prof.submit_ns; r.flight_ns = prof.flight_ns;
        return r;
    }
    void reset_stats() { std::lock_guard<std::mutex> lk(mu); memset(&st, 0, sizeof st); prof.clear(); }

private:
    std::unique_ptr<StagedSlot[]> slots;
    std::unique_ptr<std::atomic<int>[]> done_gens;
    struct SlotRange { StagedSlot *b, *e; StagedSlot *begin() const { return b; } StagedSlot *end() const { return e; } };
    SlotRange all() { return SlotRange{ slots.get(), slots.get() + nslots }; }
    std::mutex mu;
    std::condition_variable cv_engine, cv_free, cv_ready;
    std::thread th;
    bool stop = false, ready = false;
    std::atomic<int> engine_sleeping{0};
    EngineStats st;
    DevBuf cells;

    static inline void cpu_relax()
    {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#else
        std::this_thread::yield();
#endif
    }
    int done_gen(StagedSlot *s) { return done_gens[(size_t)(s - slots.get())].load(std::memory_order_acquire); }

    void wake_engine()
    {
        if (engine_sleeping.load(std::memory_order_acquire)) cv_engine.notify_one();
    }

    void init_device()
    {
        CK(cudaSetDevice(dev));
        CK(cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        for (StagedSlot &s : all()) {
            const size_t rec = (size_t)cap_pairs * 24, res = (size_t)cap_pairs * sizeof(bsw_kswr_t), seq = (size_t)cap_bytes + 256;
            CK(cudaHostAlloc((void **)&s.h_rec, rec, cudaHostAllocPortable));
            CK(cudaHostAlloc((void **)&s.h_seq, seq, cudaHostAllocPortable));
            CK(cudaHostAlloc((void **)&s.h_res, res, cudaHostAllocPortable));
            memset(s.h_seq, 0, seq);
            CK(cudaMalloc((void **)&s.d_rec, rec)); CK(cudaMalloc((void **)&s.d_seq, seq)); CK(cudaMalloc((void **)&s.d_res, res));
            CK(cudaMemset(s.d_seq, 0, seq));
            CK(cudaMalloc((void **)&s.d_sorted, (size_t)cap_pairs * 4));
            CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
            CK(cudaStreamCreateWithFlags(&s.stream2, cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&s.done_ev, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&s.fork_ev, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&s.join_ev, cudaEventDisableTiming));
        }
        cells.reserve(sizeof(unsigned long long));
        CK(cudaMemset(cells.p, 0, sizeof(unsigned long long)));
    }

    void submit(StagedSlot &s);

    void run()
    {
        init_device();
        { std::lock_guard<std::mutex> lk(mu); ready = true; }
        cv_ready.notify_all();
        std::vector<StagedSlot *> todo, flying, landed;
        auto last_activity = std::chrono::steady_clock::now();
        std::unique_lock<std::mutex> lk(mu);
        while (!stop) {
            bool progress = false;
            /* 1. completions: the event queries run without the engine mutex (callers reserve and release under it) */
            if (!flying.empty()) {
                lk.unlock();
                landed.clear();
                for (size_t k = 0; k < flying.size();) {
                    if (cudaEventQuery(flying[k]->done_ev) == cudaSuccess) { landed.push_back(flying[k]); flying[k] = flying.back(); flying.pop_back(); }
                    else ++k;
                }
                const long long t = landed.empty() ? 0 : now_ns();
                lk.lock();
                for (StagedSlot *s : landed) { s->state = StagedSlot::DONE; prof.flight_ns += t - s->t_submit; }
                if (!landed.empty()) {
                    lk.unlock();
                    for (StagedSlot *s : landed) {
                        done_gens[(size_t)(s - slots.get())].store(s->gen, std::memory_order_release);
                        { std::lock_guard<std::mutex> g(s->done_mu); }
                        s->done_cv.notify_all();
                    }
                    lk.lock();
                    progress = true;
                }
            }
            /* 2. close what the device has room for */
            todo.clear();
            for (StagedSlot &s : all())
                if (s.state == StagedSlot::OPEN && s.npairs > 0 && (s.want_close || (int)(flying.size() + todo.size()) < max_inflight)) {
                    s.state = StagedSlot::CLOSED;
                    todo.push_back(&s);
                }
            if (!todo.empty()) {
                cv_free.notify_all();                      /* callers waiting for a full slot to leave the OPEN set */
                lk.unlock();
                for (StagedSlot *s : todo) {
                    const long long t0 = now_ns();
                    while (s->packed.load(std::memory_order_acquire) < s->npairs) cpu_relax();   /* packers are quick and never block */
                    const long long t1 = now_ns();
                    submit(*s);
                    s->t_submit = now_ns();
                    prof.packers_ns += t1 - t0; prof.submit_ns += s->t_submit - t1;
                }
                lk.lock();
                for (StagedSlot *s : todo) {
                    s->state = StagedSlot::INFLIGHT;
                    flying.push_back(s);
                    st.batches++; st.pairs += s->npairs; st.max_batch_pairs = std::max<long long>(st.max_batch_pairs, s->npairs);
                }
                progress = true;
            }
            if (progress) { last_activity = std::chrono::steady_clock::now(); continue; }
            bool pending = !flying.empty();
            for (StagedSlot &s : all()) if (s.state == StagedSlot::OPEN && s.npairs > 0) pending = true;
            if (pending || std::chrono::steady_clock::now() - last_activity < std::chrono::microseconds(300)) {
                lk.unlock();                               /* busy period: poll (events + new arrivals) */
                for (int k = 0; k < 32; ++k) cpu_relax();
                /* give the core to a runnable caller if there is one (returns at once otherwise): with more worker threads than
                 * cores -- or one engine per GPU on an 8-GPU host -- a polling engine would otherwise starve the packers */
                std::this_thread::yield();
                lk.lock();
            } else {
                engine_sleeping.store(1, std::memory_order_release);
                cv_engine.wait_for(lk, std::chrono::milliseconds(50));
                engine_sleeping.store(0, std::memory_order_release);
                last_activity = std::chrono::steady_clock::now();
            }
        }
    }
};

/* one closed batch -> the device.  Engine thread only. */
inline void Engine::submit(StagedSlot &s)
{
    const int n = s.npairs;
    const BatchKey &key = s.key;
    cudaStream_t q = s.stream;
    CK(cudaMemcpyAsync(s.d_rec, s.h_rec, (size_t)n * 24, cudaMemcpyHostToDevice, q));
    if (s.nbytes > 0) CK(cudaMemcpyAsync(s.d_seq, s.h_seq, (size_t)s.nbytes + 16, cudaMemcpyHostToDevice, q));
    long long launches = 0;
    /* class order = sorted order.  Which kernel serves the packed-eligible bins of a BSW batch depends on its size:
     *   n <= wide_lat   every bin on the lanes-per-pair kernel, shapes with more lanes (a pair takes 50-100 us);
     *   n <= wide_max   the bins above wide_split x 16 columns on the lanes-per-pair kernel (throughput shapes), the
     *                   shorter ones on the thread-owns-two-pairs kernel, side by side on two streams.  Measured
     *                   (profiles/r02_stream_matrix.jsonl): a thread needs ~0.4 ms even for <= 76-column pairs when the
     *                   device is this empty, so the default split is 0 = every bin on the lanes-per-pair kernel;
     *   larger          every bin on the thread kernel (throughput; the batch is big enough to hide its latency).
     * Thread-kernel classes: queries with an N first, then the plain ones, each longest bin first (one multi-bin launch
     * per kind).  Lanes-per-pair classes: bins longest first with the N and plain pairs of a bin adjacent, the bins of
     * one configuration forming one segment of ONE launch.  Then the wide-only classes, then the general pairs. */
    ClassStart cs; memset(&cs, 0, sizeof cs);
    BinTable tb[2]; memset(tb, 0, sizeof tb);          /* [0]: N classes (BSW only), [1]: plain classes */
    int split = BSW_NBINS;
    if (key.mode == MODE_BSW && n <= key.wide_max && wide_eligible(key.c, 0, 0, 1)) split = n <= key.wide_lat ? 0 : std::max(0, std::min(key.wide_split, BSW_NBINS));
    const bool use_wide = split < BSW_NBINS;
    int pos = 0, maxne[2] = { 0, 0 };
    std::vector<WideSeg> segs;
    for (int pass = 0; pass < 2; ++pass) {
        BinTable &t = tb[pass];
        int blocks = 0;
        for (int qb = split; qb >= 1; --qb) {
            const int cls = pass == 0 ? BSW_NBINS + qb : qb, cnt = s.cls_cnt[cls];
            cs.start[cls] = pos;
            if (cnt <= 0) continue;
            t.first_block[t.nbins] = blocks; t.first_pair[t.nbins] = pos; t.ne[t.nbins] = qb * 16;
            maxne[pass] = std::max(maxne[pass], qb * 16);
            t.nbins++;
            blocks += (cnt + 63) / 64; pos += cnt;
        }
        t.first_block[t.nbins] = blocks; t.first_pair[t.nbins] = pos;
    }
    const int n_thread = pos;
    for (int qb = BSW_NBINS; qb > split; --qb)
        for (int pass = 0; pass < 2; ++pass) {
            const int cls = pass == 0 ? BSW_NBINS + qb : qb, cnt = s.cls_cnt[cls], cf = wide_cfg_of_bin(qb);
            cs.start[cls] = pos;
            if (cnt <= 0) continue;
            if (!segs.empty() && segs.back().cfg == cf) segs.back().count += cnt; else segs.push_back(WideSeg{ cf, pos, cnt });
            pos += cnt;
        }
    const int n_packed = n_thread;   /* kswv: the packed classes */
    for (int cf = BSW_WIDE_NCFG - 1; cf >= 0; --cf) {      /* wide-only classes */
        const int cnt = s.cls_cnt[CLS_WIDE0 + cf];
        cs.start[CLS_WIDE0 + cf] = pos;
        if (cnt > 0) segs.push_back(WideSeg{ cf, pos, cnt });
        pos += cnt;
    }
    cs.start[CLS_GENERAL] = pos;
    const int n_general = s.cls_cnt[CLS_GENERAL], gen_first = pos;
    class_scatter_kernel<<<(n + 255) / 256, 256, 0, q>>>(s.d_rec, n, cs, s.d_sorted);
    ++launches;
    const PairIO io = pairio_compact(s.d_rec, s.d_res);
    const uint8_t *seq = (const uint8_t *)s.d_seq;
    unsigned long long *d_cells = (unsigned long long *)cells.p;
    if (key.mode == MODE_BSW) {
        const bool sym = key.c.oe_del == key.c.oe_ins;
        /* the lanes-per-pair segments run beside the thread kernel on the slot's second stream */
        const bool fork = !segs.empty() && n_thread > 0;
        if (fork) { CK(cudaEventRecord(s.fork_ev, q)); CK(cudaStreamWaitEvent(s.stream2, s.fork_ev, 0)); }
        for (int pass = 0; pass < 2; ++pass) {
            const BinTable &t = tb[pass];
            const int blocks = t.first_block[t.nbins];
            if (blocks <= 0) continue;
            auto kern = pass == 0 ? (sym ? bsw_packed_kernel<false, true, true> : bsw_packed_kernel<false, false, true>)
                                  : (sym ? bsw_packed_kernel<false, true, false> : bsw_packed_kernel<false, false, false>);
            kern<<<blocks, 32, maxne[pass] * BSW_WARP_ELEM_BYTES, q>>>(io, seq, seq, s.d_sorted, t, key.c, key.w, d_cells);
            ++launches;
        }
        if (!segs.empty()) {
            launches += launch_wide_segs(segs.data(), (int)segs.size(), n <= key.wide_lat, sym, io, seq, seq, s.d_sorted, key.c, key.w,
                                         fork ? s.stream2 : q);
            if (fork) { CK(cudaEventRecord(s.join_ev, s.stream2)); CK(cudaStreamWaitEvent(q, s.join_ev, 0)); }
        }
        if (n_general > 0) {
            s.scratch.reserve((size_t)n_general * ((size_t)s.max_len2_general + 2) * sizeof(int2));
            bsw_general_kernel<<<(n_general + 63) / 64, 64, 0, q>>>(io, seq, seq, s.d_sorted, gen_first, n_general, key.c, key.w,
                                                                    (int2 *)s.scratch.p, d_cells);
            ++launches;
        }
    } else {
        bsw_kswr_t *d_aln = (bsw_kswr_t *)s.d_res;
        if (n_packed > 0) {
            /* kswv_sys.cuh.  A small batch (more lanes per group, symmetric gap sums) is ONE launch over all its packed
             * pairs -- every query length, with or without N (then the general score table for everybody): two or three
             * bin launches one after another would add their latencies.  Otherwise one launch per non-empty bin: queries
             * with an N first, longest bin first. */
            const bool sym = key.c.oe_del == key.c.oe_ins;
            const int lanes = sym ? kswv_sys_lanes_for(n_packed) : 8;
            if (lanes > 8) {
                launch_kswv_sys(std::max(maxne[0], maxne[1]), sym, tb[0].nbins > 0, lanes, 0, n_packed, std::max(1, s.max_len1_packed), io, seq, seq,
                                d_aln, s.d_sorted, key.c, key.is8, 0, d_cells, q);
                ++launches;
            } else
                for (int pass = 0; pass < 2; ++pass) {
                    const BinTable &t = tb[pass];
                    for (int b = 0; b < t.nbins; ++b) {
                        launch_kswv_sys(t.ne[b], sym, pass == 0, 8, t.first_pair[b], t.first_pair[b + 1] - t.first_pair[b],
                                        std::max(1, s.max_len1_packed), io, seq, seq, d_aln, s.d_sorted, key.c, key.is8, 0, d_cells, q);
                        ++launches;
                    }
                }
        }
        for (const WideSeg &sg : segs) {               /* the 16-bit classes (kswv_sys.cuh, W), largest configuration first */
            launch_kswv_sysw(sg.cfg, key.c.oe_del == key.c.oe_ins, sg.first, sg.count, std::max(1, s.max_len1_packed), io, seq, seq, d_aln,
                             s.d_sorted, key.c, key.is8, 0, d_cells, q);
            ++launches;
        }
        if (n_general > 0) {
            const size_t he_b = (size_t)n_general * ((size_t)s.max_len2_general + 2) * sizeof(int2);
            s.scratch.reserve(he_b + (size_t)n_general * ((size_t)s.max_len1_general + 2) * sizeof(int));
            kswv_general_kernel<<<(n_general + 63) / 64, 64, 0, q>>>(io, seq, seq, d_aln, s.d_sorted, gen_first, n_general, key.c,
                                                                     key.is8, 0, (int2 *)s.scratch.p, (int *)((char *)s.scratch.p + he_b), d_cells);
            ++launches;
        }
    }
    CK(cudaGetLastError());
    const size_t res_b = (size_t)n * (key.mode == MODE_BSW ? 24 : sizeof(bsw_kswr_t));
    CK(cudaMemcpyAsync(s.h_res, s.d_res, res_b, cudaMemcpyDeviceToHost, q));
    CK(cudaEventRecord(s.done_ev, q));
    std::lock_guard<std::mutex> lk(mu);
    st.launches += launches;
    if (use_wide) st.wide_batches++;
    st.h2d_bytes += (long long)n * 24 + s.nbytes;
    st.d2h_bytes += (long long)res_b;
}

} /* namespace bswh */
