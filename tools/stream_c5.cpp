// tools/stream_c5.cpp -- BASELINE.json config[4] ("C5") stand-in and the drop-in scenario of SURVEY.md 8b:
// T host threads, each pulling 512-read tasks and doing what mem_chain2aln_across_reads_V2 does around the
// boundary (bwamem.cpp) -- through the reference-named C++ shim classes of
// bwa-mem2-arm_b200/bandedSWA_b200.h, so this file is also the compile + run test of that header:
//   per task: BandedPairWiseSW bswLeft(...), bswRight(...)           (constructed per task, as upstream does)
//             pack the task's SeqPairs + bases into the thread's own arenas (pageable malloc or pinned)
//             left : tier split -> getScores8 + getScores16 with band w; retry with 2w for the pairs with
//                    score != previous && max_off >= (w >> 1) + (w >> 2)              (MAX_BAND_TRY = 2)
//             right: h0 = score of the read's left extension; same two attempts
// `--mode kswv` replays mem_sam_pe_batch instead (bwamem_pair.cpp): per task a kswv object,
// phase 0 (getScores8 then getScores16), the host hand-off (reversed prefixes, h0 = KSW_XSTOP | score),
// phase 1 (getScores16 then getScores8).
// Input: a binary workload written by bench.py; output: the result records, for an all-pairs compare
// with the CPU restatement in bench.py.  This program never touches the CPU restatement itself.
// Build: tools/Makefile (g++, links libbswcuda.so).
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include "../bwa-mem2-arm_b200/bandedSWA_b200.h"

typedef std::chrono::steady_clock clk;
static double secs(clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); }

struct Workload {
    int64_t nL = 0, nR = 0, nT = 0, lrefB = 0, lqerB = 0, rrefB = 0, rqerB = 0, mode = 0;
    std::vector<int64_t> ltask, rtask;
    std::vector<SeqPair> L, R;            // kswv mode: L = the pairs, R unused
    std::vector<uint8_t> lref, lqer, rref, rqer;
    std::vector<kswr_t> aln;              // kswv results
};

template <class T> static void rd(FILE *f, std::vector<T> &v, size_t n)
{
    v.resize(n);
    if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short workload file\n"); exit(2); }
}

static void load(const char *path, Workload &w)
{
    FILE *f = fopen(path, "rb");
    if (!f) { perror(path); exit(2); }
    int64_t h[9];
    if (fread(h, 8, 9, f) != 9 || h[0] != 0x35434d4145525453ll) { fprintf(stderr, "bad workload header\n"); exit(2); }
    w.mode = h[1]; w.nL = h[2]; w.nR = h[3]; w.nT = h[4]; w.lrefB = h[5]; w.lqerB = h[6]; w.rrefB = h[7]; w.rqerB = h[8];
    rd(f, w.ltask, (size_t)w.nT + 1); rd(f, w.rtask, (size_t)w.nT + 1);
    rd(f, w.L, (size_t)w.nL); rd(f, w.R, (size_t)w.nR);
    rd(f, w.lref, (size_t)w.lrefB); rd(f, w.lqer, (size_t)w.lqerB); rd(f, w.rref, (size_t)w.rrefB); rd(f, w.rqer, (size_t)w.rqerB);
    fclose(f);
}

struct Arena {               // a worker thread's own buffers (mem_cache of bwa-mem2), pageable or pinned
    bool pinned;
    std::vector<void *> blocks;
    explicit Arena(bool p) : pinned(p) {}
    void *get(size_t bytes)
    {
        void *p = pinned ? bsw_host_alloc(bytes) : malloc(bytes);
        if (!p) { fprintf(stderr, "out of memory\n"); exit(2); }
        blocks.push_back(p);
        return p;
    }
    ~Arena() { for (void *p : blocks) { if (pinned) bsw_host_free(p); else free(p); } }
};

struct Tally { double create_s = 0, call_s = 0; long calls = 0, creates = 0, pairs = 0, retried = 0; };

struct Opt { int o_del = 6, e_del = 1, o_ins = 6, e_ins = 1, zdrop = 100, clip5 = 5, clip3 = 5, a = 1, b = 4, w = 100, min_seed_len = 19; };

// tier rule of the caller (SURVEY.md A.1 note 2): 8-bit / 16-bit / scalar
static int tier_of(const SeqPair &p, int a)
{
    const int mn = p.h0 + (p.len1 < p.len2 ? p.len1 : p.len2) * a;
    if (p.len1 < MAX_SEQ_LEN8 && p.len2 < MAX_SEQ_LEN8 && mn < MAX_SEQ_LEN8) return 0;
    if (p.len1 < MAX_SEQ_LEN16 && p.len2 < MAX_SEQ_LEN16 && mn < MAX_SEQ_LEN16) return 1;
    return 2;
}

// one side of a task: the two attempts of the band loop over `n` pairs held in the thread's arenas
static void run_side(BandedPairWiseSW &bsw, SeqPair *p, int n, uint8_t *ref, uint8_t *qer, const Opt &o, Tally &t,
                     std::vector<SeqPair> &t8, std::vector<SeqPair> &t16, std::vector<SeqPair> &t1, std::vector<int> &where,
                     std::vector<int> &todo, std::vector<int> &prev)
{
    todo.resize(n); prev.assign(n, 0);
    for (int i = 0; i < n; ++i) todo[i] = i;
    for (int attempt = 0; attempt < 2 && !todo.empty(); ++attempt) {            // MAX_BAND_TRY = 2
        const int w = o.w << attempt;
        t8.clear(); t16.clear(); t1.clear(); where.resize(todo.size());
        for (size_t k = 0; k < todo.size(); ++k) {
            const SeqPair &sp = p[todo[k]];
            prev[todo[k]] = sp.score;
            const int tr = tier_of(sp, o.a);
            std::vector<SeqPair> &v = tr == 0 ? t8 : tr == 1 ? t16 : t1;
            where[k] = (int)v.size() | (tr << 28);
            v.push_back(sp);
        }
        auto t0 = clk::now();
        bsw.getScores8(t8.data(), ref, qer, (int32_t)t8.size(), 1, w);
        bsw.getScores16(t16.data(), ref, qer, (int32_t)t16.size(), 1, w);
        if (!t1.empty()) bsw.scalarBandedSWAWrapper(t1.data(), ref, qer, (int)t1.size(), 1, w);
        t.call_s += secs(t0, clk::now()); t.calls += 2 + !t1.empty(); t.pairs += (long)todo.size();
        if (attempt) t.retried += (long)todo.size();
        size_t keep = 0;
        for (size_t k = 0; k < todo.size(); ++k) {
            const int tr = where[k] >> 28, ix = where[k] & 0x0fffffff, i = todo[k];
            p[i] = (tr == 0 ? t8 : tr == 1 ? t16 : t1)[ix];
            if (p[i].score != prev[i] && p[i].max_off >= (w >> 1) + (w >> 2)) todo[keep++] = i;
        }
        todo.resize(keep);
    }
}

static void bsw_worker(Workload &W, const Opt &o, std::atomic<int64_t> &next, bool pinned, bool per_task, Tally &t)
{
    Arena ar(pinned);
    size_t maxp = 0, maxrb = 0, maxqb = 0;
    for (int64_t k = 0; k < W.nT; ++k) {
        maxp = std::max<size_t>(maxp, std::max(W.ltask[k + 1] - W.ltask[k], W.rtask[k + 1] - W.rtask[k]));
        size_t rb = 0, qb = 0;
        for (int64_t i = W.ltask[k]; i < W.ltask[k + 1]; ++i) { rb += W.L[i].len1; qb += W.L[i].len2; }
        maxrb = std::max(maxrb, rb); maxqb = std::max(maxqb, qb);
        rb = qb = 0;
        for (int64_t i = W.rtask[k]; i < W.rtask[k + 1]; ++i) { rb += W.R[i].len1; qb += W.R[i].len2; }
        maxrb = std::max(maxrb, rb); maxqb = std::max(maxqb, qb);
    }
    SeqPair *lp = (SeqPair *)ar.get((maxp + 64) * sizeof(SeqPair)), *rp = (SeqPair *)ar.get((maxp + 64) * sizeof(SeqPair));
    uint8_t *lr = (uint8_t *)ar.get(maxrb + 64), *lq = (uint8_t *)ar.get(maxqb + 64);
    uint8_t *rr = (uint8_t *)ar.get(maxrb + 64), *rq = (uint8_t *)ar.get(maxqb + 64);
    std::vector<SeqPair> t8, t16, t1; std::vector<int> where, todo, prev;
    BandedPairWiseSW *keepL = nullptr, *keepR = nullptr;
    const int8_t mat[25] = { 0 };
    for (;;) {
        const int64_t k = next.fetch_add(1);
        if (k >= W.nT) break;
        auto t0 = clk::now();
        BandedPairWiseSW *bl = keepL, *br = keepR;
        if (!bl) {
            bl = new BandedPairWiseSW(o.o_del, o.e_del, o.o_ins, o.e_ins, o.zdrop, o.clip5, mat, (int8_t)o.a, (int8_t)o.b, 1);
            br = new BandedPairWiseSW(o.o_del, o.e_del, o.o_ins, o.e_ins, o.zdrop, o.clip3, mat, (int8_t)o.a, (int8_t)o.b, 1);
            t.creates += 2;
        }
        t.create_s += secs(t0, clk::now());
        // pack the task into the thread's arenas (what the caller does before the boundary)
        const int64_t l0 = W.ltask[k], nl = W.ltask[k + 1] - l0, r0 = W.rtask[k], nr = W.rtask[k + 1] - r0;
        int ro = 0, qo = 0;
        for (int64_t i = 0; i < nl; ++i) {
            SeqPair sp = W.L[l0 + i];
            memcpy(lr + ro, W.lref.data() + sp.idr, (size_t)sp.len1); memcpy(lq + qo, W.lqer.data() + sp.idq, (size_t)sp.len2);
            sp.idr = ro; sp.idq = qo; ro += sp.len1; qo += sp.len2;
            lp[i] = sp;
        }
        run_side(*bl, lp, (int)nl, lr, lq, o, t, t8, t16, t1, where, todo, prev);
        ro = qo = 0;
        for (int64_t i = 0; i < nr; ++i) {
            SeqPair sp = W.R[r0 + i];
            memcpy(rr + ro, W.rref.data() + sp.idr, (size_t)sp.len1); memcpy(rq + qo, W.rqer.data() + sp.idq, (size_t)sp.len2);
            sp.idr = ro; sp.idq = qo; ro += sp.len1; qo += sp.len2;
            if (sp.regid >= 0) sp.h0 = lp[sp.regid - l0].score;             // right h0 = left score (row a9)
            rp[i] = sp;
        }
        run_side(*br, rp, (int)nr, rr, rq, o, t, t8, t16, t1, where, todo, prev);
        for (int64_t i = 0; i < nl; ++i) memcpy(&W.L[l0 + i].score, &lp[i].score, 24);
        for (int64_t i = 0; i < nr; ++i) { W.R[r0 + i].h0 = rp[i].h0; memcpy(&W.R[r0 + i].score, &rp[i].score, 24); }
        t0 = clk::now();
        if (per_task) { delete bl; delete br; } else { keepL = bl; keepR = br; }
        t.create_s += secs(t0, clk::now());
    }
    delete keepL; delete keepR;
}

static void kswv_worker(Workload &W, const Opt &o, std::atomic<int64_t> &next, bool pinned, bool per_task, Tally &t)
{
    Arena ar(pinned);
    size_t maxp = 0, maxrb = 0, maxqb = 0;
    for (int64_t k = 0; k < W.nT; ++k) {
        maxp = std::max<size_t>(maxp, W.ltask[k + 1] - W.ltask[k]);
        size_t rb = 0, qb = 0;
        for (int64_t i = W.ltask[k]; i < W.ltask[k + 1]; ++i) { rb += W.L[i].len1; qb += W.L[i].len2; }
        maxrb = std::max(maxrb, rb); maxqb = std::max(maxqb, qb);
    }
    SeqPair *p = (SeqPair *)ar.get((maxp + 64) * sizeof(SeqPair)), *p1 = (SeqPair *)ar.get((maxp + 64) * sizeof(SeqPair));
    kswr_t *aln = (kswr_t *)ar.get((maxp + 64) * sizeof(kswr_t));
    uint8_t *r = (uint8_t *)ar.get(maxrb + 64), *q = (uint8_t *)ar.get(maxqb + 64);
    uint8_t *r1 = (uint8_t *)ar.get(maxrb + 64), *q1 = (uint8_t *)ar.get(maxqb + 64);
    std::vector<int> order;
    kswv *keep = nullptr;
    for (;;) {
        const int64_t k = next.fetch_add(1);
        if (k >= W.nT) break;
        auto t0 = clk::now();
        kswv *ks = keep ? keep : new kswv(o.o_del, o.e_del, o.o_ins, o.e_ins, (int8_t)o.a, (int8_t)o.b, 1, 2048, 256);
        if (!keep) t.creates++;
        t.create_s += secs(t0, clk::now());
        const int64_t f0 = W.ltask[k]; const int n = (int)(W.ltask[k + 1] - f0);
        // sort_classify: the pairs with KSW_XBYTE first (R:docs-archive/WEEK1_BENCHMARK_RESULTS.md:97-103)
        order.clear();
        for (int i = 0; i < n; ++i) if (W.L[f0 + i].h0 & KSW_XBYTE) order.push_back(i);
        const int pcnt8 = (int)order.size();
        for (int i = 0; i < n; ++i) if (!(W.L[f0 + i].h0 & KSW_XBYTE)) order.push_back(i);
        int ro = 0, qo = 0;
        for (int i = 0; i < n; ++i) {
            SeqPair sp = W.L[f0 + order[i]];
            memcpy(r + ro, W.lref.data() + sp.idr, (size_t)sp.len1); memcpy(q + qo, W.lqer.data() + sp.idq, (size_t)sp.len2);
            sp.idr = ro; sp.idq = qo; ro += sp.len1; qo += sp.len2; sp.regid = i;
            p[i] = sp;
        }
        t0 = clk::now();
        ks->getScores8(p, r, q, aln, pcnt8, 1, 0);
        ks->getScores16(p + pcnt8, r, q, aln + pcnt8, n - pcnt8, 1, 0);
        t.call_s += secs(t0, clk::now()); t.calls += 2; t.pairs += n;
        // hand-off (bwa ksw_align2): XSTART && !(XSUBO && score < minsc) -> reversed prefixes, h0 = KSW_XSTOP | score
        int n1 = 0, n1_8 = 0; ro = qo = 0;
        for (int pass = 0; pass < 2; ++pass)                 // 8-bit pairs first so that phase 1 can split them again
            for (int i = pass ? pcnt8 : 0; i < (pass ? n : pcnt8); ++i) {
                const int xtra = p[i].h0;
                if (!(xtra & KSW_XSTART) || ((xtra & KSW_XSUBO) && aln[i].score < (xtra & 0xffff))) continue;
                SeqPair sp = p[i];
                const int tl = aln[i].te + 1, ql = aln[i].qe + 1;
                for (int j = 0; j < tl; ++j) r1[ro + j] = r[p[i].idr + tl - 1 - j];
                for (int j = 0; j < ql; ++j) q1[qo + j] = q[p[i].idq + ql - 1 - j];
                sp.idr = ro; sp.idq = qo; sp.len1 = tl; sp.len2 = ql; sp.h0 = KSW_XSTOP | aln[i].score; sp.regid = i;
                ro += tl; qo += ql;
                p1[n1++] = sp;
                if (!pass) n1_8 = n1;
            }
        t0 = clk::now();
        ks->getScores16(p1 + n1_8, r1, q1, aln, n1 - n1_8, 1, 1);
        ks->getScores8(p1, r1, q1, aln, n1_8, 1, 1);
        t.call_s += secs(t0, clk::now()); t.calls += 2; t.pairs += n1; t.retried += n1;
        for (int i = 0; i < n; ++i) W.aln[f0 + order[i]] = aln[i];
        t0 = clk::now();
        if (per_task) delete ks; else keep = ks;
        t.create_s += secs(t0, clk::now());
    }
    delete keep;
}

int main(int argc, char **argv)
{
    const char *in = nullptr, *out = nullptr;
    int threads = 8, pinned = 0, per_task = 1, reps = 1;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto val = [&]() { if (i + 1 >= argc) { fprintf(stderr, "missing value for %s\n", a.c_str()); exit(2); } return argv[++i]; };
        if (a == "--in") in = val();
        else if (a == "--out") out = val();
        else if (a == "--threads") threads = atoi(val());
        else if (a == "--pinned") pinned = atoi(val());
        else if (a == "--per-task-handles") per_task = atoi(val());
        else if (a == "--reps") reps = atoi(val());
        else { fprintf(stderr, "usage: stream_c5 --in W.bin [--out R.bin] [--threads T] [--pinned 0|1] [--per-task-handles 0|1] [--reps N]\n"); return 2; }
    }
    if (!in) { fprintf(stderr, "--in is required\n"); return 2; }
    Workload W; load(in, W);
    Opt o;
    if (W.mode == 1) W.aln.assign((size_t)W.nL, kswr_t{ 0, -1, -1, -1, -1, -1, -1 });
    {   // warm-up: first use builds the per-device runtime (streams, engine, pinned staging); not part of any figure
        const int8_t mat[25] = { 0 };
        BandedPairWiseSW w0(o.o_del, o.e_del, o.o_ins, o.e_ins, o.zdrop, o.clip5, mat, (int8_t)o.a, (int8_t)o.b, 1);
        SeqPair sp; memset(&sp, 0, sizeof sp); sp.len1 = sp.len2 = 4; sp.h0 = 5;
        uint8_t z[32] = { 0 };
        w0.getScores8(&sp, z, z, 1, 1, o.w);
        kswv k0(o.o_del, o.e_del, o.o_ins, o.e_ins, (int8_t)o.a, (int8_t)o.b, 1, 2048, 256);
        kswr_t a0; sp.h0 = KSW_XBYTE | KSW_XSUBO | 19;
        k0.getScores8(&sp, z, z, &a0, 1, 1, 0);
    }
    double best = 1e30; std::vector<Tally> tl;
    for (int rep = 0; rep < reps; ++rep) {
        std::atomic<int64_t> next(0);
        std::vector<Tally> tally((size_t)threads);
        for (int d = 0; d < bsw_device_count(); ++d) bsw_engine_reset_stats(d);
        std::vector<std::thread> th;
        auto t0 = clk::now();
        for (int t = 0; t < threads; ++t)
            th.emplace_back([&, t] {
                if (W.mode == 1) kswv_worker(W, o, next, pinned != 0, per_task != 0, tally[(size_t)t]);
                else bsw_worker(W, o, next, pinned != 0, per_task != 0, tally[(size_t)t]);
            });
        for (auto &x : th) x.join();
        const double s = secs(t0, clk::now());
        if (s < best) { best = s; tl = tally; }
    }
    Tally sum;
    for (const Tally &t : tl) { sum.create_s += t.create_s; sum.call_s += t.call_s; sum.calls += t.calls; sum.creates += t.creates; sum.pairs += t.pairs; sum.retried += t.retried; }
    bsw_engine_stats_t es[8]; int nd = bsw_device_count(); if (nd > 8) nd = 8;
    long long eb = 0, ep = 0, el = 0, emax = 0, ec = 0, eh2d = 0, ed2h = 0;
    double pf[8] = { 0 };
    for (int d = 0; d < nd; ++d) {
        bsw_engine_get_stats(d, &es[d]); eb += es[d].batches; ep += es[d].pairs; el += es[d].kernel_launches; ec += es[d].calls; eh2d += es[d].h2d_bytes; ed2h += es[d].d2h_bytes;
        emax = std::max<long long>(emax, es[d].max_batch_pairs);
        const int64_t v[8] = { es[d].scan_ns, es[d].reserve_ns, es[d].pack_ns, es[d].wait_ns, es[d].unpack_ns, es[d].packers_ns, es[d].submit_ns, es[d].flight_ns };
        for (int i = 0; i < 8; ++i) pf[i] += (double)v[i];
    }
    int used = 0; for (int d = 0; d < nd; ++d) used += es[d].batches > 0;      /* devices whose engine ran a batch */
    const double pc = ec ? 1e-3 / (double)ec : 0.0, pb = eb ? 1e-3 / (double)eb : 0.0;
    printf("{\"mode\": \"%s\", \"threads\": %d, \"pinned\": %d, \"per_task_handles\": %d, \"tasks\": %lld, \"pairs_submitted\": %ld, "
           "\"pairs_retried_or_phase1\": %ld, \"wall_s\": %.6f, \"pairs_per_s\": %.1f, \"calls\": %ld, \"us_per_call\": %.2f, "
           "\"handles_created\": %ld, \"us_per_create_destroy\": %.2f, \"engine_batches\": %lld, \"engine_pairs_per_batch\": %.1f, "
           "\"engine_max_batch\": %lld, \"engine_launches\": %lld, \"reps\": %d, "
           "\"us_per_call_scan_reserve_pack_wait_unpack\": [%.1f, %.1f, %.1f, %.1f, %.1f], "
           "\"us_per_batch_packers_submit_flight\": [%.1f, %.1f, %.1f], \"devices\": %d, "
           "\"engine_h2d_bytes\": %lld, \"engine_d2h_bytes\": %lld}\n",
           W.mode == 1 ? "kswv" : "bsw", threads, pinned, per_task, (long long)W.nT, sum.pairs, sum.retried, best, sum.pairs / best,
           sum.calls, sum.calls ? sum.call_s / sum.calls * 1e6 : 0.0, sum.creates, sum.creates ? sum.create_s / sum.creates * 1e6 : 0.0,
           eb, eb ? (double)ep / eb : 0.0, emax, el, reps, pf[0] * pc, pf[1] * pc, pf[2] * pc, pf[3] * pc, pf[4] * pc,
           pf[5] * pb, pf[6] * pb, pf[7] * pb, used, eh2d, ed2h);
    if (out) {
        FILE *f = fopen(out, "wb");
        if (!f) { perror(out); return 2; }
        if (W.mode == 1) fwrite(W.aln.data(), sizeof(kswr_t), W.aln.size(), f);
        else { fwrite(W.L.data(), sizeof(SeqPair), W.L.size(), f); fwrite(W.R.data(), sizeof(SeqPair), W.R.size(), f); }
        fclose(f);
    }
    return 0;
}
