// tests/emu/emu.cpp -- TEST-ONLY host build of the kernel's per-thread DP
// (bwa-mem2-arm_b200/csrc/bsw_core.cuh, kswv_core.cuh) with the DPX intrinsics
// emulated in plain C.  Lets `pytest -m "not gpu"` check the kernel LOGIC against
// the oracle in a container without a GPU.  Never linked into libbswcuda.so.
#include <vector>
#include <cstring>
#include <cstdint>
#include <algorithm>
#include "warp_emu.h"
#include "../../bwa-mem2-arm_b200/csrc/bsw_core.cuh"
#include "../../bwa-mem2-arm_b200/csrc/bsw_core2.cuh"
#include "../../bwa-mem2-arm_b200/csrc/kswv_core.cuh"
#include "../../bwa-mem2-arm_b200/csrc/bsw_wide.cuh"
#include "../../bwa-mem2-arm_b200/csrc/kswv_sys.cuh"
#include "../../bwa-mem2-arm_b200/csrc/global_core.cuh"
#include "../../include/bsw_cuda.h"

using namespace bswk;

static Const mk(int o_del, int e_del, int o_ins, int e_ins, int zdrop, int end_bonus, int a, int b, int ambig)
{
    Const c; c.o_del = o_del; c.e_del = e_del; c.o_ins = o_ins; c.e_ins = e_ins;
    c.oe_del = o_del + e_del; c.oe_ins = o_ins + e_ins; c.zdrop = zdrop; c.end_bonus = end_bonus;
    c.a = a; c.b = b; c.ambig = ambig; c.maxsc = a > 0 ? a : 0;
    return c;
}

static void put(bsw_seqpair_t &sp, const PairOut &o)
{ sp.score = o.score; sp.tle = o.tle; sp.gtle = o.gtle; sp.qle = o.qle; sp.gscore = o.gscore; sp.max_off = o.max_off; }

// lanes-per-pair kernel (bsw_wide.cuh): one emulated warp per 32 / G pairs.  min_cfg forces a larger configuration than
// the query needs (more idle lanes, same results).  Returns the number of pairs it took; the others are left untouched.
template <int G, int R>
static void wide_warp(const Const &c, int w, const bsw_seqpair_t *pairs, const int *ids, int cnt, const uint8_t *ref,
                      const uint8_t *qer, bsw_seqpair_t *out)
{
    PairOut res[32 / G];
    warp_emu::run([&](int lane) {
        const int g = lane / G, gl = lane % G;
        PairIn P = { 0, 0, 0, 0, 0, 0 };
        if (g < cnt) { const bsw_seqpair_t &p = pairs[ids[g]]; P = { p.idr, p.idq, p.len1, p.len2, p.h0, 1 }; }
        PairOut o;
        if (c.oe_del == c.oe_ins) bsw_wide_lane<G, R, true>(c, w, P, ref, qer, gl, o);
        else bsw_wide_lane<G, R, false>(c, w, P, ref, qer, gl, o);
        if (gl == 0 && g < cnt) res[g] = o;
    });
    for (int g = 0; g < cnt; ++g) put(out[ids[g]], res[g]);
}

int emu_bsw_wide_impl(bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, int n,
                            int o_del, int e_del, int o_ins, int e_ins, int zdrop, int end_bonus,
                            int a, int b, int ambig, int w, int min_cfg, unsigned char *taken)
{
    Const c = mk(o_del, e_del, o_ins, e_ins, zdrop, end_bonus, a, b, ambig);
    std::vector<int> by_cfg[BSW_WIDE_NCFG];
    const bool latency = (min_cfg & 0x100) != 0;      // the shapes with more lanes per pair (bsw_kernels.cuh wide_latency_shape)
    min_cfg &= 0xff;
    int tot = 0;
    for (int i = 0; i < n; ++i) {
        taken[i] = 0;
        if (!wide_eligible(c, pairs[i].len1, pairs[i].len2, pairs[i].h0)) continue;
        int k = wide_cfg_for(pairs[i].len2);
        if (k < 0) continue;
        if (k < min_cfg) k = min_cfg;
        by_cfg[k].push_back(i); taken[i] = 1; ++tot;
    }
    for (int k = 0; k < BSW_WIDE_NCFG; ++k) {
        const std::vector<int> &v = by_cfg[k];
        // the library's shape table (bsw_kernels.cuh): 0..7 = the configurations, 8..10 = their latency alternates
        const int shape = latency ? (k == 0 ? 8 : k <= 2 ? 9 : k <= 4 ? 10 : k) : k;
        const int lanes = shape <= 7 ? wide_cfg_lanes(shape) : shape == 8 ? 8 : shape == 9 ? 16 : 32;
        const int per = 32 / lanes;
        for (size_t f = 0; f < v.size(); f += per) {
            const int cnt = (int)std::min<size_t>(per, v.size() - f);
            switch (shape) {
            case 0: wide_warp<4, 10>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            case 1: wide_warp<4, 16>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            case 2: wide_warp<8, 9>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            case 3: wide_warp<8, 16>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            case 4: wide_warp<16, 9>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            case 5: wide_warp<16, 16>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            case 6: wide_warp<32, 9>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            case 7: wide_warp<32, 16>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            case 8: wide_warp<8, 5>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            case 9: wide_warp<16, 5>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            default: wide_warp<32, 5>(c, w, pairs, v.data() + f, cnt, ref, qer, pairs); break;
            }
        }
    }
    return tot;
}

extern "C" {

// returns number of pairs that went through the packed path; cells_out nullable
int emu_bsw(bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, int n,
            int o_del, int e_del, int o_ins, int e_ins, int zdrop, int end_bonus,
            int a, int b, int ambig, int w, int force_general, long long *cells_out)
{
    const bool count_cells = cells_out != nullptr;
    const bool force_qn = (force_general & 2) != 0;   // bit 1: run every thread through the general-selector variant
    force_general &= 1;
    Const c = mk(o_del, e_del, o_ins, e_ins, zdrop, end_bonus, a, b, ambig);
    std::vector<int> el, ge;
    for (int i = 0; i < n; ++i)
        (!force_general && packed_eligible(c, pairs[i].len1, pairs[i].len2, pairs[i].h0) ? el : ge).push_back(i);
    for (size_t k = 0; k < el.size(); k += 2) {
        PairIn A, B; PairOut oa, ob;
        int ia = el[k], ib = k + 1 < el.size() ? el[k + 1] : -1;
        A = { pairs[ia].idr, pairs[ia].idq, pairs[ia].len1, pairs[ia].len2, pairs[ia].h0, 1 };
        if (ib >= 0) B = { pairs[ib].idr, pairs[ib].idq, pairs[ib].len1, pairs[ib].len2, pairs[ib].h0, 1 };
        else B = { 0, 0, 0, 0, 0, 0 };
        int qm = A.len2 > B.len2 ? A.len2 : B.len2;
        std::vector<uint32_t> he(qm + 8); std::vector<uint16_t> qs(qm + 8);
        const bool sym = c.oe_del == c.oe_ins;
        {
            // the library's prep_kernel routes pairs whose query holds an N (or an invalid base) to the QN = true variant;
            // the cell-counting variant always runs QN = true (bsw_api.cu launch_bsw_dp)
            uint32_t lut5[5]; for (int t = 0; t < 5; ++t) lut5[t] = bsw_lutrow(((uint32_t)(-c.b) & 0xffu) * 0x01010101u, ((uint32_t)c.a ^ (uint32_t)(-c.b)) & 0xffu, lut_hi(c), t * 8u);
            bool qn = false;
            for (int j = 0; j < A.len2; ++j) qn |= qer[A.idq + j] > 3;
            for (int j = 0; ib >= 0 && j < B.len2; ++j) qn |= qer[B.idq + j] > 3;
            if (force_qn || count_cells) qn = true;
#define EMU_RUN(SYM, QN, COUNT) bsw_packed_thread2<1, SYM, QN, COUNT>(c, w, A, B, ref, qer, he.data(), qs.data(), lut5, oa, ob)
            if (count_cells) { if (sym) EMU_RUN(true, true, true); else EMU_RUN(false, true, true); }
            else if (sym) { if (qn) EMU_RUN(true, true, false); else EMU_RUN(true, false, false); }
            else { if (qn) EMU_RUN(false, true, false); else EMU_RUN(false, false, false); }
#undef EMU_RUN
        }
        put(pairs[ia], oa); if (cells_out) cells_out[ia] = oa.cells;
        if (ib >= 0) { put(pairs[ib], ob); if (cells_out) cells_out[ib] = ob.cells; }
    }
    for (int i : ge) {
        PairIn P = { pairs[i].idr, pairs[i].idq, pairs[i].len1, pairs[i].len2, pairs[i].h0, 1 };
        PairOut o;
        std::vector<int2> eh(P.len2 + 2);
        bsw_general_thread<int>(c, w, P, ref, qer, eh.data(), 1, o);
        put(pairs[i], o); if (cells_out) cells_out[i] = o.cells;
    }
    return (int)el.size();
}


int emu_bsw_wide(bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, int n,
                 int o_del, int e_del, int o_ins, int e_ins, int zdrop, int end_bonus,
                 int a, int b, int ambig, int w, int min_cfg, unsigned char *taken)
{ return emu_bsw_wide_impl(pairs, ref, qer, n, o_del, e_del, o_ins, e_ins, zdrop, end_bonus, a, b, ambig, w, min_cfg, taken); }

static void kput(bsw_kswr_t *aln, const bsw_seqpair_t &sp, int i, const KswOut &o, int phase)
{
    if (phase == 0) { bsw_kswr_t r = { o.score, o.te, o.qe, o.score2, o.te2, -1, -1 }; aln[i] = r; }
    else { bsw_kswr_t *a = &aln[sp.regid]; if (a->score == o.score) { a->tb = a->te - o.te; a->qb = a->qe - o.qe; } }
}

}   // extern "C"

// registers-and-shuffles kernel (kswv_sys.cuh): one emulated warp per 32 / G groups of two pairs
template <int G, int C, bool SYM, bool QN, bool W = false>
static void sys_warp(const Const &c, int is8, const bsw_seqpair_t *pairs, const int *ids, int cnt, const uint8_t *ref,
                     const uint8_t *qer, KswOut *res)
{
    constexpr int NG = 32 / G;
    int maxrows = 0;
    for (int k = 0; k < cnt; ++k) maxrows = std::max(maxrows, pairs[ids[k]].len1);
    const int cap = kswv_sys_list_cap(maxrows);
    std::vector<uint32_t> bl((size_t)NG * 2 * cap, 0xdeadbeefu);
    warp_emu::run([&](int lane) {
        const int g = lane / G, gl = lane % G;
        PairIn P[2];
        for (int u = 0; u < 2; ++u) {
            const int k = g * 2 + u;
            if (k < cnt) { const bsw_seqpair_t &p = pairs[ids[k]]; P[u] = { p.idr, p.idq, p.len1, p.len2, p.h0, 1 }; }
            else P[u] = { 0, 0, 0, 0, 0, 0 };
        }
        KswOut oa, ob;
        kswv_sys_lane<G, C, SYM, QN, W>(c, is8, P[0], P[1], ref, qer, gl, bl.data() + (size_t)(g * 2) * cap, bl.data() + (size_t)(g * 2 + 1) * cap, oa, ob);
        if (gl == G - 1) { if (g * 2 < cnt) res[g * 2] = oa; if (g * 2 + 1 < cnt) res[g * 2 + 1] = ob; }
    });
}

template <int G, int C>
static void sys_warp_v(const Const &c, int is8, bool qn, const bsw_seqpair_t *pairs, const int *ids, int cnt, const uint8_t *ref,
                       const uint8_t *qer, KswOut *res)
{
    const bool sym = c.oe_del == c.oe_ins;
    if (sym) { if (qn) sys_warp<G, C, true, true>(c, is8, pairs, ids, cnt, ref, qer, res); else sys_warp<G, C, true, false>(c, is8, pairs, ids, cnt, ref, qer, res); }
    else { if (qn) sys_warp<G, C, false, true>(c, is8, pairs, ids, cnt, ref, qer, res); else sys_warp<G, C, false, false>(c, is8, pairs, ids, cnt, ref, qer, res); }
}

extern "C" {

// the 16-bit tier (kswv_sys.cuh, W): configuration cfg, general score table
static void sysw_warp(int cfg, const Const &c, int is8, const bsw_seqpair_t *pairs, const int *ids, int cnt, const uint8_t *ref,
                      const uint8_t *qer, KswOut *res)
{
    const bool sym = c.oe_del == c.oe_ins;
#define SYSW(G, CC) if (sym) sys_warp<G, CC, true, true, true>(c, is8, pairs, ids, cnt, ref, qer, res); \
                    else sys_warp<G, CC, false, true, true>(c, is8, pairs, ids, cnt, ref, qer, res)
    switch (cfg) {
    case 0: SYSW(8, 8); break;   case 1: SYSW(8, 16); break;  case 2: SYSW(8, 20); break;  case 3: SYSW(8, 24); break;
    case 4: SYSW(8, 32); break;  case 5: SYSW(16, 20); break; case 6: SYSW(16, 26); break; default: SYSW(16, 32); break;
    }
#undef SYSW
}

int emu_kswv(const bsw_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, bsw_kswr_t *aln, int n,
             int o_del, int e_del, int o_ins, int e_ins, int a, int b, int ambig,
             int is8, int phase, int force_general, long long *cells_out)
{
    Const c = mk(o_del, e_del, o_ins, e_ins, 0, 0, a, b, ambig);
    std::vector<int> ge;
    if ((force_general & 0xff) != 1) {
        // the library's routing: the registers-and-shuffles kernel (kswv_sys.cuh) for every pair it admits, the general kernel
        // for the rest (force_general = 1: the general kernel for all).  Bits 8..15:
        // minimum columns per lane (a larger shape than the query needs: more pad columns, same results); bit 16: run
        // every group through the QN variant; bits 17..18: lanes per group.  Pairs are taken in input order, two per group, like the sorted order
        // of the library only that neighbouring lengths differ more.
        const int min_c = (force_general >> 8) & 0xff;
        const bool all_qn = (force_general >> 16) & 1;
        const int lanes = 8 << ((force_general >> 17) & 3);      // bits 17..18: 8 / 16 / 32 lanes per group (the small-batch shapes)
        std::vector<int> by[2][17];
        std::vector<int> byw[KSWV_SYSW_NCFG];
        const bool force_w = (force_general >> 19) & 1;      // bit 19: every admissible pair through the 16-bit shapes
        for (int i = 0; i < n; ++i) {
            if (force_w || !kswv_sys_eligible(c, pairs[i].len1, pairs[i].len2)) {
                if (kswv_sysw_eligible(c, is8, pairs[i].len1, pairs[i].len2)) byw[kswv_sysw_cfg_for(pairs[i].len2)].push_back(i);
                else ge.push_back(i);
                continue;
            }
            bool qn = all_qn;
            for (int j = 0; j < pairs[i].len2; ++j) qn |= qer[pairs[i].idq + j] > 3;
            int qb = (pairs[i].len2 + 15) >> 4; qb = qb < 1 ? 1 : qb;
            while (2 * qb < min_c && qb < 16) ++qb;
            by[qn][qb].push_back(i);
        }
        int tot = 0;
        for (int qn = 0; qn < 2; ++qn)
            for (int qb = 1; qb <= 16; ++qb) {
                const std::vector<int> &v = by[qn][qb];
                const int per = 2 * (32 / lanes);
                for (size_t f = 0; f < v.size(); f += per) {
                    const int cnt = (int)std::min<size_t>(per, v.size() - f);
                    KswOut res[8];
                    if (lanes == 32) {
                        switch ((qb + 1) / 2) {      // 32-column steps
#define SYS_CASE(CC) case CC: sys_warp_v<32, CC>(c, is8, qn != 0, pairs, v.data() + f, cnt, ref, qer, res); break;
                        SYS_CASE(1) SYS_CASE(2) SYS_CASE(3) SYS_CASE(4) SYS_CASE(5) SYS_CASE(6) SYS_CASE(7) SYS_CASE(8)
#undef SYS_CASE
                        }
                    } else if (lanes == 16) {
                        switch ((qb + 1) / 2) {
#define SYS_CASE(CC) case CC: sys_warp_v<16, 2 * CC>(c, is8, qn != 0, pairs, v.data() + f, cnt, ref, qer, res); break;
                        SYS_CASE(1) SYS_CASE(2) SYS_CASE(3) SYS_CASE(4) SYS_CASE(5) SYS_CASE(6) SYS_CASE(7) SYS_CASE(8)
#undef SYS_CASE
                        }
                    } else
                    switch (qb) {
#define SYS_CASE(QB) case QB: sys_warp_v<8, 2 * QB>(c, is8, qn != 0, pairs, v.data() + f, cnt, ref, qer, res); break;
                    SYS_CASE(1) SYS_CASE(2) SYS_CASE(3) SYS_CASE(4) SYS_CASE(5) SYS_CASE(6) SYS_CASE(7) SYS_CASE(8)
                    SYS_CASE(9) SYS_CASE(10) SYS_CASE(11) SYS_CASE(12) SYS_CASE(13) SYS_CASE(14) SYS_CASE(15) SYS_CASE(16)
#undef SYS_CASE
                    }
                    for (int k = 0; k < cnt; ++k) {
                        const int i = v[f + k];
                        kput(aln, pairs[i], i, res[k], phase); if (cells_out) cells_out[i] = res[k].cells;
                    }
                    tot += cnt;
                }
            }
        for (int cfg = 0; cfg < KSWV_SYSW_NCFG; ++cfg) {
            const std::vector<int> &v = byw[cfg];
            const int per = cfg < 5 ? 8 : 4;
            for (size_t f = 0; f < v.size(); f += per) {
                const int cnt = (int)std::min<size_t>(per, v.size() - f);
                KswOut res[8];
                sysw_warp(cfg, c, is8, pairs, v.data() + f, cnt, ref, qer, res);
                for (int k = 0; k < cnt; ++k) {
                    const int i = v[f + k];
                    kput(aln, pairs[i], i, res[k], phase); if (cells_out) cells_out[i] = res[k].cells;
                }
                tot += cnt;
            }
        }
        for (int i : ge) {
            PairIn P = { pairs[i].idr, pairs[i].idq, pairs[i].len1, pairs[i].len2, pairs[i].h0, 1 };
            KswOut o;
            std::vector<int2> he(P.len2 + 2); std::vector<int> rm(P.len1 + 2);
            kswv_general_thread<int>(c, is8, P, ref, qer, he.data(), 1, rm.data(), 1, o);
            kput(aln, pairs[i], i, o, phase); if (cells_out) cells_out[i] = o.cells;
        }
        return tot;
    }
    for (int i = 0; i < n; ++i) ge.push_back(i);
    for (int i : ge) {
        PairIn P = { pairs[i].idr, pairs[i].idq, pairs[i].len1, pairs[i].len2, pairs[i].h0, 1 };
        KswOut o;
        std::vector<int2> he(P.len2 + 2); std::vector<int> rm(P.len1 + 2);
        kswv_general_thread<int>(c, is8, P, ref, qer, he.data(), 1, rm.data(), 1, o);
        kput(aln, pairs[i], i, o, phase); if (cells_out) cells_out[i] = o.cells;
    }
    return 0;
}


// CIGAR generation (global_core.cuh): jobs as the library lays them out -- `nthreads` jobs per "launch" sharing an interleaved
// rolling-row scratch of M = 2 * max_w + 2 slots, direction words per job, count pass, scan, emit pass.  Returns total ops.
long long emu_global2(bsw_global_job_t *jobs, int n, const uint8_t *ref, const uint8_t *qer,
                      int o_del, int e_del, int o_ins, int e_ins, int a, int b, int ambig, uint32_t *cigar, long long cap, int nthreads)
{
    Const c = mk(o_del, e_del, o_ins, e_ins, 0, 0, a, b, ambig);
    long long total = 0;
    for (int first = 0; first < n; first += nthreads) {
        const int cnt = std::min(nthreads, n - first);
        int max_w = 0;
        for (int t = 0; t < cnt; ++t) max_w = std::max(max_w, jobs[first + t].w);
        const int M = 2 * max_w + 2;
        std::vector<int2> eh((size_t)M * cnt);
        std::vector<std::vector<uint32_t>> z(cnt);
        for (int t = 0; t < cnt; ++t) {
            bsw_global_job_t &g = jobs[first + t];
            z[t].assign((size_t)global2_zwords(g.tlen, g.qlen, g.w) + 1, 0xdeadbeefu);
            long long cl = 0;
            g.score = global2_dp(c, g.tlen, g.qlen, g.w, ref + g.idr, qer + g.idq, eh.data() + t, (long long)cnt, M, z[t].data(), cl);
            g.n_cigar = global2_backtrack<false>(g.tlen, g.qlen, g.w, z[t].data(), nullptr, 0);
        }
        for (int t = 0; t < cnt; ++t) {
            bsw_global_job_t &g = jobs[first + t];
            g.cigar_off = (int)total;
            if (total + g.n_cigar <= cap) global2_backtrack<true>(g.tlen, g.qlen, g.w, z[t].data(), cigar + total, g.n_cigar);
            total += g.n_cigar;
        }
    }
    return total;
}

}
