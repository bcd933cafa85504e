/*
 * bsw_core2.cuh -- second-generation row loop of the packed banded-extension kernel
 * (BandedPairWiseSW::getScores8/16; semantics = bwa ksw_extend2).
 *
 * Same mapping, cell arithmetic and shared-memory row layout as bsw_core.cuh (one thread, two pairs,
 * s16x2 DPX ops, byte-packed state word per column); what changed is everything AROUND the cells,
 * which the r01 source-level profile showed to be 63 % of the warp time:
 *   - a row is swept in FULL groups of four columns, always up to and including column `end` of the
 *     longer lane: each lane's f chain is cut at its own exact `end` inside a "cut group" (per-cell
 *     AND masks from one PRMT), after which the lane computes zeros over zero entries.  No partial
 *     groups, no per-cell bounds, no separate store of eh[end];
 *   - the per-row bookkeeping of both lanes (band, trim, arg-max decode, z-drop, gscore) is
 *     straight-line select code; the rare events (a lane dies, a right-trim scan is needed, the lanes'
 *     start columns differ, a lane is not "clean") are collected into ONE branch per row;
 *   - target bases come from a 4-row window per lane that is refilled by a warp-uniform branch every
 *     fourth row (one aligned word load per lane, issued four rows before it is needed).
 * Rows the fast path does not admit run the masked generic cell loop (exact for any band geometry).
 * COUNT additionally tracks every lane's exact trimmed `beg` (A.1's left trim, which the sweep itself
 * replaces by the lazy common start `t`) and sums end - beg over the rows entered: the oracle's DP-cell
 * count.  Results are identical with and without it.
 *
 * Requires columns [0, max(len2)+3] of he/qs to exist (the kernel sizes the bin with 4 columns of slack).
 */
#pragma once
#include "bsw_core.cuh"

namespace bswk {

#if defined(__CUDA_ARCH__)
/* shift amounts >= 32 (incl. "negative" ones) give 0: PTX shl.b32 clamps the amount */
__device__ __forceinline__ uint32_t shl_clamp(uint32_t v, int s) { uint32_t r; asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"((uint32_t)s)); return r; }
__device__ __forceinline__ uint32_t p_minu4(uint32_t a, uint32_t b) { return __vminu4(a, b); }
__device__ __forceinline__ uint32_t funnel_r(uint32_t lo, uint32_t hi, int s) { return __funnelshift_r(lo, hi, (uint32_t)s); }
__device__ __forceinline__ uint32_t ldg32(const uint32_t *p) { return __ldg(p); }
#else
static inline uint32_t shl_clamp(uint32_t v, int s) { return (uint32_t)s < 32u ? v << s : 0u; }
static inline uint32_t p_minu4(uint32_t a, uint32_t b)
{
    uint32_t r = 0;
    for (int k = 0; k < 4; ++k) { uint32_t x = (a >> (8 * k)) & 0xff, y = (b >> (8 * k)) & 0xff; r |= (x < y ? x : y) << (8 * k); }
    return r;
}
static inline uint32_t funnel_r(uint32_t lo, uint32_t hi, int s) { s &= 31; return s ? (lo >> s) | (hi << (32 - s)) : lo; }
static inline uint32_t ldg32(const uint32_t *p) { return *p; }
#endif

/* 4-byte asynchronous global -> shared copy (LDGSTS) and the wait for all of this thread's copies */
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ void stage32(uint32_t *dst, const uint32_t *src)
#if defined(BSW2_STAGE_LDST)
{ *dst = __ldg(src); }
__device__ __forceinline__ void stage_wait() {}
#else
{ asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory"); }
__device__ __forceinline__ void stage_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }
#endif
#else
static inline void stage32(uint32_t *dst, const uint32_t *src) { *dst = *src; }
static inline void stage_wait() {}
#endif

/* Row context of the second-generation kernel.  QN = false (no N / invalid base in either query, decided by
 * prep_kernel): the substitution score of both lanes is ONE PRMT into a per-row byte table, LA = scores of lane
 * A's target base against query bases 0..3, LB = the same for lane B; the stored selector picks byte q (and its
 * sign).  QN = true: the general form of bsw_core.cuh (LOP3 index + PRMT into the constant 8-entry table). */
struct RowCtx2 : RowCtx { uint32_t LA, LB; };

template <bool SYM, bool QN>
BSW_HD void bsw_cell2(const RowCtx2 &r, uint32_t w, uint32_t q, uint32_t &f, uint32_t &h, uint32_t &en)
{
    const uint32_t e = st_e(w), hs = st_h2(w, e);
    const uint32_t s = QN ? p_perm(r.LUTLO, r.LUTHI, p_xsel(q, r.T)) : p_perm(r.LA, r.LB, q);
    const uint32_t M = p_addmin(hs, s, hs * r.GATE);
    h = p_max3(M, e, f);
    if (SYM) {
        const uint32_t u = p_add(M, r.NOD);
        en = p_addmax_relu(e, r.NED, u);
        f = p_addmax_relu(f, r.NEI, u);
    } else {
        en = p_addmax_relu(M, r.NOD, p_add(e, r.NED));
        f = p_addmax_relu(M, r.NOI, p_add(f, r.NEI));
    }
}

/* selector of query column bases (qa, qb), both already clamped to 0..4 */
template <bool QN>
BSW_HD uint16_t bsw_selector(uint32_t qa, uint32_t qb)
{
    if (QN) return (uint16_t)(((8u | qb) << 12) | (qb << 8) | ((8u | qa) << 4) | qa);
    qa &= 3u; qb &= 3u;
    return (uint16_t)(((12u | qb) << 12) | ((4u | qb) << 8) | ((8u | qa) << 4) | qa);
}

/* per-row score table of one lane (QN = false): byte q = score(t, q) for q = 0..3; t8 = 8 * target base (32 = N) */
BSW_HD uint32_t bsw_lutrow(uint32_t mis, uint32_t x, uint32_t amb, uint32_t t8)
{
    return t8 >= 32u ? amb : (mis ^ shl_clamp(x, (int)t8));
}

/* one unchecked group of four columns (both lanes): cells, state stores, row-max key */
#define BSW2_GROUP(V0, V1, V2, V3, Q0, Q1, Q2, Q3)                                                 \
    {                                                                                              \
        uint32_t ha, hb, hc, hd, ea, eb, ec, ed;                                                   \
        bsw_cell2<SYM, QN>(r, V0, Q0, f, ha, ea); hp[(k + 0) * S] = st_pack(h1, ea);                    \
        bsw_cell2<SYM, QN>(r, V1, Q1, f, hb, eb); hp[(k + 1) * S] = st_pack(ha, eb);                    \
        bsw_cell2<SYM, QN>(r, V2, Q2, f, hc, ec); hp[(k + 2) * S] = st_pack(hb, ec);                    \
        bsw_cell2<SYM, QN>(r, V3, Q3, f, hd, ed); hp[(k + 3) * S] = st_pack(hc, ed);                    \
        h1 = hd;                                                                                   \
        const uint32_t g = p_maxu(p_max3u((ha << 8) + c0, (hb << 8) + c1, (hc << 8) + c2), (hd << 8) + c3); \
        key = p_maxu(key, g + (uint32_t)k * 0x00010001u);                                          \
    }

/* one cell of a cut group: CAP = lanes whose exact `end` is this column.  Those lanes hand their running h
 * to h1save (it is eh[end].h and the epilogue's h1), get f cut, and -- when end == qlen (QF) -- h1 zeroed so
 * that entry qlen stays zero; the cell itself then stores eh[end] = { h1, 0 } and yields h = 0. */
#define BSW2_CUT_CELL(C, V, Q)                                                                     \
    {                                                                                              \
        const uint32_t cap = p_perm(ZA, ZB, 0x4400u + (uint32_t)(C) * 0x1111u);                    \
        h1save |= h1 & cap;                                                                        \
        h1 &= ~(cap & QF);                                                                         \
        f &= ~cap;                                                                                 \
        uint32_t h_, e_;                                                                           \
        bsw_cell2<SYM, QN>(r, V, Q, f, h_, e_);                                                         \
        hp[(k + (C)) * S] = st_pack(h1, e_);                                                       \
        h1 = h_;                                                                                   \
        gk[C] = (h_ << 8);                                                                         \
    }

/*
 * Sweep of columns [j0, j0 + 4*ng) for both lanes, ng = (max(cA, cB) - j0) / 4 + 1, i.e. up to and including
 * the cut column of the longer lane.  cA / cB (> j0) are the lanes' exact row ends: the f chain of a lane is
 * cut there.  Everything at or right of a lane's cut column must be zero state (clean lane) -- then the lane
 * computes zeros from its cut on.  QF: lane mask of the lanes whose end == qlen.
 * Out: h1save = per lane the h of its last in-range cell; key = max over (h << 8 | j).
 */
template <int S, bool SYM, bool QN>
BSW_HD void bsw_sweep2(const RowCtx2 &r, uint32_t *he, const uint16_t *qs, int j0, int cA, int cB, uint32_t QF,
                       uint32_t &f, uint32_t &h1, uint32_t &key, uint32_t &h1save)
{
    uint32_t *hp = he + j0 * S;
    const uint16_t *qp = qs + j0 * S;
    const uint32_t c0 = (uint32_t)j0 * 0x00010001u, c1 = c0 + 0x00010001u, c2 = c0 + 0x00020002u, c3 = c0 + 0x00030003u;
    const int dA = cA - j0, dB = cB - j0;                 /* relative cut columns, >= 1 */
    const int K1 = mn(dA, dB) & ~3, K2 = mx(dA, dB) & ~3; /* first column of the group(s) holding a cut */
    int k = 0, kend = K1;
    uint32_t v0 = hp[0], v1 = hp[S], v2 = hp[2 * S], v3 = hp[3 * S];
    uint32_t q0 = qp[0], q1 = qp[S], q2 = qp[2 * S], q3 = qp[3 * S];
    for (;;) {
        /* unchecked groups.  Each state word / selector register is reloaded with the next group's column (at the
         * latest the cut group's) right after the cell that consumed it: prefetch about one group ahead, in place,
         * with no second register set and no moves (a two-groups-per-iteration unroll was measured slower: the
         * remainder group runs once per row for the whole warp as soon as one thread has an odd group count) */
        while (k < kend) {
            uint32_t ha, hb, hc, hd, ea, eb, ec, ed;
            bsw_cell2<SYM, QN>(r, v0, q0, f, ha, ea); v0 = hp[(k + 4) * S]; q0 = qp[(k + 4) * S]; hp[(k + 0) * S] = st_pack(h1, ea);
            bsw_cell2<SYM, QN>(r, v1, q1, f, hb, eb); v1 = hp[(k + 5) * S]; q1 = qp[(k + 5) * S]; hp[(k + 1) * S] = st_pack(ha, eb);
            bsw_cell2<SYM, QN>(r, v2, q2, f, hc, ec); v2 = hp[(k + 6) * S]; q2 = qp[(k + 6) * S]; hp[(k + 2) * S] = st_pack(hb, ec);
            bsw_cell2<SYM, QN>(r, v3, q3, f, hd, ed); v3 = hp[(k + 7) * S]; q3 = qp[(k + 7) * S]; hp[(k + 3) * S] = st_pack(hc, ed);
            h1 = hd;
            /* two 3-input maxima instead of max3 + max + max: the column offset of the group is added on the FMA pipe */
            const uint32_t kk = (uint32_t)k * 0x00010001u;
            key = p_max3u(key, p_max3u((ha << 8) + c0, (hb << 8) + c1, (hc << 8) + c2) + kk, (hd << 8) + (c3 + kk));
            k += 4;
        }
        {   /* cut group */
            const uint32_t ZA = shl_clamp(0xffu, 8 * (dA - k)), ZB = shl_clamp(0xffu, 8 * (dB - k));
            uint32_t gk[4];
            BSW2_CUT_CELL(0, v0, q0)
            BSW2_CUT_CELL(1, v1, q1)
            BSW2_CUT_CELL(2, v2, q2)
            BSW2_CUT_CELL(3, v3, q3)
            const uint32_t g = p_maxu(p_max3u(gk[0] + c0, gk[1] + c1, gk[2] + c2), gk[3] + c3);
            key = p_maxu(key, g + (uint32_t)k * 0x00010001u);
        }
        if (k >= K2) break;
        k += 4; kend = K2;
        v0 = hp[k * S]; v1 = hp[(k + 1) * S]; v2 = hp[(k + 2) * S]; v3 = hp[(k + 3) * S];
        q0 = qp[k * S]; q1 = qp[(k + 1) * S]; q2 = qp[(k + 2) * S]; q3 = qp[(k + 3) * S];
    }
}

/* target bases of one lane: aligned words of the reference buffer, a window of the next four rows */
struct TWin {
    const uint32_t *wp;   /* aligned word holding the pair's first base */
    int sh, nwords;       /* bit offset of the first base in that word; words that hold at least one base */
    uint32_t lo, hi;      /* words r and r+1 at refill r */
};

BSW_HD void twin_init(TWin &t, const uint8_t *ref, int idr, int tlen, bool act)
{
    const uintptr_t a = reinterpret_cast<uintptr_t>(ref + (act ? idr : 0));
    t.wp = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
    t.sh = (int)(a & 3) * 8;
    t.nwords = act ? (int)(((a & 3) + (uintptr_t)tlen + 3) >> 2) : 0;
    t.lo = t.nwords > 0 ? ldg32(t.wp) : 0u;
    t.hi = t.nwords > 1 ? ldg32(t.wp + 1) : 0u;
}

/* refill r (row i = 4r): returns the bases of rows 4r .. 4r+3 (clamped to 0..4) and fetches word r+2 */
BSW_HD uint32_t twin_refill(TWin &t, int r)
{
    const uint32_t win = funnel_r(t.lo, t.hi, t.sh);
#if defined(__CUDA_ARCH__)
    /* rotate and load IN PLACE: written as `hi = *p` the compiler loads into a temporary and moves it, and the move
     * waits for the load right here (5 % of the kernel time in profile s1) instead of four rows later */
    asm volatile("{ .reg .pred q; setp.lt.s32 q, %3, %4; mov.b32 %0, %1; mov.b32 %1, 0; @q ld.global.nc.u32 %1, [%2]; }"
                 : "+r"(t.lo), "+r"(t.hi) : "l"(t.wp + r + 2), "r"(r + 2), "r"(t.nwords));
#else
    t.lo = t.hi;
    t.hi = (r + 2 < t.nwords) ? ldg32(t.wp + r + 2) : 0u;
#endif
    return p_minu4(win, 0x04040404u);
}

/*
 * One thread, two pairs (see bsw_core.cuh for the layout of he / qs and the exactness notes).
 */
template <int S, bool SYM, bool QN, bool COUNT>
BSW_HD void bsw_packed_thread2(const Const &c, int w_in, const PairIn &A, const PairIn &B,
                               const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                               uint32_t *he, uint16_t *qs, const uint32_t *lut5, PairOut &oa, PairOut &ob)
{
    const PairIn *in[2] = { &A, &B };
    int qlen[2], tlen[2], h0[2], wb[2], end[2], smax[2], max_i[2], max_j[2], max_ie[2], gscore[2], max_off[2], clean[2];
    int act = 0;                                   /* bit p: lane p is live */
    int beg[2] = { 0, 0 };                         /* COUNT: A.1's exact beg of each lane */
    long long cells[2] = { 0, 0 };
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const PairIn &x = *in[p];
        qlen[p] = x.valid ? x.len2 : 0;
        tlen[p] = x.valid ? x.len1 : 0;
        h0[p] = x.valid ? x.h0 : 0;
        wb[p] = eff_band(c, qlen[p], w_in);
        end[p] = qlen[p];
        smax[p] = h0[p]; max_i[p] = -1; max_j[p] = -1; max_ie[p] = -1; gscore[p] = -1; max_off[p] = 0;
        if (x.valid && x.len1 > 0) act |= 1 << p;
        clean[p] = 1;
    }

    RowCtx2 r;
    r.LUTLO = launder(lut_lo(c)); r.LUTHI = launder(lut_hi(c));
    r.LA = r.LB = 0;
    const uint32_t MISW = ((uint32_t)(-c.b) & 0xffu) * 0x01010101u, AMBW = lut_hi(c), XW = ((uint32_t)c.a ^ (uint32_t)(-c.b)) & 0xffu;
    r.NOD = pack2(-c.oe_del, -c.oe_del); r.NOI = pack2(-c.oe_ins, -c.oe_ins);
    r.NED = pack2(-c.e_del, -c.e_del);   r.NEI = pack2(-c.e_ins, -c.e_ins);
    r.GATE = (uint32_t)(c.a + 1);
    r.T = 0;

    /* ---- query selectors and first row of H (A.1), E = 0; columns qmax+1 .. qmax+3 zeroed (sweep slack) ---- */
    {
        const int qmax = mx(qlen[0], qlen[1]);
        /* (1) the raw query words of both lanes are staged in the (still unused) state rows -- lane A's words in
         *     he[0 .. nwA), lane B's behind them -- by asynchronous copies issued back to back: one memory round
         *     trip instead of one per four columns (the dependent word loads were 7 % of the kernel time, profile s2) */
        const uint32_t *wq[2]; int shq[2], nwq[2];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const PairIn &x = *in[p];
            const uintptr_t aq = reinterpret_cast<uintptr_t>(qer + (x.valid ? x.idq : 0));
            wq[p] = reinterpret_cast<const uint32_t *>(aq & ~(uintptr_t)3);
            shq[p] = (int)(aq & 3) * 8;
            nwq[p] = qlen[p] > 0 ? (int)(((aq & 3) + (uintptr_t)qlen[p] + 3) >> 2) : 0;
        }
        for (int g = 0; g < nwq[0]; ++g) stage32(he + g * S, wq[0] + g);
        for (int g = 0; g < nwq[1]; ++g) stage32(he + (nwq[0] + g) * S, wq[1] + g);
        stage_wait();
        /* (2) selectors; a word index past a lane's staged words reads other staged / unused words: masked by j < qlen */
        for (int j0 = 0; j0 <= qmax + 3; j0 += 4) {
            const int g = j0 >> 2;
            const uint32_t by0 = funnel_r(he[g * S], he[(g + 1) * S], shq[0]);
            const uint32_t by1 = funnel_r(he[(nwq[0] + g) * S], he[(nwq[0] + g + 1) * S], shq[1]);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = j0 + u;
                uint32_t qa = (j < qlen[0]) ? ((by0 >> (8 * u)) & 0xffu) : 0u;
                uint32_t qb = (j < qlen[1]) ? ((by1 >> (8 * u)) & 0xffu) : 0u;
                qa = qa > 4u ? 4u : qa; qb = qb > 4u ? 4u : qb;
                if (j <= qmax + 3) qs[j * S] = bsw_selector<QN>(qa, qb);
            }
        }
        /* (3) H(-1, j): h0, relu(h0 - oe_ins), then relu(previous - e_ins); zero right of the lane's query */
        int lastnz0 = 0, lastnz1 = 0;
        uint32_t V = pack2(h0[0], h0[1]);
        he[0] = V;
        V = p_addmax_relu(V, r.NOI, 0u);
        for (int j = 1; j <= qmax + 3; ++j) {
            V &= (j <= qlen[0] ? 0x0000ffffu : 0u) | (j <= qlen[1] ? 0xffff0000u : 0u);
            he[j * S] = V;
            lastnz0 = (V & 0x0000ffffu) ? j : lastnz0;
            lastnz1 = (V & 0xffff0000u) ? j : lastnz1;
            V = p_addmax_relu(V, r.NEI, 0u);
        }
        /* clean: every non-zero initial entry is inside row 0's range [0, min(qlen, w+1)] */
        clean[0] = lastnz0 <= mn(qlen[0], wb[0] + 1);
        clean[1] = lastnz1 <= mn(qlen[1], wb[1] + 1);
        /* the initial entry at row 0's end is dead (row 0 overwrites eh[end] before anything reads it): zero it, so that
         * "every entry at or right of a clean lane's end is zero" already holds in row 0 (bsw_sweep2 relies on it) */
        if (clean[0]) he[mn(qlen[0], wb[0] + 1) * S] &= 0xffff0000u;
        if (clean[1]) he[mn(qlen[1], wb[1] + 1) * S] &= 0x0000ffffu;
    }

    const int BIG = 0x3fffffff;

    TWin tw[2];
    twin_init(tw[0], ref, A.idr, tlen[0], (act & 1) != 0);
    twin_init(tw[1], ref, B.idr, tlen[1], (act & 2) != 0);
    uint32_t winA = 0, winB = 0;

    /* first-column boundary H(i, -1) = max(h0 - o_del - e_del*(i+1), 0) of both lanes, one packed op per row */
    uint32_t hcol = pack2(h0[0] - c.o_del, h0[1] - c.o_del);
    if (!(act & 1)) hcol &= 0xffff0000u;
    if (!(act & 2)) hcol &= 0x0000ffffu;

    int t = 0;          /* lazy common start: all entries < t are zero in both lanes */
    int i = 0;
    while (act) {
        /* ---------------- row prologue ---------------- */
        if ((i & 3) == 0) { winA = twin_refill(tw[0], i >> 2); winB = twin_refill(tw[1], i >> 2); }
        if (QN) r.T = (winA & 0xffu) * 0x0011u | (winB & 0xffu) * 0x1100u;
#if defined(BSW2_LUTSMEM)
        else {      /* per-row score tables of the two lanes: one shared-memory load each (lut5[t]) */
            r.LA = lut5[winA & 0xffu]; r.LB = lut5[winB & 0xffu];
        }
#else
        else { r.LA = bsw_lutrow(MISW, XW, AMBW, (winA & 0xffu) * 8u); r.LB = bsw_lutrow(MISW, XW, AMBW, (winB & 0xffu) * 8u); }
        (void)lut5;
#endif
        winA >>= 8; winB >>= 8;
        hcol = p_addmax_relu(hcol, r.NED, 0u);

        int lb[2], st[2];
        int empty = 0;
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            lb[p] = mx(i - wb[p], 0);
#if defined(BSW2_T25)
            end[p] = mn(end[p], i + wb[p] + 1);                 /* end <= qlen always holds (set that way below) */
            st[p] = mx(t, i - wb[p]);
#else
            end[p] = mn(mn(end[p], i + wb[p] + 1), qlen[p]);
            st[p] = mx(t, lb[p]);
#endif
            if (COUNT && ((act >> p) & 1)) {                    /* A.1: beg = max(beg, i - w); cells += end - beg */
                beg[p] = mx(beg[p], i - wb[p]);
                if (end[p] > beg[p]) cells[p] += end[p] - beg[p];
            }
            if (((act >> p) & 1) && st[p] >= end[p]) empty |= 1 << p;
        }
        if (empty) {
            /* no cell in this row for that lane: m == 0, the lane stops here (A.1: gscore check, then break) */
#pragma unroll
            for (int p = 0; p < 2; ++p)
                if (empty & (1 << p)) {
                    const int h1 = st[p] == 0 ? half_s(hcol, p) : 0;
                    const int jafter = lb[p] < end[p] ? end[p] : lb[p];
                    if (jafter == qlen[p]) {
                        max_ie[p] = gscore[p] > h1 ? max_ie[p] : i;
                        gscore[p] = gscore[p] > h1 ? gscore[p] : h1;
                    }
                    const uint32_t keep = p ? 0x0000ffffu : 0xffff0000u;
                    for (int j = 0; j <= qlen[p]; ++j) he[j * S] &= keep;
                    hcol &= keep;
                }
            act &= ~empty;
            if (!act) break;
        }
        const int both = act == 3;
        const bool la = (act & 1) != 0;                            /* lane A is live (else lane B is); selects, not array
                                                                    * indexing by a run-time lane (that put st[]/clean[] in local memory) */
        uint32_t f = 0, key = 0, h1save = 0;
        const bool fast = (both ? (clean[0] & clean[1] & (st[0] == st[1])) : (la ? clean[0] : clean[1])) != 0;
        if (fast) {
            const int j0 = la ? st[0] : st[1];
            const int cA = (act & 1) ? end[0] : end[1], cB = (act & 2) ? end[1] : end[0];   /* a dead lane mirrors the live one */
            const uint32_t QF = (end[0] == qlen[0] ? 0x0000ffffu : 0u) | (end[1] == qlen[1] ? 0xffff0000u : 0u);
            uint32_t h1 = j0 == 0 ? hcol : 0u;
            bsw_sweep2<S, SYM, QN>(r, he, qs, j0, cA, cB, QF, f, h1, key, h1save);
        } else {
            /* generic masked row: each lane exactly over its own [st, end); entries outside are left untouched */
            int sp[2], ep[2];
#pragma unroll
            for (int p = 0; p < 2; ++p) { const bool l = (act >> p) & 1; sp[p] = l ? st[p] : BIG; ep[p] = l ? end[p] : 0; }
            uint32_t h1 = pack2(sp[0] == 0 ? half_s(hcol, 0) : 0, sp[1] == 0 ? half_s(hcol, 1) : 0);
            const int ls = mn(sp[0], sp[1]), e2 = mx(ep[0], ep[1]);
            for (int j = ls; j < e2; ++j) {
                const uint32_t MK = ((j >= sp[0] && j < ep[0]) ? 0x0000ffffu : 0u) | ((j >= sp[1] && j < ep[1]) ? 0xffff0000u : 0u);
                const uint32_t v = he[j * S];
                uint32_t h_, en, fn = f;
                bsw_cell2<SYM, QN>(r, v, qs[j * S], fn, h_, en);
                he[j * S] = st_pack(h1 & MK, en & MK) | (v & ~MK);
                f = fn & MK;
                h1 = (h_ & MK) | (h1 & ~MK);
                key = p_maxu(key, (((h_ & MK) << 8) + ((uint32_t)j * 0x00010001u)) & MK);
            }
#pragma unroll
            for (int p = 0; p < 2; ++p)
                if ((act >> p) & 1) {   /* eh[end] = { h1, 0 }; entry qlen of a clean lane is kept zero (never read as a diagonal) */
                    const uint32_t hm = p ? 0xffff0000u : 0x0000ffffu;
                    const int h1p = half_s(h1, p);
                    const uint32_t v = he[end[p] * S];
                    he[end[p] * S] = (v & ~hm) | ((end[p] == qlen[p] && clean[p]) ? 0u : (pack2(h1p, h1p) & hm));
                }
            h1save = h1;
        }

        /* ---------------- row epilogue (A.1), both lanes, straight-line ---------------- */
#ifdef BSW2_TRACE
        printf("row %d fast %d st %d %d end %d %d key %08x h1save %08x act %d t %d\n", i, (int)fast, st[0], st[1], end[0], end[1], key, h1save, act, t);
#endif
        int dying = 0, scan = 0;
        const int eold[2] = { end[0], end[1] };
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const bool live = (act >> p) & 1;
            const int h1p = half_s(h1save, p), kp = half_u(key, p);
            const int m = kp >> 8, mj = kp & 0xff;
            const bool atq = live && end[p] == qlen[p];            /* a computing lane has beg < end: A.1's j after the loop is end */
            max_ie[p] = (atq && !(gscore[p] > h1p)) ? i : max_ie[p];
            gscore[p] = (atq && h1p > gscore[p]) ? h1p : gscore[p];
            const bool upd = live && m > smax[p];
            int off = mj - i; off = off < 0 ? -off : off;
            const int di = i - max_i[p], dj = mj - max_j[p];
            const int pen = di > dj ? (di - dj) * c.e_del : (dj - di) * c.e_ins;
#if defined(BSW2_ZPRE)
            /* z-drop needs max - m > zdrop at least (the gap term only subtracts): then the lane takes the rare branch, which
             * repeats this row's exact test */
            const bool zpre = !upd && c.zdrop > 0 && smax[p] - m > c.zdrop;
            const bool zd = zpre && smax[p] - m - pen > c.zdrop;
#else
            const bool zd = !upd && c.zdrop > 0 && smax[p] - m - pen > c.zdrop;
#endif
            const bool stop = (m == 0) | zd | (i + 1 >= tlen[p]);
            max_off[p] = (upd && off > max_off[p]) ? off : max_off[p];
            smax[p] = upd ? m : smax[p]; max_i[p] = upd ? i : max_i[p]; max_j[p] = upd ? mj : max_j[p];
            if (live && stop) dying |= 1 << p;
            if (live && !stop && h1p == 0) scan |= 1 << p;
            if (COUNT && live && !stop) {                          /* A.1's left trim: first non-zero entry in [beg, end) */
                const uint32_t hm = p ? 0xffff0000u : 0x0000ffffu;
                int jj = beg[p];
                while (jj < eold[p] && !(he[jj * S] & hm)) ++jj;
                beg[p] = jj;
            }
            end[p] = mn(end[p] + 2, qlen[p]);                      /* right trim when eh[end].h != 0; else the scan below */
        }
        if (scan | dying) {                                        /* the rare events of a row, one branch */
#pragma unroll
            for (int p = 0; p < 2; ++p)
                if (scan & (1 << p)) {
                    /* eh[end] is zero: find the last non-zero entry left of it (m > 0 guarantees one at or right of the row start) */
                    const uint32_t hm = p ? 0xffff0000u : 0x0000ffffu;
                    int jj = eold[p] - 1;
                    while (jj >= 0) { if (he[jj * S] & hm) break; --jj; }
                    end[p] = mn(jj + 2, qlen[p]);
                }
            if (dying) {
                act &= ~dying;
                if (!act) break;
                /* zero the dead lane so that it computes zeros from now on */
#pragma unroll
                for (int p = 0; p < 2; ++p)
                    if (dying & (1 << p)) {
                        const uint32_t keep = p ? 0x0000ffffu : 0xffff0000u;
                        for (int j = 0; j <= qlen[p]; ++j) he[j * S] &= keep;
                        hcol &= keep;
                    }
            }
        }
        /* lazy left trim: skip leading columns that are zero in both lanes (dead lanes are zero) */
        if ((i & 3) == 3) {
            const int lim = mn((act & 1) ? end[0] : BIG, (act & 2) ? end[1] : BIG);
            int steps = 0;
            while (steps < 8 && t < lim) { if (he[t * S]) break; ++t; ++steps; }
        }
        ++i;
    }

    PairOut *out[2] = { &oa, &ob };
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        out[p]->score = smax[p]; out[p]->qle = max_j[p] + 1; out[p]->tle = max_i[p] + 1;
        out[p]->gtle = max_ie[p] + 1; out[p]->gscore = gscore[p]; out[p]->max_off = max_off[p];
        out[p]->cells = cells[p];
    }
}

} /* namespace bswk */
