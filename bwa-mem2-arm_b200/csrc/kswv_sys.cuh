/*
 * kswv_sys.cuh -- batched mate-rescue local alignment with the DP state in REGISTERS: G lanes of a warp form a
 * systolic pipeline over the query columns of TWO pairs (kswv::getScores8/16, bwa-mem2 src/kswv.cpp; semantics =
 * bwa ksw_u8 / ksw_i16 in exact arithmetic).
 *
 * Why not the thread-owns-two-pairs mapping of the BSW kernel (rounds 1-2 shipped one, kswv_packed_kernel): its rows cost
 * 6 bytes per column and thread of shared memory, which capped a 150-column launch at 7 warps per SM (ncu: 57 % of the
 * issue slots used, three LSU instructions per cell vector, 0.40 of the DPX peak), and it kept a row-maximum scratch of
 * rows x pairs x 2 B in HBM.  ksw_u8 has no band and no trimming -- row i + 1 needs nothing from row i but the cells
 * above it -- so the rows can be skewed over the lanes:
 *
 *   lane t of the group owns the padded columns [t*C, (t+1)*C): Hc[k] = H(i-1, c), E[k] = E(i, c), Q[k] = PRMT
 *   selector of the query bases of both pairs (pair A in the low, pair B in the high 16 bits of every register, as
 *   everywhere in this library).  In step s lane t sweeps its C columns of row i = s - t; what crosses a lane
 *   boundary -- H(i, tC-1), F(i, tC), the row's running (max << 8 | 255 - column) key and the per-row score tables
 *   (or the target bases, QN) -- travels by one-lane shuffles (five to seven per step).  No shared memory for the
 *   state, no loads or stores in the cell loop, occupancy set by registers only.  ncu: ALU pipe 85 % busy.
 *
 * Exactness:
 *   - the query is RIGHT-aligned in the G*C columns: column c holds query base c - pad (pad = G*C - qlen, per pair).
 *     The pad columns come first and use a selector whose score is <= 0 for every target base (sign bytes of the row
 *     table: 0 or -1; the N code when QN), so they evaluate to H = E = F = 0 exactly like the j = -1 boundary;
 *     nothing needs a mask and the two pairs of a group may have different lengths;
 *   - lanes that have no row yet (s < t) see all-zero score tables, so their state stays zero; rows past a pair's
 *     target are fed the N base (scores <= 0: no value grows, the 16-bit halves cannot carry into each other) and are
 *     ignored by the bookkeeping, which the LAST lane of the group does in row order: row maximum and its first
 *     column from the key, (gmax, te, qe), the stop conditions (KSW_XSTOP score, 8-bit overflow), and ksw_u8's
 *     `b` list -- kept as a list ((score << 16 | row) entries in shared memory, at most rows / 2 + 1 of them) and
 *     resolved once te and score are final, as ksw_u8 does.  A pair that stops early only shortens the loop.
 * Admission (kswv_sys_eligible): scores and columns < 256 (the key is max << 8 | column in 16 bits) and at most
 * KSWV_SYS_MAX_ROWS target bases (sizes the list).  Everything else goes to the general 32-bit kernel.
 *
 * __host__ __device__: tests/emu/warp_emu.h runs the same source on an emulated warp.
 */
#pragma once
#include "bsw_core.cuh"
#include "bsw_core2.cuh"
#include "kswv_core.cuh"
#include "bsw_wide.cuh"

namespace bswk {

#define KSWV_SYS_MAX_ROWS 2046
#define KSWV_SYS_MAX_COLS 256

BSW_HD bool kswv_sys_eligible(const Const &c, int len1, int len2)
{
    return kswv_packed_eligible(c, len1, len2) && len1 <= KSWV_SYS_MAX_ROWS && c.a <= 120 && c.b <= 120 && c.ambig <= 0;
}
/* The 16-bit tier (template flag W): scores < 2048, queries of up to 512 bases -- what kswv::getScores16 is for (reads
 * longer than 255 bp, match scores > 1).  Eight shapes by query length: 8 lanes x 8 / 16 / 20 / 24 / 32 columns, then
 * 16 lanes x 20 / 26 / 32 columns.  The 8-bit entry point keeps ksw_u8's saturation rule and is not admitted. */
#define KSWV_SYSW_NCFG 8
#define KSWV_SYSW_MAX_SCORE 2047
BSW_HD int kswv_sysw_cols(int cfg)
{   /* spaced for the usual read lengths: 150 -> 160, 250 -> 256, 300 -> 320 */
    return cfg == 0 ? 64 : cfg == 1 ? 128 : cfg == 2 ? 160 : cfg == 3 ? 192 : cfg == 4 ? 256 : cfg == 5 ? 320 : cfg == 6 ? 416 : 512;
}
BSW_HD int kswv_sysw_cfg_for(int len2)
{
    for (int k = 0; k < KSWV_SYSW_NCFG; ++k) if (len2 <= kswv_sysw_cols(k)) return k;
    return -1;
}
BSW_HD bool kswv_sysw_eligible(const Const &c, int is8, int len1, int len2)
{
    if (is8 || len1 < 0 || len2 < 0 || len2 > kswv_sysw_cols(KSWV_SYSW_NCFG - 1) || len1 > KSWV_SYS_MAX_ROWS) return false;
    if (c.a > 120 || c.b > 120 || c.ambig > 0) return false;
    return (long long)mn(len1, len2) * c.a <= KSWV_SYSW_MAX_SCORE;
}
/* `b`-list entries a pair of `rows` target bases can produce, plus the open one */
BSW_HD int kswv_sys_list_cap(int rows) { return (rows >> 1) + 2; }

/*
 * One lane of a G-lane group; gl = its index.  Every lane of the warp must call this together.  blA / blB: the
 * group's `b`-list storage for pair A / B (kswv_sys_list_cap(max rows) words each; only the last lane touches it).
 * Results are valid in lane G - 1 only.
 */
template <int G, int C, bool SYM, bool QN, bool W = false>
BSW_HD void kswv_sys_lane(const Const &c, int is8, const PairIn &A, const PairIn &B,
                          const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer, int gl,
                          uint32_t *blA, uint32_t *blB, KswOut &oa, KswOut &ob)
{
    /* key of a cell: max << 8 | 255 - global column (scores and columns < 256), or -- W, the 16-bit tier -- max << 5 |
     * 31 - column within the lane (scores < 2048, any number of columns), the lane travelling beside it */
    static_assert(W ? C <= 32 : G * C <= KSWV_SYS_MAX_COLS, "column index must fit the key");
    const PairIn *in[2] = { &A, &B };
    uint32_t *bl[2] = { blA, blB };
    const int shift = c.b > -c.ambig ? c.b : -c.ambig;       /* |min(mat)| */
    int qlen[2], tlen[2], pad[2], lim[2], minsc[2], endsc[2], gmax[2], te[2], qk[2], bn[2];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const PairIn &x = *in[p];
        qlen[p] = x.valid ? x.len2 : 0;
        tlen[p] = x.valid ? x.len1 : 0;
        pad[p] = G * C - qlen[p];
        lim[p] = tlen[p] > 0 ? tlen[p] : 0;                  /* rows the bookkeeping still wants */
        minsc[p] = (x.h0 & KSW_XSUBO_) ? (x.h0 & 0xffff) : 0x10000;
        endsc[p] = (x.h0 & KSW_XSTOP_) ? (x.h0 & 0xffff) : 0x10000;
        gmax[p] = 0; te[p] = -1; qk[p] = 0; bn[p] = 0;
    }

    /* ---- query selectors of this lane's columns; state = 0 ---- */
    uint32_t Hc[C], E[C], Q[C];
#pragma unroll
    for (int k = 0; k < C; ++k) {
        const int col = gl * C + k;
        uint32_t qv[2];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const int j = col - pad[p];
            uint32_t q = 4u;                                 /* pad column */
            if (j >= 0) { q = qer[(*in[p]).idq + j]; q = q > 4u ? 4u : q; }
            qv[p] = q;
        }
        if (QN) Q[k] = bsw_selector<true>(qv[0], qv[1]);
        else {
            /* in-range: byte q of the pair's row table and its sign; pad: the sign of byte 0 twice (0 or -1) */
            const uint32_t sa = qv[0] > 3u ? 0x88u : (((8u | qv[0]) << 4) | qv[0]);
            const uint32_t sb = qv[1] > 3u ? 0xCCu : (((12u | qv[1]) << 4) | (4u | qv[1]));
            Q[k] = (sb << 8) | sa;
        }
        Hc[k] = 0u; E[k] = 0u;
    }

    const uint32_t MISW = ((uint32_t)(-c.b) & 0xffu) * 0x01010101u, AMBW = lut_hi(c), XW = ((uint32_t)c.a ^ (uint32_t)(-c.b)) & 0xffu;
    const uint32_t LUTLO = launder(lut_lo(c)), LUTHI = launder(AMBW);
    /* kept in registers: left to itself ptxas rebuilds these four from the kernel arguments in every step (nine instructions) */
    const uint32_t NOD = launder(pack2(-c.oe_del, -c.oe_del)), NOI = launder(pack2(-c.oe_ins, -c.oe_ins));
    const uint32_t NED = launder(pack2(-c.e_del, -c.e_del)), NEI = launder(pack2(-c.e_ins, -c.e_ins));
    const uint32_t KOFF = (uint32_t)(gl * C) * 0x00010001u;
    /* h * 256 + column as an IMAD (the FMA pipe idles, the ALU pipe is the bound): the multiplier must be a value ptxas
     * cannot fold, or it emits an ALU-pipe LEA */
    const uint32_t K256 = (W ? 32u : 256u) + ((uint32_t)is8 & 2u);
    const uint32_t KCOL = W ? 0x001f001fu : 0x00ff00ffu;

    /* target bases, G rows at a time: in the steps with s % G == 0 lane l of the group turns the bases of row s + l (fetched
     * one batch ahead) into that row's score tables; lane 0 picks row s's up with one indexed shuffle per table.  Rows
     * past a pair's target read as N. */
    uint32_t TA = 0u, TB = 0u, nb[2];
#pragma unroll
    for (int p = 0; p < 2; ++p) nb[p] = gl < tlen[p] ? (uint32_t)ref[(*in[p]).idr + gl] : 4u;

    /* bookkeeping state of the last lane, both pairs packed: rows left, best (max << 8 | 255 - column) key and its max, te */
    uint32_t R = pack2(lim[0], lim[1]), GM = 0u, BK = 0u, TE = 0xffffffffu;
    uint32_t BS = 0u, BR = 0xfffefffeu;                   /* open `b`-list entry: score, row (0xfffe: none yet) */
    const uint32_t ONE2 = 0x00010001u, NEG1 = 0xffffffffu;
    uint32_t NTH, NMS;
    {
        int th[2], ms[2];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            th[p] = mn(mn(endsc[p], is8 ? 255 - shift : 0x7fff), 0x7fff);     /* stop: gmax >= endsc or the 8-bit overflow */
            ms[p] = mn(minsc[p], 0x7fff);
        }
        NTH = pack2(-th[0], -th[1]); NMS = pack2(-ms[0], -ms[1]);
    }

    /* row tables of the step (QN: LA = target selector T, LB unused).  A lane without a row yet: all-zero tables / the N code */
    uint32_t LA = QN ? 0x4444u : 0u, LB = 0u;
    uint32_t hout = 0u, fout = 0u, kout = 0u, dprev = 0u, lout = 0u, BL = 0u;    /* W: lane of the running key / of the best key */
    int s = 0;
    while (w_any(gl == G - 1 && R != 0u)) {
        if ((s & (G - 1)) == 0) {
            const uint32_t ta = nb[0] > 4u ? 4u : nb[0], tb = nb[1] > 4u ? 4u : nb[1];
            if (QN) TA = ta * 0x0011u | tb * 0x1100u;
            else { TA = bsw_lutrow(MISW, XW, AMBW, ta * 8u); TB = bsw_lutrow(MISW, XW, AMBW, tb * 8u); }
#pragma unroll
            for (int p = 0; p < 2; ++p) { const int rn = s + G + gl; nb[p] = rn < tlen[p] ? (uint32_t)ref[(*in[p]).idr + rn] : 4u; }
        }
        /* ---- row s's tables (lane 0) / what the left neighbour produced in the previous step (the others) ---- */
        const uint32_t fA = w_shfl(TA, s & (G - 1), G);
        uint32_t nLA = w_shfl_up(LA, 1, G), nLB = 0u, fB = 0u;
        if (!QN) { fB = w_shfl(TB, s & (G - 1), G); nLB = w_shfl_up(LB, 1, G); }
        uint32_t hin = w_shfl_up(hout, 1, G), f = w_shfl_up(fout, 1, G), kin = w_shfl_up(kout, 1, G);
        uint32_t lin = 0u;
        if (W) lin = w_shfl_up(lout, 1, G);
        if (gl == 0) { nLA = fA; nLB = fB; hin = 0u; f = 0u; kin = 0u; lin = 0u; }
        LA = nLA; LB = nLB;

        /* ---- this lane's C cells of row s - gl (A.2) ----
         * Hc[k] = H(i-1, c) is updated in place: the first operation of cell k + 1 -- t = max(H(i-1, c) + s, E, 0), all
         * of the cell that does not hang on the f chain -- is issued before cell k overwrites Hc[k] with H(i, c). */
        uint32_t key = 0u, kprev = 0u;
        uint32_t t = p_addmax_relu(dprev, QN ? p_perm(LUTLO, LUTHI, p_xsel(Q[0], LA)) : p_perm(LA, LB, Q[0]), E[0]);
        dprev = hin;                                     /* H(i, c0 - 1): the diagonal of the next row's first cell */
#pragma unroll
        for (int k = 0; k < C; ++k) {
            uint32_t tn = 0u;
            if (k + 1 < C) {
                const uint32_t sc = QN ? p_perm(LUTLO, LUTHI, p_xsel(Q[k + 1], LA)) : p_perm(LA, LB, Q[k + 1]);
                tn = p_addmax_relu(Hc[k], sc, E[k + 1]);
            }
            const uint32_t h = p_max_relu(t, f);
            if (SYM) {
                const uint32_t u = p_add(h, NOD);
                E[k] = p_addmax_relu(E[k], NED, u);
                f = p_addmax_relu(f, NEI, u);
            } else {
                E[k] = p_addmax_relu(h, NOD, p_add(E[k], NED));
                f = p_addmax_relu(h, NOI, p_add(f, NEI));
            }
            Hc[k] = h;
            const uint32_t kc = h * K256 + (KCOL - (uint32_t)k * 0x00010001u);           /* IMAD: the FMA pipe idles here */
            if (k & 1) key = p_max3u(key, kprev, kc); else kprev = kc;
            t = tn;
        }
        if (C & 1) key = p_maxu(key, kprev);
        hout = Hc[C - 1]; fout = f;
        if (!W) kout = p_maxu(kin, key - KOFF);      /* column byte 255 - (gl*C + k): ties keep the smaller column */
        else {                                       /* mine only on a strictly larger maximum: ties keep the earlier lane */
            const uint32_t KI = kin | 0x001f001fu;
            const uint32_t WM = p_minu(p_maxu(KI, key) ^ KI, ONE2) * 0xffffu;
            kout = (kin & ~WM) | (key & WM);
            lout = (lin & ~WM) | (((uint32_t)gl * ONE2) & WM);
        }

        /* ---- bookkeeping of row s - (G-1), in the last lane (A.2), both pairs at once ---- */
        if (gl == G - 1 && s >= G - 1) {
            const int row = s - (G - 1);
            const uint32_t LM = p_minu(R, ONE2) * 0xffffu;           /* 0xffff per pair that still wants this row */
            const uint32_t IM = (W ? (kout >> 5) & 0x07ff07ffu : st_e(kout)) & LM;     /* row maxima */
            const uint32_t nGM = p_maxu(GM, IM);
            const uint32_t chg = nGM ^ GM;
            if (chg) {                                               /* imax > gmax: new (gmax, te, qe); stop conditions */
                const uint32_t CM = p_minu(chg, ONE2) * 0xffffu;
                TE = (TE & ~CM) | (((uint32_t)row * ONE2) & CM);
                BK = (BK & ~CM) | (kout & CM);
                if (W) BL = (BL & ~CM) | (lout & CM);
                GM = nGM;
                const uint32_t stopb = ~p_add(nGM, NTH) & CM & 0x80008000u;
                if (stopb & 0x00008000u) { lim[0] = row + 1; R &= 0xffff0000u; }
                if (stopb & 0x80000000u) { lim[1] = row + 1; R &= 0x0000ffffu; }
            }
            /* ksw_u8's `b` list, both pairs at once.  A row with imax >= minsc either extends the open entry (its row is
             * brow + 1: the entry keeps the larger score and, on a strictly larger one, takes this row) or opens a new one;
             * the entry it replaces goes to the list (rare: once per run of rows). */
            const uint32_t ROW2 = (uint32_t)row * ONE2;
            const uint32_t QM = LM & ~p_perm(p_add(IM, NMS), 0u, 0xBB99u);        /* 0xffff where imax >= minsc */
            const uint32_t Z = (BR + ONE2) ^ ROW2;                                 /* non-zero half: not adjacent */
            const uint32_t NEWM = p_minu(Z, ONE2) * 0xffffu & QM;
            if (NEWM) {
#pragma unroll
                for (int p = 0; p < 2; ++p)
                    if (half_u(NEWM, p) && half_u(BR, p) != 0xfffeu)
                        bl[p][bn[p]++] = ((uint32_t)half_u(BS, p) << 16) | (uint32_t)half_u(BR, p);
            }
            const uint32_t UM = p_minu((p_maxu(BS, IM) ^ BS) | Z, ONE2) * 0xffffu & QM;   /* new entry, or a strictly larger score */
            BS = (BS & ~UM) | (IM & UM);
            BR = (BR & ~UM) | (ROW2 & UM);
            R = p_addmax_relu(R, NEG1, 0u);
        }
        ++s;
    }
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        gmax[p] = half_u(GM, p); te[p] = half_s(TE, p);
        /* 255 - global column of the best key */
        qk[p] = W ? 255 - (half_u(BL, p) * C + 31 - (half_u(BK, p) & 31)) : half_u(BK, p) & 0xff;
    }

    KswOut *out[2] = { &oa, &ob };
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        KswOut &o = *out[p];
        const bool ovf = is8 && gmax[p] + shift >= 255;
        o.score = ovf ? 255 : gmax[p]; o.te = te[p]; o.qe = -1; o.score2 = -1; o.te2 = -1;
        o.cells = (long long)lim[p] * qlen[p];
        if (gl == G - 1 && !ovf) {
            /* ksw_u8: qe = smallest j attaining the maximum of row te; Hmax stays all-zero when gmax == 0 */
            o.qe = qlen[p] > 0 ? (gmax[p] > 0 ? 255 - qk[p] - pad[p] : 0) : -1;
            if (minsc[p] < 0x10000) {
                const int d = (o.score + c.maxsc - 1) / c.maxsc, low = te[p] - d, high = te[p] + d;
                if (half_u(BR, p) != 0xfffeu) bl[p][bn[p]++] = ((uint32_t)half_u(BS, p) << 16) | (uint32_t)half_u(BR, p);
                for (int k = 0; k < bn[p]; ++k) {
                    const int e = (int)(bl[p][k] & 0xffffu), sc2 = (int)(bl[p][k] >> 16);
                    if ((e < low || e > high) && sc2 > o.score2) { o.score2 = sc2; o.te2 = e; }
                }
            }
        }
    }
}

} /* namespace bswk */
