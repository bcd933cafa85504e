/*
 * kswv_core.cuh -- per-thread dynamic programming of the batched mate-rescue
 * local alignment (kswv::getScores8/16, bwa-mem2 src/kswv.cpp; semantics = bwa
 * ksw_u8 / ksw_i16 evaluated in exact arithmetic).
 *
 * Same mapping as bsw_core.cuh: one thread owns two pairs in the two 16-bit
 * lanes of each register and sweeps full rows (no band):
 *     h  = max(Hd + s, e, f, 0)              VIADD.16x2 + VIMNMX3.S16x2.RELU
 *     e' = max(h - oe_del, e - e_del, 0)     VIADD.16x2 + VIADDMNMX.S16x2.RELU
 *     f' = max(h - oe_ins, f - e_ins, 0)     VIADD.16x2 + VIADDMNMX.S16x2.RELU
 *     (imax, qe) = max over (h << 8 | 255-j) IMAD + VIMNMX.U16x2 (ties -> smallest j)
 * Per-row maxima go to a [row][thread] scratch in HBM (coalesced 2-byte
 * stores); the second-best scan of ksw_u8 (the `b` list) is replayed from it
 * once te and score are final.
 */
#pragma once
#include "bsw_core.cuh"

namespace bswk {

#define KSW_XBYTE_  0x10000
#define KSW_XSTOP_  0x20000
#define KSW_XSUBO_  0x40000
#define KSW_XSTART_ 0x80000

struct KswOut { int score, te, qe, score2, te2; long long cells; };

BSW_HD bool kswv_packed_eligible(const Const &c, int len1, int len2)
{
    if (len1 < 0 || len2 < 0 || len2 > BSW_PACKED_MAX_QLEN) return false;
    return (long long)mn(len1, len2) * c.a <= BSW_PACKED_MAX_SCORE;
}

#if defined(__CUDA_ARCH__)
__device__ __forceinline__ uint32_t p_max3_relu(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_s16x2_relu(a, b, c); }
__device__ __forceinline__ uint32_t p_max_relu(uint32_t a, uint32_t b) { return __vimax_s16x2_relu(a, b); }
#else
static inline uint32_t p_max_relu(uint32_t a, uint32_t b)
{ return pk(imax(imax(lo16(a), lo16(b)), 0), imax(imax(hi16(a), hi16(b)), 0)); }
static inline uint32_t p_max3_relu(uint32_t a, uint32_t b, uint32_t c)
{ return pk(imax(imax(imax(lo16(a), lo16(b)), lo16(c)), 0), imax(imax(imax(hi16(a), hi16(b)), hi16(c)), 0)); }
#endif

/* replay of ksw_u8's b-list rule over the stored row maxima (A.2) */
struct BList {
    int n, sc, row, low, high, score2, te2;
    BSW_HDM void init(int lo, int hi) { n = 0; sc = 0; row = 0; low = lo; high = hi; score2 = -1; te2 = -1; }
    BSW_HDM void flush() { if (n && (row < low || row > high) && sc > score2) { score2 = sc; te2 = row; } }
    BSW_HDM void push(int imax, int i)
    {
        if (n == 0 || row + 1 != i) { flush(); n = 1; sc = imax; row = i; }
        else if (sc < imax) { sc = imax; row = i; }
    }
};

struct KState {
    int qlen, tlen, idr, minsc, endsc, gmax, te, qe, rows, act;
};

/*
 * he[j*S] = row state word (st_pack: H(i-1, j) low byte, E(i, j) high byte of each lane), qs[j*S] = selector,
 * rm[i*RS] = imaxA | imaxB << 8 (this thread's slot of the row-max scratch).
 */
template <int S, bool SYM>
BSW_HD void kswv_packed_thread(const Const &c, int is8, const PairIn &A, const PairIn &B,
                               const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                               uint32_t *he, uint16_t *qs, uint16_t *rm, long long RS,
                               KswOut &oa, KswOut &ob)
{
    KState ks[2];
    const PairIn *in[2] = { &A, &B };
    const int shift = c.b > -c.ambig ? c.b : -c.ambig;       /* |min(mat)| */
    const int mxs = c.maxsc;
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const PairIn &x = *in[p];
        const int xtra = x.h0;
        ks[p].qlen = x.valid ? x.len2 : 0;
        ks[p].tlen = x.valid ? x.len1 : 0;
        ks[p].idr = x.idr;
        ks[p].minsc = (xtra & KSW_XSUBO_) ? (xtra & 0xffff) : 0x10000;
        ks[p].endsc = (xtra & KSW_XSTOP_) ? (xtra & 0xffff) : 0x10000;
        ks[p].gmax = 0; ks[p].te = -1; ks[p].qe = -1; ks[p].rows = 0;
        ks[p].act = (x.valid && x.len1 > 0) ? 1 : 0;
    }
    const int qmin = mn(ks[0].qlen, ks[1].qlen), qmax = mx(ks[0].qlen, ks[1].qlen);
    /* row state = 0; query selectors from aligned 32-bit words fetched one word ahead (see bsw_core.cuh) */
    {
        const uint32_t *wq[2]; int shq[2]; uint32_t cw0[2], cw1[2];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const PairIn &x = *in[p];
            const uintptr_t aq = reinterpret_cast<uintptr_t>(qer + (x.valid ? x.idq : 0));
            wq[p] = reinterpret_cast<const uint32_t *>(aq & ~(uintptr_t)3);
            shq[p] = (int)(aq & 3) * 8;
            cw0[p] = cw1[p] = 0u;
            if (x.valid && x.len2 > 0) { cw0[p] = wq[p][0]; cw1[p] = wq[p][1]; }
        }
        for (int j0 = 0; j0 < qmax; j0 += 4) {
            uint32_t nw[2], by[2];
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                const PairIn &x = *in[p];
                nw[p] = (x.valid && j0 + 4 < x.len2) ? wq[p][(j0 >> 2) + 2] : 0u;
                by[p] = shq[p] ? ((cw0[p] >> shq[p]) | (cw1[p] << (32 - shq[p]))) : cw0[p];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = j0 + u;
                if (j < qmax) {
                    he[j * S] = 0u;
                    uint32_t qa = (A.valid && j < A.len2) ? ((by[0] >> (8 * u)) & 0xffu) : 0u;
                    uint32_t qb = (B.valid && j < B.len2) ? ((by[1] >> (8 * u)) & 0xffu) : 0u;
                    qa = qa > 4u ? 4u : qa; qb = qb > 4u ? 4u : qb;
                    qs[j * S] = (uint16_t)(((8u | qb) << 12) | (qb << 8) | ((8u | qa) << 4) | qa);
                }
            }
            cw0[0] = cw1[0]; cw1[0] = nw[0]; cw0[1] = cw1[1]; cw1[1] = nw[1];
        }
    }
    const uint32_t LUTLO = launder(lut_lo(c)), LUTHI = launder(lut_hi(c));
    const uint32_t NEG_OE_DEL = pack2(-c.oe_del, -c.oe_del), NEG_OE_INS = pack2(-c.oe_ins, -c.oe_ins);
    const uint32_t NEG_E_DEL = pack2(-c.e_del, -c.e_del), NEG_E_INS = pack2(-c.e_ins, -c.e_ins);
    const uint32_t LONGER = ks[0].qlen > ks[1].qlen ? 0x0000ffffu : 0xffff0000u;

    /* target bases: aligned words one word (4 rows) ahead, predicated in-place refill (bsw_core.cuh) */
    uint32_t wcur[2], wnxt[2];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        wcur[p] = wnxt[p] = 0u;
        if (ks[p].act) {
            const uintptr_t a = reinterpret_cast<uintptr_t>(ref + ks[p].idr);
            const uint32_t *wp = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
            wcur[p] = wp[0]; wnxt[p] = wp[1];
        }
    }

/* one local-SW cell of both lanes (A.2) */
#define KSW_CELL(WV, Q, H, EN)                                                        \
    {                                                                                 \
        const uint32_t ve_ = st_e(WV);                                                \
        const uint32_t s_ = p_perm(LUTLO, LUTHI, p_xsel(Q, T));                       \
        H = p_max3_relu(p_add(hd, s_), ve_, f);                                       \
        if (SYM) {   /* equal gap-open+extend sums: h - oe shared by e' and f' */     \
            const uint32_t u_ = p_add(H, NEG_OE_DEL);                                 \
            EN = p_addmax_relu(ve_, NEG_E_DEL, u_);                                   \
            f = p_addmax_relu(f, NEG_E_INS, u_);                                      \
        } else {                                                                      \
            EN = p_addmax_relu(H, NEG_OE_DEL, p_add(ve_, NEG_E_DEL));                 \
            f = p_addmax_relu(H, NEG_OE_INS, p_add(f, NEG_E_INS));                    \
        }                                                                             \
        hd = st_h2(WV, ve_);                                                          \
    }

    /* one local-SW cell with unpacked inputs (trailing row of a two-row pass): HD = H(i, j-1), E = E(i+1, j) */
#define KSW_CELL_HE(TSEL, HD, E, Q, FF, H, EN)                                        \
    {                                                                                 \
        const uint32_t s_ = p_perm(LUTLO, LUTHI, p_xsel(Q, TSEL));                    \
        H = p_max3_relu(p_add(HD, s_), E, FF);                                        \
        EN = p_addmax_relu(H, NEG_OE_DEL, p_add(E, NEG_E_DEL));                       \
        FF = p_addmax_relu(H, NEG_OE_INS, p_add(FF, NEG_E_INS));                      \
    }

    /* target base selector of both lanes for row `row` (advances the word queues) */
    auto fetch_T = [&](int row) -> uint32_t {
        uint32_t tb[2];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            tb[p] = 0u;
            if (ks[p].act) {
                const uintptr_t adr = reinterpret_cast<uintptr_t>(ref + ks[p].idr) + (uintptr_t)row;
                const uint32_t t = (wcur[p] >> (8 * (int)(adr & 3))) & 0xffu;
                tb[p] = t > 4u ? 4u : t;
                rot_ld_if(wcur[p], wnxt[p], reinterpret_cast<const uint32_t *>(adr & ~(uintptr_t)3) + 2, (adr & 3) == 3);
            }
        }
        return tb[0] * 0x0011u | tb[1] * 0x1100u;
    };

    /* columns [j, qmax) of row `row` with masks (one lane finished, the tail of the longer query, or the
     * columns past the last full group), then the row's bookkeeping (A.2) */
    auto finish_row = [&](int row, uint32_t T, int j, uint32_t f, uint32_t hd, uint32_t key) {
        const uint32_t ACT = (ks[0].act ? 0x0000ffffu : 0u) | (ks[1].act ? 0xffff0000u : 0u);
        uint32_t jv = 0x00ff00ffu - (uint32_t)j * 0x00010001u;      /* 255 - j per lane */
        for (; j < qmax; ++j) {
            const uint32_t MK = ACT & (j < qmin ? 0xffffffffu : LONGER);
            const uint32_t wv = he[j * S];
            const uint32_t ve = st_e(wv);
            uint32_t q_ = qs[j * S];
            uint32_t x_ = p_xsel(q_, T);
            uint32_t s_ = p_perm(LUTLO, LUTHI, x_);
            uint32_t h_ = p_max3_relu(p_add(hd, s_), ve, f) & MK;
            uint32_t en = p_addmax_relu(h_, NEG_OE_DEL, p_add(ve, NEG_E_DEL)) & MK;
            f = p_addmax_relu(h_, NEG_OE_INS, p_add(f, NEG_E_INS)) & MK;
            hd = st_h(wv) & MK;
            he[j * S] = st_pack(h_, en);
            key = p_maxu(key, ((h_ << 8) + jv) & MK);
            jv -= 0x00010001u;
        }
        {
            const int ia = ks[0].act ? (int)((key >> 8) & 0xff) : 0;
            const int ib = ks[1].act ? (int)(key >> 24) : 0;
            rm[(long long)row * RS] = (uint16_t)(ia | (ib << 8));
        }
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            KState &s = ks[p];
            if (!s.act) continue;
            const int kp = half_u(key, p);
            const int imax = s.qlen > 0 ? (kp >> 8) : 0;
            s.rows = row + 1;
            if (imax > s.gmax) {
                s.gmax = imax; s.te = row; s.qe = 255 - (kp & 0xff);
                if ((is8 && s.gmax + shift >= 255) || s.gmax >= s.endsc) { s.act = 0; continue; }
            }
            if (row + 1 >= s.tlen) s.act = 0;
        }
    };

    const uint32_t c0 = 0x00ff00ffu, c1 = 0x00fe00feu, c2 = 0x00fd00fdu, c3 = 0x00fc00fcu;
    const int n4 = qmin & ~3;
    int i = 0;
    while (ks[0].act | ks[1].act) {
        const bool both = ks[0].act && ks[1].act;
        const uint32_t T = fetch_T(i);
        uint32_t f = 0, hd = 0, key = 0;
        int j = 0;
        if (both && n4 >= 8 && i + 1 < ks[0].tlen && i + 1 < ks[1].tlen) {
            /* ---- two rows per pass: the leading row i runs one group of four columns ahead and hands
             * H(i, .), E(i+1, .) to the trailing row i+1 in registers: per column one state load, one
             * selector load and one state store serve two rows, and two f chains are in flight.  A lane
             * that stops at row i (score threshold / 8-bit overflow) simply has its row i+1 ignored. */
            const uint32_t T2 = fetch_T(i + 1);
            const int G = n4 >> 2;
            uint32_t fT = 0, keyT = 0, h1T = 0;            /* trailing chain; h1T = H(i, 4g-1) */
            uint32_t q0 = qs[0], q1 = qs[S], q2 = qs[2 * S], q3 = qs[3 * S];
            uint32_t ha, hb, hc, hd4, ea, eb, ec, ed;
            {
                const uint32_t v0 = he[0], v1 = he[S], v2 = he[2 * S], v3 = he[3 * S];
                KSW_CELL(v0, q0, ha, ea) KSW_CELL(v1, q1, hb, eb) KSW_CELL(v2, q2, hc, ec) KSW_CELL(v3, q3, hd4, ed)
                key = p_maxu(p_max3u((ha << 8) + c0, (hb << 8) + c1, (hc << 8) + c2), (hd4 << 8) + c3);
            }
            uint32_t w0 = he[4 * S], w1 = he[5 * S], w2 = he[6 * S], w3 = he[7 * S];
            uint32_t r0 = qs[4 * S], r1 = qs[5 * S], r2 = qs[6 * S], r3 = qs[7 * S];
            for (int g = 0; g + 1 < G; ++g) {
                const int kl = 4 * g + 4, kt = 4 * g;
                uint32_t x0 = 0, x1 = 0, x2 = 0, x3 = 0, y0 = 0, y1 = 0, y2 = 0, y3 = 0;
                if (g + 2 < G) {                           /* loads of leading group g+2 */
                    x0 = he[(kl + 4) * S]; x1 = he[(kl + 5) * S]; x2 = he[(kl + 6) * S]; x3 = he[(kl + 7) * S];
                    y0 = qs[(kl + 4) * S]; y1 = qs[(kl + 5) * S]; y2 = qs[(kl + 6) * S]; y3 = qs[(kl + 7) * S];
                }
                uint32_t na, nb, nc, nd, fa, fb, fc, fd;
                KSW_CELL(w0, r0, na, fa) KSW_CELL(w1, r1, nb, fb) KSW_CELL(w2, r2, nc, fc) KSW_CELL(w3, r3, nd, fd)
                key = p_maxu(key, p_maxu(p_max3u((na << 8) + c0, (nb << 8) + c1, (nc << 8) + c2), (nd << 8) + c3) - (uint32_t)kl * 0x00010001u);
                uint32_t ta, tb_, tc, td, ua, ub, uc, ud;
                KSW_CELL_HE(T2, h1T, ea, q0, fT, ta, ua) he[(kt + 0) * S] = st_pack(ta, ua);
                KSW_CELL_HE(T2, ha, eb, q1, fT, tb_, ub) he[(kt + 1) * S] = st_pack(tb_, ub);
                KSW_CELL_HE(T2, hb, ec, q2, fT, tc, uc)  he[(kt + 2) * S] = st_pack(tc, uc);
                KSW_CELL_HE(T2, hc, ed, q3, fT, td, ud)  he[(kt + 3) * S] = st_pack(td, ud);
                keyT = p_maxu(keyT, p_maxu(p_max3u((ta << 8) + c0, (tb_ << 8) + c1, (tc << 8) + c2), (td << 8) + c3) - (uint32_t)kt * 0x00010001u);
                h1T = hd4; ha = na; hb = nb; hc = nc; hd4 = nd; ea = fa; eb = fb; ec = fc; ed = fd;
                q0 = r0; q1 = r1; q2 = r2; q3 = r3;
                w0 = x0; w1 = x1; w2 = x2; w3 = x3; r0 = y0; r1 = y1; r2 = y2; r3 = y3;
            }
            {   /* trailing group G-1 */
                const int kt = n4 - 4;
                uint32_t ta, tb_, tc, td, ua, ub, uc, ud;
                KSW_CELL_HE(T2, h1T, ea, q0, fT, ta, ua) he[(kt + 0) * S] = st_pack(ta, ua);
                KSW_CELL_HE(T2, ha, eb, q1, fT, tb_, ub) he[(kt + 1) * S] = st_pack(tb_, ub);
                KSW_CELL_HE(T2, hb, ec, q2, fT, tc, uc)  he[(kt + 2) * S] = st_pack(tc, uc);
                KSW_CELL_HE(T2, hc, ed, q3, fT, td, ud)  he[(kt + 3) * S] = st_pack(td, ud);
                keyT = p_maxu(keyT, p_maxu(p_max3u((ta << 8) + c0, (tb_ << 8) + c1, (tc << 8) + c2), (td << 8) + c3) - (uint32_t)kt * 0x00010001u);
            }
            /* columns past the last full group: row i first (its outputs feed row i+1), then row i+1 */
            finish_row(i, T, n4, f, hd, key);
            finish_row(i + 1, T2, n4, fT, hd4, keyT);
            i += 2;
            continue;
        }
        if (both && qmin >= 4) {
            /* both lanes live: unmasked groups of four columns, next group's loads issued first.
             * key = max over (h << 8 | 255 - j): row maximum and its FIRST column. */
            uint32_t v0 = he[0], v1 = he[S], v2 = he[2 * S], v3 = he[3 * S];
            uint32_t q0 = qs[0], q1 = qs[S], q2 = qs[2 * S], q3 = qs[3 * S];
#define KSW_GROUP(V0, V1, V2, V3, Q0, Q1, Q2, Q3)                                     \
            {                                                                         \
                uint32_t ha, hb, hc, hd4, ea, eb, ec, ed;                             \
                KSW_CELL(V0, Q0, ha, ea) he[(j + 0) * S] = st_pack(ha, ea);           \
                KSW_CELL(V1, Q1, hb, eb) he[(j + 1) * S] = st_pack(hb, eb);           \
                KSW_CELL(V2, Q2, hc, ec) he[(j + 2) * S] = st_pack(hc, ec);           \
                KSW_CELL(V3, Q3, hd4, ed) he[(j + 3) * S] = st_pack(hd4, ed);         \
                const uint32_t g = p_maxu(p_max3u((ha << 8) + c0, (hb << 8) + c1, (hc << 8) + c2), (hd4 << 8) + c3); \
                key = p_maxu(key, g - (uint32_t)j * 0x00010001u);                     \
            }
            while (j + 8 <= n4) {
                const uint32_t w0 = he[(j + 4) * S], w1 = he[(j + 5) * S], w2 = he[(j + 6) * S], w3 = he[(j + 7) * S];
                const uint32_t r0 = qs[(j + 4) * S], r1 = qs[(j + 5) * S], r2 = qs[(j + 6) * S], r3 = qs[(j + 7) * S];
                KSW_GROUP(v0, v1, v2, v3, q0, q1, q2, q3)
                v0 = w0; v1 = w1; v2 = w2; v3 = w3; q0 = r0; q1 = r1; q2 = r2; q3 = r3;
                j += 4;
            }
            KSW_GROUP(v0, v1, v2, v3, q0, q1, q2, q3)
            j += 4;
#undef KSW_GROUP
        }
        finish_row(i, T, j, f, hd, key);
        ++i;
    }

    KswOut *out[2] = { &oa, &ob };
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        KState &s = ks[p];
        KswOut &o = *out[p];
        const bool ovf = is8 && s.gmax + shift >= 255;
        o.score = ovf ? 255 : s.gmax; o.te = s.te; o.qe = -1; o.score2 = -1; o.te2 = -1;
        o.cells = (long long)s.rows * s.qlen;
        if (!ovf) {
            /* ksw_u8: qe = smallest j attaining max(Hmax); Hmax stays all-zero when gmax == 0 */
            o.qe = s.qlen > 0 ? (s.gmax > 0 ? s.qe : 0) : -1;
            if (s.minsc < 0x10000) {
                BList bl; const int d = (o.score + mxs - 1) / mxs;
                bl.init(s.te - d, s.te + d);
                for (int r = 0; r < s.rows; ++r) {
                    const int v = rm[(long long)r * RS];
                    const int im = p ? (v >> 8) : (v & 0xff);
                    if (im >= s.minsc) bl.push(im, r);
                }
                bl.flush();
                o.score2 = bl.score2; o.te2 = bl.te2;
            }
        }
    }
}

/* general 32-bit path: one thread, one pair; H/E rows and row maxima in HBM */
template <typename IDX>
BSW_HD void kswv_general_thread(const Const &c, int is8, const PairIn &P,
                                const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                                int2 *he, IDX S, int *rm, IDX RS, KswOut &o)
{
    const int qlen = P.len2, tlen = P.len1, xtra = P.h0;
    const uint8_t *query = qer + P.idq, *target = ref + P.idr;
    const int shift = c.b > -c.ambig ? c.b : -c.ambig, mxs = c.maxsc;
    const int minsc = (xtra & KSW_XSUBO_) ? (xtra & 0xffff) : 0x10000;
    const int endsc = (xtra & KSW_XSTOP_) ? (xtra & 0xffff) : 0x10000;
    int gmax = 0, te = -1, qe = -1, rows = 0;
    for (int j = 0; j < qlen; ++j) he[(IDX)j * S] = make_int2(0, 0);
    for (int i = 0; i < tlen; ++i) {
        int f = 0, hd = 0, imax = 0, jmax = -1;
        int t = target[i]; t = t > 4 ? 4 : t;
        for (int j = 0; j < qlen; ++j) {
            int2 v = he[(IDX)j * S];
            int q = query[j]; q = q > 4 ? 4 : q;
            int sc = (q == 4 || t == 4) ? c.ambig : (q == t ? c.a : -c.b);
            int h = hd + sc, e = v.y, tt;
            hd = v.x;
            h = h > 0 ? h : 0; h = h > e ? h : e; h = h > f ? h : f;
            if (h > imax) { imax = h; jmax = j; }
            e -= c.e_del; tt = h - c.oe_del; e = e > tt ? e : tt; e = e > 0 ? e : 0;
            f -= c.e_ins; tt = h - c.oe_ins; f = f > tt ? f : tt; f = f > 0 ? f : 0;
            he[(IDX)j * S] = make_int2(h, e);
        }
        rm[(IDX)i * RS] = imax;
        rows = i + 1;
        if (imax > gmax) {
            gmax = imax; te = i; qe = jmax;
            if ((is8 && gmax + shift >= 255) || gmax >= endsc) break;
        }
    }
    const bool ovf = is8 && gmax + shift >= 255;
    o.score = ovf ? 255 : gmax; o.te = te; o.qe = -1; o.score2 = -1; o.te2 = -1;
    o.cells = (long long)rows * qlen;
    if (!ovf) {
        o.qe = qlen > 0 ? (gmax > 0 ? qe : 0) : -1;
        if (minsc < 0x10000) {
            BList bl; const int d = (o.score + mxs - 1) / mxs;
            bl.init(te - d, te + d);
            for (int r = 0; r < rows; ++r) { const int im = rm[(IDX)r * RS]; if (im >= minsc) bl.push(im, r); }
            bl.flush();
            o.score2 = bl.score2; o.te2 = bl.te2;
        }
    }
}

} /* namespace bswk */
