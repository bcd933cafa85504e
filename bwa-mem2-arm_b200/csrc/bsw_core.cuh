/*
 * bsw_core.cuh -- per-thread dynamic programming of the banded extension
 * (BandedPairWiseSW::getScores8/16, bwa-mem2 src/bandedSWA.cpp; semantics =
 * bwa ksw_extend2).
 *
 * Mapping (DESIGN.md "BSW kernel"): inter-pair SIMD-in-word.  One thread owns
 * TWO sequence pairs, lane A in the low and lane B in the high 16 bits of every
 * 32-bit register, and sweeps them row-synchronously, exactly like one
 * 2-lane slice of the reference's inter-sequence vector kernels.  A warp is a
 * 64-pair vector.  All cell arithmetic is packed signed 16-bit DPX:
 *     M  = min(Hd + s, (a+1)*Hd)             IMAD + VIADDMNMX.S16x2 (zero gate of A.1)
 *     h  = max3(M, e, f)                     VIMNMX3.S16x2
 *     e' = max(M - oe_del, e - e_del, 0)     VIADD.16x2 + VIADDMNMX.S16x2.RELU
 *     f' = max(M - oe_ins, f - e_ins, 0)     VIADD.16x2 + VIADDMNMX.S16x2.RELU
 *     (m, mj) = max over (h << 8 | j)        IMAD + VIMNMX.U16x2  (ties -> last j)
 * The substitution score of both lanes comes from one LOP3 + one PRMT used as
 * an 8-entry byte LUT with sign replication.
 * The DP row (H shifted by one column, E) lives in shared memory as one 32-bit
 * word per [column][thread] (st_pack), so every access is bank-conflict free no
 * matter which column a thread is at; rows are never re-read from HBM.
 *
 * The functions are __host__ __device__: tests/emu builds them with g++ (DPX
 * emulated in plain C) to check the kernel logic against the oracle without a
 * GPU.  The product library only ever runs the __device__ instantiation.
 */
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define BSW_HD __host__ __device__ __forceinline__
#define BSW_HDM __host__ __device__ __forceinline__
#else
#define BSW_HD static inline
#define BSW_HDM inline
struct uint2 { uint32_t x, y; };
static inline uint2 make_uint2(uint32_t x, uint32_t y) { uint2 r; r.x = x; r.y = y; return r; }
struct int2 { int32_t x, y; };
static inline int2 make_int2(int32_t x, int32_t y) { int2 r; r.x = x; r.y = y; return r; }
#endif

namespace bswk {

struct Const {
    int o_del, e_del, o_ins, e_ins, oe_del, oe_ins;
    int zdrop, end_bonus;
    int a;        /* match score  (> 0)              */
    int b;        /* mismatch penalty (> 0)          */
    int ambig;    /* score against N (<= 0, ref: -1) */
    int maxsc;    /* max entry of the scoring matrix */
};

struct PairIn  { int idr, idq, len1, len2, h0, valid; };
struct PairOut { int score, tle, gtle, qle, gscore, max_off; long long cells; };

/* ---------------------------------------------------------------- packed ops */
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ uint32_t p_addmin(uint32_t a, uint32_t b, uint32_t c) { return __viaddmin_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t p_max3(uint32_t a, uint32_t b, uint32_t c)   { return __vimax3_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t p_addmax_relu(uint32_t a, uint32_t b, uint32_t c) { return __viaddmax_s16x2_relu(a, b, c); }
__device__ __forceinline__ uint32_t p_add(uint32_t a, uint32_t b)                { return __vadd2(a, b); }
__device__ __forceinline__ uint32_t p_maxu(uint32_t a, uint32_t b)               { return __vmaxu2(a, b); }
/* raw PRMT (default mode): selector nibble bit 3 replicates the sign of the selected byte.
 * NOT __byte_perm, which masks the selector with 0x7777 and so drops the sign replication. */
__device__ __forceinline__ uint32_t p_perm(uint32_t a, uint32_t b, uint32_t s)
{ uint32_t r; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(s)); return r; }
/* LUT index of both lanes in ONE LOP3: (q ^ t) | ((q | t) & 0x4444), truth table 0xBC */
__device__ __forceinline__ uint32_t p_xsel(uint32_t q, uint32_t t)
{ uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0xBC;" : "=r"(r) : "r"(q), "r"(t), "r"(0x4444u)); return r; }
#else
static inline int16_t lo16(uint32_t v) { return (int16_t)(v & 0xffff); }
static inline int16_t hi16(uint32_t v) { return (int16_t)(v >> 16); }
static inline uint32_t pk(int lo, int hi) { return ((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16); }
static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline uint32_t p_addmin(uint32_t a, uint32_t b, uint32_t c)
{ return pk(imin((int16_t)(lo16(a) + lo16(b)), lo16(c)), imin((int16_t)(hi16(a) + hi16(b)), hi16(c))); }
static inline uint32_t p_max3(uint32_t a, uint32_t b, uint32_t c)
{ return pk(imax(imax(lo16(a), lo16(b)), lo16(c)), imax(imax(hi16(a), hi16(b)), hi16(c))); }
static inline uint32_t p_addmax_relu(uint32_t a, uint32_t b, uint32_t c)
{ return pk(imax(imax((int16_t)(lo16(a) + lo16(b)), lo16(c)), 0), imax(imax((int16_t)(hi16(a) + hi16(b)), hi16(c)), 0)); }
static inline uint32_t p_add(uint32_t a, uint32_t b) { return pk(lo16(a) + lo16(b), hi16(a) + hi16(b)); }
static inline uint32_t p_maxu(uint32_t a, uint32_t b)
{ uint32_t l = (a & 0xffff) > (b & 0xffff) ? (a & 0xffff) : (b & 0xffff), h = (a >> 16) > (b >> 16) ? (a >> 16) : (b >> 16); return l | (h << 16); }
static inline uint32_t p_xsel(uint32_t q, uint32_t t) { return (q ^ t) | ((q | t) & 0x4444u); }
static inline uint32_t p_perm(uint32_t a, uint32_t b, uint32_t s)
{
    uint64_t v = ((uint64_t)b << 32) | a; uint32_t r = 0;
    for (int k = 0; k < 4; ++k) {
        uint32_t n = (s >> (4 * k)) & 0xf;
        uint32_t byte = (uint32_t)(v >> (8 * (n & 7))) & 0xff;
        if (n & 8) byte = (byte & 0x80) ? 0xff : 0x00;
        r |= byte << (8 * k);
    }
    return r;
}
#endif

BSW_HD uint32_t pack2(int lo, int hi) { return ((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16); }
BSW_HD int half_s(uint32_t v, int p) { return p ? (int)(int16_t)(v >> 16) : (int)(int16_t)(v & 0xffff); }
BSW_HD int half_u(uint32_t v, int p) { return p ? (int)(v >> 16) : (int)(v & 0xffff); }
BSW_HD int mn(int a, int b) { return a < b ? a : b; }
BSW_HD int mx(int a, int b) { return a > b ? a : b; }

/* limits of the packed kernel: every cell value h <= h0 + min(len1,len2)*a must
 * fit 8 bits and every column index must fit 8 bits, so that (h << 8 | j) fits
 * one unsigned 16-bit lane. */
#define BSW_PACKED_MAX_SCORE 255
#define BSW_PACKED_MAX_QLEN  255
/* the BSW sweep runs up to 3 columns past the query end (bsw_core2.cuh): column indices must still fit 8 bits */
#define BSW_PACKED_MAX_QLEN_BSW 252
#define BSW_SWEEP_SLACK 4

BSW_HD bool packed_eligible(const Const &c, int len1, int len2, int h0)
{
    if (len1 < 0 || len2 < 0 || h0 <= 0) return false;
    if (len2 > BSW_PACKED_MAX_QLEN_BSW) return false;
    long long bound = (long long)h0 + (long long)mn(len1, len2) * c.a;
    if (c.a > 120 || c.b > 120) return false;
    return bound <= BSW_PACKED_MAX_SCORE;
}

/* effective band of one pair (A.1: w = min(w, max_ins, max_del)) */
BSW_HD int eff_band(const Const &c, int qlen, int w)
{
    int max_ins = (int)((double)(qlen * c.maxsc + c.end_bonus - c.o_ins) / c.e_ins + 1.);
    max_ins = max_ins > 1 ? max_ins : 1;
    w = w < max_ins ? w : max_ins;
    int max_del = (int)((double)(qlen * c.maxsc + c.end_bonus - c.o_del) / c.e_del + 1.);
    max_del = max_del > 1 ? max_del : 1;
    w = w < max_del ? w : max_del;
    return w;
}

/* score LUT for PRMT: index = (q ^ t) | ((q | t) & 4):  0 match, 1..3 mismatch, 4..7 ambiguous */
BSW_HD uint32_t lut_lo(const Const &c)
{ uint32_t m = (uint32_t)c.a & 0xff, x = (uint32_t)(-c.b) & 0xff; return m | (x << 8) | (x << 16) | (x << 24); }
BSW_HD uint32_t lut_hi(const Const &c)
{ uint32_t n = (uint32_t)c.ambig & 0xff; return n * 0x01010101u; }

/* per-pair scalar state of the row loop */
struct PState {
    int qlen, tlen, h0, w, idr;
    int end, max, max_i, max_j, max_ie, gscore, max_off;
    int act, clean;
    int beg;             /* true (trimmed) beg: only maintained when COUNT */
    long long cells;
};

struct RowCtx { uint32_t T, LUTLO, LUTHI, GATE, NOD, NOI, NED, NEI; };

#if defined(__CUDA_ARCH__)
/* predicated in-place global load: w = *p iff c (the register keeps its value otherwise) */
__device__ __forceinline__ void rot_ld_if(uint32_t &cur, uint32_t &nxt, const uint32_t *p, bool c)
{   /* if (c) { cur = nxt; nxt = *p; } with both steps predicated: nxt is only READ on rows that rotate */
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %3, 0; @q mov.b32 %0, %1; @q ld.global.nc.u32 %1, [%2]; }"
                 : "+r"(cur), "+r"(nxt) : "l"(p), "r"((uint32_t)c));
}
__device__ __forceinline__ uint32_t p_max3u(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_u16x2(a, b, c); }
/* keeps a warp-uniform constant in a vector register (PRMT cannot take a uniform register as data) */
__device__ __forceinline__ uint32_t launder(uint32_t v) { uint32_t r; asm volatile("mov.b32 %0, %1;" : "=r"(r) : "r"(v)); return r; }
#else
static inline uint32_t p_max3u(uint32_t a, uint32_t b, uint32_t c) { return p_maxu(p_maxu(a, b), c); }
static inline uint32_t launder(uint32_t v) { return v; }
static inline void rot_ld_if(uint32_t &cur, uint32_t &nxt, const uint32_t *p, bool c) { if (c) { cur = nxt; nxt = *p; } }
#endif

/* one DP cell of both lanes (A.1 inner loop body) */
/* DP row state of one column, both lanes, in ONE 32-bit word: lane A = bits 0..15, lane B = bits
 * 16..31, each lane = H(i-1, j-1) in its low byte and E(i, j) in its high byte.  The packed kernel
 * only admits pairs whose values fit 8 bits (packed_eligible), so nothing is lost; the row costs
 * 4 B + a 2 B selector per column per thread of shared memory instead of 8 B + 2 B, which is what
 * sets the number of resident warps per SM (DESIGN.md "occupancy"). */
#if defined(__CUDA_ARCH__)
/* E of both lanes in one PRMT (ALU pipe); H and the re-pack as IMADs (FMA pipe, which idles here) */
__device__ __forceinline__ uint32_t st_e(uint32_t w) { return p_perm(w, 0u, 0x4341u); }
#else
static inline uint32_t st_e(uint32_t w) { return (w >> 8) & 0x00ff00ffu; }
#endif
BSW_HD uint32_t st_h2(uint32_t w, uint32_t e) { return w - e * 256u; }
BSW_HD uint32_t st_h(uint32_t w) { return w & 0x00ff00ffu; }
BSW_HD uint32_t st_pack(uint32_t h, uint32_t e) { return e * 256u + h; }

/* SYM: o_del + e_del == o_ins + e_ins (bwa-mem's default): M - oe is computed once for both gap
 * directions, e' = max(e - e_del, M - oe, 0), f' = max(f - e_ins, M - oe, 0): one packed add less. */
template <bool SYM>
BSW_HD void bsw_cell(const RowCtx &r, uint32_t w, uint32_t q, uint32_t &f, uint32_t &h, uint32_t &en)
{
    const uint32_t e = st_e(w), hs = st_h2(w, e);
    const uint32_t x = p_xsel(q, r.T);
    const uint32_t s = p_perm(r.LUTLO, r.LUTHI, x);
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

/* up to four cells with a bound and the f-cut check: used for the group that holds the cut column
 * and for the final partial group of a sweep (everything else runs the unchecked BSW_GROUP) */
template <int S, bool SYM>
BSW_HD void bsw_group_generic(const RowCtx &r, uint32_t *hp, int k, int cnt, int kcut, uint32_t keepf, uint32_t keeph,
                              uint32_t v0, uint32_t v1, uint32_t v2, uint32_t v3,
                              uint32_t q0, uint32_t q1, uint32_t q2, uint32_t q3, uint32_t cbase,
                              uint32_t &f, uint32_t &h1, uint32_t &key, uint32_t &h1save)
{
    const uint32_t vv[4] = { v0, v1, v2, v3 }, qq[4] = { q0, q1, q2, q3 };
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        if (c < cnt) {
            if (k + c == kcut) { h1save = h1; f &= keepf; h1 &= keeph; }
            uint32_t h, en;
            bsw_cell<SYM>(r, vv[c], qq[c], f, h, en);
            hp[(k + c) * S] = st_pack(h1, en);
            h1 = h;
            key = p_maxu(key, (h << 8) + cbase + (uint32_t)(k + c) * 0x00010001u);
        }
    }
}

/*
 * Unmasked sweep of columns [j0, j1) for both lanes, software-pipelined in groups of four
 * columns: the next group's state words and selectors are loaded before the current group's
 * dependent chain (f, h1, key) is evaluated.
 * key = max over (h << 8 | j): row maximum and its LAST column (ties -> larger j).
 * jcut (j0 <= jcut < j1, or < 0 for none): before column jcut the running h1 is saved to h1save,
 * f &= keepf and h1 &= keeph -- the exact right-trim cut of the lane that ends first, after which
 * that lane sweeps on over zero-state entries.
 */
template <int S, bool SYM>
BSW_HD void bsw_sweep_cut(const RowCtx &r, uint32_t *he, const uint16_t *qs, int j0, int j1,
                          int jcut, uint32_t keepf, uint32_t keeph,
                          uint32_t &f, uint32_t &h1, uint32_t &key, uint32_t &h1save)
{
    const int n = j1 - j0;
    if (n <= 0) return;
    uint32_t *hp = he + j0 * S;
    const uint16_t *qp = qs + j0 * S;
    const uint32_t c0 = (uint32_t)j0 * 0x00010001u, c1 = c0 + 0x00010001u, c2 = c0 + 0x00020002u, c3 = c0 + 0x00030003u;
    const int kcut = jcut - j0;                       /* relative cut column (negative: none) */
    const int n4 = n & ~3;
    int k = 0;
    int stop = kcut >= 0 ? (kcut & ~3) : n4;          /* unchecked groups run up to the group holding the cut */
    for (;;) {
        if (k < stop) {
            uint32_t v0 = hp[k * S], v1 = hp[(k + 1) * S], v2 = hp[(k + 2) * S], v3 = hp[(k + 3) * S];
            uint32_t q0 = qp[k * S], q1 = qp[(k + 1) * S], q2 = qp[(k + 2) * S], q3 = qp[(k + 3) * S];
#define BSW_GROUP(V0, V1, V2, V3, Q0, Q1, Q2, Q3)                                              \
            {                                                                                  \
                uint32_t ha, hb, hc, hd, ea, eb, ec, ed;                                       \
                bsw_cell<SYM>(r, V0, Q0, f, ha, ea); hp[(k + 0) * S] = st_pack(h1, ea);             \
                bsw_cell<SYM>(r, V1, Q1, f, hb, eb); hp[(k + 1) * S] = st_pack(ha, eb);             \
                bsw_cell<SYM>(r, V2, Q2, f, hc, ec); hp[(k + 2) * S] = st_pack(hb, ec);             \
                bsw_cell<SYM>(r, V3, Q3, f, hd, ed); hp[(k + 3) * S] = st_pack(hc, ed);             \
                h1 = hd;                                                                       \
                const uint32_t g = p_maxu(p_max3u((ha << 8) + c0, (hb << 8) + c1, (hc << 8) + c2), (hd << 8) + c3); \
                key = p_maxu(key, g + (uint32_t)k * 0x00010001u);                              \
            }
            while (k + 8 <= stop) {
                const uint32_t w0 = hp[(k + 4) * S], w1 = hp[(k + 5) * S], w2 = hp[(k + 6) * S], w3 = hp[(k + 7) * S];
                const uint32_t r0 = qp[(k + 4) * S], r1 = qp[(k + 5) * S], r2 = qp[(k + 6) * S], r3 = qp[(k + 7) * S];
                BSW_GROUP(v0, v1, v2, v3, q0, q1, q2, q3)
                v0 = w0; v1 = w1; v2 = w2; v3 = w3; q0 = r0; q1 = r1; q2 = r2; q3 = r3;
                k += 4;
            }
            BSW_GROUP(v0, v1, v2, v3, q0, q1, q2, q3)
            k += 4;
#undef BSW_GROUP
        }
        if (k >= n) break;
        {   /* the group holding the cut column, or the final partial group */
            const int cnt = n - k < 4 ? n - k : 4;
            const uint32_t v0 = hp[k * S], q0 = qp[k * S];
            const uint32_t v1 = cnt > 1 ? hp[(k + 1) * S] : 0u, q1 = cnt > 1 ? qp[(k + 1) * S] : 0u;
            const uint32_t v2 = cnt > 2 ? hp[(k + 2) * S] : 0u, q2 = cnt > 2 ? qp[(k + 2) * S] : 0u;
            const uint32_t v3 = cnt > 3 ? hp[(k + 3) * S] : 0u, q3 = cnt > 3 ? qp[(k + 3) * S] : 0u;
            bsw_group_generic<S, SYM>(r, hp, k, cnt, kcut, keepf, keeph, v0, v1, v2, v3, q0, q1, q2, q3, c0, f, h1, key, h1save);
            k += cnt;
        }
        if (k >= n) break;
        stop = n4;
    }
}

/*
 * One thread, two pairs.  he[j*S] = row state word of column j (see st_pack),
 * qs[j*S] = PRMT selector of query column j for both lanes.  S = element stride
 * (32 in the kernel: [column][lane]).
 *
 * Exactness notes (all relative to A.1):
 *  - left trimming is value-neutral, so the sweep starts at a lazily advanced common start
 *    `t` (every entry left of t is zero in both lanes) instead of each lane's exact `beg`;
 *  - right trimming is value-affecting: a lane's f chain is cut at its exact `end`; beyond it
 *    the lane keeps sweeping over zero-state entries (writes zeros over zeros) while the
 *    longer lane finishes, which is exact as long as everything right of `end` is zero
 *    ("clean" lane: no initial-row value survives right of row 0's range).  Lanes that are
 *    not clean, or whose band limit is crossed, take the masked path;
 *  - a lane that has terminated is zeroed once, after which it computes zeros for free.
 * COUNT additionally tracks the exact `beg` to report the oracle's DP-cell count.
 */
template <int S, bool COUNT, bool SYM>
BSW_HD void bsw_packed_thread(const Const &c, int w_in, const PairIn &A, const PairIn &B,
                              const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                              uint32_t *he, uint16_t *qs, PairOut &oa, PairOut &ob)
{
    PState ps[2];
    const PairIn *in[2] = { &A, &B };
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const PairIn &x = *in[p];
        ps[p].qlen = x.valid ? x.len2 : 0;
        ps[p].tlen = x.valid ? x.len1 : 0;
        ps[p].h0 = x.valid ? x.h0 : 0;
        ps[p].idr = x.idr;
        ps[p].w = eff_band(c, ps[p].qlen, w_in);
        ps[p].beg = 0; ps[p].end = ps[p].qlen;
        ps[p].max = ps[p].h0; ps[p].max_i = -1; ps[p].max_j = -1; ps[p].max_ie = -1;
        ps[p].gscore = -1; ps[p].max_off = 0; ps[p].cells = 0;
        ps[p].act = (x.valid && x.len1 > 0) ? 1 : 0;
        ps[p].clean = 1;
    }

    /* ---- first row of H (A.1), E = 0, query selectors ----
     * The query bases are read as aligned 32-bit words, one word (4 columns) ahead of their use
     * (funnel-shifted to the pair's byte offset): a byte load per column stalls on every column
     * (10.8 % of the kernel time in profile r1g).  Reads stay within 12 bytes of the last base. */
    {
        const int qmax = mx(ps[0].qlen, ps[1].qlen);
        int hp0 = 0, hp1 = 0, lastnz0 = 0, lastnz1 = 0;
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
        for (int j0 = 0; j0 <= qmax; j0 += 4) {
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
                if (j <= qmax) {
                    int hv[2];
#pragma unroll
                    for (int p = 0; p < 2; ++p) {
                        int prev = p ? hp1 : hp0, v;
                        if (j > ps[p].qlen) v = 0;
                        else if (j == 0) v = ps[p].h0;
                        else if (j == 1) v = ps[p].h0 > c.oe_ins ? ps[p].h0 - c.oe_ins : 0;
                        else v = prev > c.e_ins ? prev - c.e_ins : 0;
                        hv[p] = v;
                    }
                    hp0 = hv[0]; hp1 = hv[1];
                    if (hv[0]) lastnz0 = j;
                    if (hv[1]) lastnz1 = j;
                    he[j * S] = pack2(hv[0], hv[1]);
                    uint32_t qa = (A.valid && j < A.len2) ? ((by[0] >> (8 * u)) & 0xffu) : 0u;
                    uint32_t qb = (B.valid && j < B.len2) ? ((by[1] >> (8 * u)) & 0xffu) : 0u;
                    qa = qa > 4u ? 4u : qa; qb = qb > 4u ? 4u : qb;
                    qs[j * S] = (uint16_t)(((8u | qb) << 12) | (qb << 8) | ((8u | qa) << 4) | qa);
                }
            }
            cw0[0] = cw1[0]; cw1[0] = nw[0]; cw0[1] = cw1[1]; cw1[1] = nw[1];
        }
        /* clean: every non-zero initial entry is inside row 0's range [0, min(qlen, w+1)] */
        ps[0].clean = lastnz0 <= mn(ps[0].qlen, ps[0].w + 1);
        ps[1].clean = lastnz1 <= mn(ps[1].qlen, ps[1].w + 1);
    }

    RowCtx r;
    r.LUTLO = launder(lut_lo(c)); r.LUTHI = launder(lut_hi(c));
    r.NOD = pack2(-c.oe_del, -c.oe_del); r.NOI = pack2(-c.oe_ins, -c.oe_ins);
    r.NED = pack2(-c.e_del, -c.e_del);   r.NEI = pack2(-c.e_ins, -c.e_ins);
    /* zero gate of A.1 (M = Hd ? Hd + s : 0) as M = min(Hd + s, (a+1)*Hd): exact because
     * s <= a <= a*Hd for Hd >= 1, and (a+1)*0 = 0 caps a dead cell at 0.  Lanes hold <= 255. */
    r.GATE = (uint32_t)(c.a + 1);
    const int BIG = 0x3fffffff;

    /* target bases: aligned 32-bit words of the reference buffer, fetched one word (4 rows) ahead;
     * reads stay within 8 bytes of the pair's last base (the boundary layer pads the buffers).
     * The refill is a predicated IN-PLACE rotate + load (rot_ld_if): written as `if (c) w = *p` the compiler loads
     * into a temporary and moves it, and the move waits for the load (13 % of kernel time, profile r1f). */
    uint32_t wcur[2], wnxt[2];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        wcur[p] = wnxt[p] = 0u;
        if (ps[p].act) {
            const uintptr_t a = reinterpret_cast<uintptr_t>(ref + ps[p].idr);
            const uint32_t *wp = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
            wcur[p] = wp[0]; wnxt[p] = wp[1];
        }
    }

    int t = 0;          /* lazy common start: all entries < t are zero in both lanes */
    int i = 0;
    while (ps[0].act | ps[1].act) {
        int sp[2], ep[2], h1i[2], lb[2];
        uint32_t tb[2];
        int died = 0;                              /* lanes that terminated at row start (empty range) */
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            PState &s = ps[p];
            sp[p] = BIG; ep[p] = 0; h1i[p] = 0; tb[p] = 0; lb[p] = 0;
            if (!s.act) continue;
            lb[p] = i - s.w > 0 ? i - s.w : 0;
            if (s.end > i + s.w + 1) s.end = i + s.w + 1;
            if (s.end > s.qlen) s.end = s.qlen;
            if (COUNT) { if (s.beg < lb[p]) s.beg = lb[p]; if (s.end > s.beg) s.cells += s.end - s.beg; }
            const int st = t > lb[p] ? t : lb[p];
            int h1 = 0;
            if (st == 0) { h1 = s.h0 - (c.o_del + c.e_del * (i + 1)); if (h1 < 0) h1 = 0; }
            {
                const uintptr_t adr = reinterpret_cast<uintptr_t>(ref + s.idr) + (uintptr_t)i;
                const uint32_t tt = (wcur[p] >> (8 * (int)(adr & 3))) & 0xffu;
                tb[p] = tt > 4u ? 4u : tt;
                const bool cross = (adr & 3) == 3;   /* next row starts a new word: rotate, fetch the one after */
                rot_ld_if(wcur[p], wnxt[p], reinterpret_cast<const uint32_t *>(adr & ~(uintptr_t)3) + 2, cross);
            }
            if (st >= s.end) {
                /* no cell in this row: m == 0, the lane stops here (A.1: gscore check, then break) */
                const int jafter = lb[p] < s.end ? s.end : lb[p];
                if (jafter == s.qlen) {
                    s.max_ie = s.gscore > h1 ? s.max_ie : i;
                    s.gscore = s.gscore > h1 ? s.gscore : h1;
                }
                s.act = 0; died |= 1 << p;
                continue;
            }
            sp[p] = st; ep[p] = s.end; h1i[p] = h1;
        }
        if (died) {
            if (!(ps[0].act | ps[1].act)) break;
            /* zero the dead lane so that it computes zeros from now on */
#pragma unroll
            for (int p = 0; p < 2; ++p)
                if (died & (1 << p)) {
                    const uint32_t keep = p ? 0x0000ffffu : 0xffff0000u;
                    for (int j = 0; j <= ps[p].qlen; ++j) he[j * S] &= keep;
                }
        }
        r.T = tb[0] * 0x0011u | tb[1] * 0x1100u;
        uint32_t f = 0, key = 0, h1 = pack2(h1i[0], h1i[1]);
        int h1cut = 0, cut = -1;                   /* lane whose f chain was cut before the other's end */

        /* generic masked cell: lanes outside their [sp, ep) are left untouched */
#define BSW_CELL_MASKED(J)                                                              \
        {                                                                               \
            const int j_ = (J);                                                         \
            const uint32_t MK = ((j_ >= sp[0] && j_ < ep[0]) ? 0x0000ffffu : 0u) |      \
                                ((j_ >= sp[1] && j_ < ep[1]) ? 0xffff0000u : 0u);       \
            const uint32_t v = he[j_ * S];                                              \
            uint32_t h_, en, fn = f;                                                    \
            bsw_cell<SYM>(r, v, qs[j_ * S], fn, h_, en);                                     \
            he[j_ * S] = st_pack(h1 & MK, en & MK) | (v & ~MK);                         \
            f = fn & MK;                                                                \
            h1 = (h_ & MK) | (h1 & ~MK);                                                \
            key = p_maxu(key, (((h_ & MK) << 8) + ((uint32_t)j_ * 0x00010001u)) & MK);  \
        }

        uint32_t h1s = 0;                          /* running h1 saved at the cut */
        if (ps[0].act && ps[1].act) {
            const int ls = mn(sp[0], sp[1]), hs = mx(sp[0], sp[1]);
            const int e1 = mn(ep[0], ep[1]), e2 = mx(ep[0], ep[1]);
            const int sh = ep[0] < ep[1] ? 0 : 1;                  /* lane that ends first */
            const int sh_clean = sh ? ps[1].clean : ps[0].clean, sh_w = sh ? ps[1].w : ps[0].w;
            const int sh_qlen = sh ? ps[1].qlen : ps[0].qlen;
            const uint32_t keep = sh ? 0x0000ffffu : 0xffff0000u;
            const bool lazy = e1 == e2 || (sh_clean && e2 <= i + sh_w + 1);
            if (ls == hs && lazy && hs < e1) {
                /* common case: one sweep, the short lane's f chain cut at its exact end */
                const int jcut = e1 < e2 ? e1 : -1;
                if (jcut >= 0) cut = sh;
                bsw_sweep_cut<S, SYM>(r, he, qs, hs, e2, jcut, keep, e1 == sh_qlen ? keep : 0xffffffffu, f, h1, key, h1s);
            } else {
                int j = ls;
                const int head_end = mn(hs, e2);
                for (; j < head_end; ++j) BSW_CELL_MASKED(j)
                if (hs < e1) { bsw_sweep_cut<S, SYM>(r, he, qs, hs, e1, -1, 0u, 0u, f, h1, key, h1s); j = e1; }
                if (j < hs) j = hs;
                if (j < e2) {
                    if (lazy && j >= e1) {
                        cut = sh; h1s = h1;
                        f &= keep;
                        if (e1 == sh_qlen) h1 &= keep;              /* entry qlen is never read: keep it zero */
                        uint32_t dummy = 0;
                        bsw_sweep_cut<S, SYM>(r, he, qs, j, e2, -1, 0u, 0u, f, h1, key, dummy);
                    } else {
                        for (; j < e2; ++j) BSW_CELL_MASKED(j)
                    }
                }
            }
            h1cut = half_s(h1s, sh);
        } else {
            /* the other lane is dead and zeroed */
            bsw_sweep_cut<S, SYM>(r, he, qs, ps[0].act ? sp[0] : sp[1], ps[0].act ? ep[0] : ep[1], -1, 0u, 0u, f, h1, key, h1s);
        }
#undef BSW_CELL_MASKED

        /* eh[end] = { h1, 0 }.  When the sweep ended at one common column for both live lanes (equal
         * ends, a lazily extended short lane -- whose half of h1 is 0 there --, or a single live lane
         * next to a zeroed dead one) this is one plain store. */
        bool plain_end = false;
        {
            const int ea = ps[0].act ? ep[0] : -1, eb = ps[1].act ? ep[1] : -1;
            const int ee = ea > eb ? ea : eb;
            const bool same = (ea < 0 || eb < 0 || ea == eb || cut >= 0);
            if (same && ee >= 0) {
                uint32_t hv = h1;
                if (ps[0].act && ea == ps[0].qlen && ps[0].clean && ea == ee) hv &= 0xffff0000u;   /* entry qlen: keep zero */
                if (ps[1].act && eb == ps[1].qlen && ps[1].clean && eb == ee) hv &= 0x0000ffffu;
                he[ee * S] = hv;
                plain_end = true;
            }
        }
        /* ---- per-pair row epilogue (A.1) ---- */
        died = 0;
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            PState &s = ps[p];
            if (!s.act) continue;
            const uint32_t hm = p ? 0xffff0000u : 0x0000ffffu;
            const int h1p = (cut == p) ? h1cut : half_s(h1, p);
            const int kp = half_u(key, p);
            const int m = kp >> 8, mj = kp & 0xff;
            if (cut != p && !plain_end) {   /* eh[end].h = h1; eh[end].e = 0 (a cut lane wrote it during the lazy sweep) */
                const uint32_t v = he[s.end * S];
                he[s.end * S] = (v & ~hm) | ((s.end == s.qlen && s.clean) ? 0u : (pack2(h1p, h1p) & hm));
            }
            if (s.end == s.qlen) {         /* a computing lane has beg < end, so A.1's j after the loop is end */
                s.max_ie = s.gscore > h1p ? s.max_ie : i;
                s.gscore = s.gscore > h1p ? s.gscore : h1p;
            }
            bool stop = (m == 0);
            if (!stop) {
                if (m > s.max) {
                    int off = mj - i; off = off < 0 ? -off : off;
                    s.max = m; s.max_i = i; s.max_j = mj;
                    s.max_off = s.max_off > off ? s.max_off : off;
                } else if (c.zdrop > 0) {
                    const int di = i - s.max_i, dj = mj - s.max_j;
                    if (di > dj) stop = s.max - m - (di - dj) * c.e_del > c.zdrop;
                    else         stop = s.max - m - (dj - di) * c.e_ins > c.zdrop;
                }
            }
            if (!stop) {
                if (COUNT) {
                    int jj = s.beg;
                    while (jj < s.end) { if (he[jj * S] & hm) break; ++jj; }
                    s.beg = jj;
                }
                int jj = s.end;
                if (h1p == 0) {
                    /* m > 0 guarantees a non-zero entry at or right of the sweep start */
                    --jj;
                    while (jj >= 0) { if (he[jj * S] & hm) break; --jj; }
                }
                s.end = jj + 2 < s.qlen ? jj + 2 : s.qlen;
                if (i + 1 >= s.tlen) stop = true;
            }
            if (stop) { s.act = 0; died |= 1 << p; }
        }
        if (died) {
            if (!(ps[0].act | ps[1].act)) break;
#pragma unroll
            for (int p = 0; p < 2; ++p)
                if (died & (1 << p)) {
                    const uint32_t keep = p ? 0x0000ffffu : 0xffff0000u;
                    for (int j = 0; j <= ps[p].qlen; ++j) he[j * S] &= keep;
                }
        }
        /* lazy left trim: skip leading columns that are zero in both lanes (dead lanes are zero) */
        if ((i & 3) == 3) {
            const int lim = mn(ps[0].act ? ps[0].end : BIG, ps[1].act ? ps[1].end : BIG);
            int steps = 0;
            while (steps < 8 && t < lim) { if (he[t * S]) break; ++t; ++steps; }
        }
        ++i;
    }

    PairOut *out[2] = { &oa, &ob };
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        out[p]->score = ps[p].max; out[p]->qle = ps[p].max_j + 1; out[p]->tle = ps[p].max_i + 1;
        out[p]->gtle = ps[p].max_ie + 1; out[p]->gscore = ps[p].gscore; out[p]->max_off = ps[p].max_off;
        out[p]->cells = ps[p].cells;
    }
}

/*
 * General 32-bit path: one thread, one pair, any length / score (pairs the
 * packed kernel does not admit, incl. the reference's scalar tier >= 32768).
 * eh[j*S] = { H(i-1,j-1), E(i,j) } in global memory.  Literal A.1.
 */
template <typename IDX>
BSW_HD void bsw_general_thread(const Const &c, int w, const PairIn &P,
                               const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                               int2 *eh, IDX S, PairOut &o)
{
    const int qlen = P.len2, tlen = P.len1, h0 = P.h0;
    const uint8_t *query = qer + P.idq, *target = ref + P.idr;
    int i, j, beg, end, max, max_i, max_j, max_ie, gscore, max_off;
    long long cells = 0;
    {
        int prev = 0;
        for (j = 0; j <= qlen; ++j) {
            int v;
            if (j == 0) v = h0;
            else if (j == 1) v = h0 > c.oe_ins ? h0 - c.oe_ins : 0;
            else v = prev > c.e_ins ? prev - c.e_ins : 0;
            prev = v;
            eh[(IDX)j * S] = make_int2(v, 0);
        }
    }
    w = eff_band(c, qlen, w);
    max = h0; max_i = max_j = -1; max_ie = -1; gscore = -1; max_off = 0;
    beg = 0; end = qlen;
    for (i = 0; i < tlen; ++i) {
        int f = 0, h1, m = 0, mj = -1;
        int t = target[i]; t = t > 4 ? 4 : t;
        if (beg < i - w) beg = i - w;
        if (end > i + w + 1) end = i + w + 1;
        if (end > qlen) end = qlen;
        if (beg == 0) { h1 = h0 - (c.o_del + c.e_del * (i + 1)); if (h1 < 0) h1 = 0; }
        else h1 = 0;
        if (end > beg) cells += end - beg;
        for (j = beg; j < end; ++j) {
            int2 v = eh[(IDX)j * S];
            int q = query[j]; q = q > 4 ? 4 : q;
            int sc = (q == 4 || t == 4) ? c.ambig : (q == t ? c.a : -c.b);
            int M = v.x ? v.x + sc : 0, e = v.y, h, tt;
            h = M > e ? M : e; h = h > f ? h : f;
            eh[(IDX)j * S].x = h1;
            h1 = h;
            mj = m > h ? mj : j; m = m > h ? m : h;
            tt = M - c.oe_del; tt = tt > 0 ? tt : 0; e -= c.e_del; e = e > tt ? e : tt;
            eh[(IDX)j * S].y = e;
            tt = M - c.oe_ins; tt = tt > 0 ? tt : 0; f -= c.e_ins; f = f > tt ? f : tt;
        }
        eh[(IDX)end * S] = make_int2(h1, 0);
        if ((beg < end ? end : beg) == qlen) {
            max_ie = gscore > h1 ? max_ie : i;
            gscore = gscore > h1 ? gscore : h1;
        }
        if (m == 0) break;
        if (m > max) {
            int off = mj - i; off = off < 0 ? -off : off;
            max = m; max_i = i; max_j = mj; max_off = max_off > off ? max_off : off;
        } else if (c.zdrop > 0) {
            if (i - max_i > mj - max_j) { if (max - m - ((i - max_i) - (mj - max_j)) * c.e_del > c.zdrop) break; }
            else                        { if (max - m - ((mj - max_j) - (i - max_i)) * c.e_ins > c.zdrop) break; }
        }
        for (j = beg; j < end; ++j) { int2 v = eh[(IDX)j * S]; if (v.x | v.y) break; }
        beg = j;
        for (j = end; j >= beg; --j) { int2 v = eh[(IDX)j * S]; if (v.x | v.y) break; }
        end = j + 2 < qlen ? j + 2 : qlen;
    }
    o.score = max; o.qle = max_j + 1; o.tle = max_i + 1; o.gtle = max_ie + 1;
    o.gscore = gscore; o.max_off = max_off; o.cells = cells;
}

} /* namespace bswk */
