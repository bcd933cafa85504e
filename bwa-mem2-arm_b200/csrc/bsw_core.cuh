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
 * This header holds the packed-op layer (with its host emulation), the row-state
 * word helpers and the general 32-bit thread-per-pair path; the row loop of the
 * packed kernel is in bsw_core2.cuh.
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
