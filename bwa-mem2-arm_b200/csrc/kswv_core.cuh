/*
 * kswv_core.cuh -- shared pieces of the batched mate-rescue local alignment
 * (kswv::getScores8/16, bwa-mem2 src/kswv.cpp; semantics = bwa ksw_u8 / ksw_i16 evaluated in exact arithmetic,
 * SURVEY.md Appendix A.2): flags, result record, the admission rule of the byte-key kernels, the replay of ksw_u8's
 * `b` list, and the general 32-bit thread-per-pair path (rows and row maxima in HBM) for what the packed kernel does not
 * admit.  The packed kernel itself -- DP state in registers, rows skewed over the lanes of a group -- is kswv_sys.cuh.
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
