/*
 * global_core.cuh -- CIGAR generation: banded global alignment with backtrack, one thread per alignment
 * (semantics = bwa ksw.c ksw_global2, which bwa-mem2's mem_reg2aln reaches through
 * bwa_gen_cigar2 once per reported alignment).
 *
 * Differences from the extension kernels: the DP is Needleman-Wunsch (negative scores, -inf outside the band, 32-bit
 * state), every cell leaves a direction byte, and the product is the op list, not a score.  A job is ~3 k cells
 * (150 x (2w + 1), w = 5 .. 40 by bwa_gen_cigar2's rule), 6x fewer than an extension, so the mapping is the plain one:
 *   - the rolling row eh[] is addressed modulo M = 2 w_max + 2: at row i only the columns [i - w, i + w + 1] are live
 *     (read in [beg, end), written at [beg, end]), so 2w + 2 slots never collide; it lives in an interleaved global
 *     scratch (slot j of thread t at eh[j * stride + t]: a warp's accesses coalesce when its jobs run in step);
 *   - direction bytes are packed four to a word before they are stored; row pitch = ceil(n_col / 4) words;
 *   - eh[qlen].h (the score) is tracked in a register, because its slot may be reused when the band never reaches the
 *     last column;
 *   - the backtrack runs twice: global2_backtrack<false> counts the ops right after the DP (same thread, its own
 *     stores), an exclusive scan over the counts places every job in the dense op array, global2_backtrack<true>
 *     walks again and writes the ops back to front -- which is the reversal ksw_global2 does at its end.
 */
#pragma once
#include "bsw_core.cuh"

namespace bswk {

#define GLOBAL_MINUS_INF (-0x40000000)

/* the job record of the C ABI (include/bsw_cuda.h bsw_global_job_t) */
struct GJob { int idr, idq, tlen, qlen, w, score, n_cigar, cigar_off; };

BSW_HD int global2_ncol(int qlen, int w) { return qlen < 2 * w + 1 ? qlen : 2 * w + 1; }
BSW_HD int global2_pitch(int qlen, int w) { const int nc = global2_ncol(qlen, w); return nc > 0 ? (nc + 3) >> 2 : 0; }
/* words of direction bytes a job needs */
BSW_HD long long global2_zwords(int tlen, int qlen, int w) { return (long long)global2_pitch(qlen, w) * (long long)(tlen > 0 ? tlen : 0); }

BSW_HD int global2_score(const Const &c, uint32_t t, uint32_t q)
{
    return (t > 3u || q > 3u) ? c.ambig : (t == q ? c.a : -c.b);
}

/* DP of one job.  eh: slot j at eh[(j % M) * eh_stride]; z: the job's direction words.  Returns the score. */
BSW_HD int global2_dp(const Const &c, int tlen, int qlen, int w, const uint8_t *__restrict__ target,
                      const uint8_t *__restrict__ query, int2 *eh, long long eh_stride, int M, uint32_t *z, long long &cells_out)
{
    const int o_del = c.o_del, e_del = c.e_del, o_ins = c.o_ins, e_ins = c.e_ins;
    const int oe_del = o_del + e_del, oe_ins = o_ins + e_ins;
    const int pitch = global2_pitch(qlen, w);
    long long cells = 0;
    /* first row: only the slots the band can read before they are written (j <= w) need their values */
    for (int j = 0; j <= qlen && j < M; ++j) {
        int2 v;
        v.x = j == 0 ? 0 : (j <= w ? -(o_ins + e_ins * j) : GLOBAL_MINUS_INF);
        v.y = GLOBAL_MINUS_INF;
        eh[(long long)j * eh_stride] = v;
    }
    int hq = qlen == 0 ? 0 : (qlen <= w ? -(o_ins + e_ins * qlen) : GLOBAL_MINUS_INF);      /* eh[qlen].h */
    for (int i = 0; i < tlen; ++i) {
        int f = GLOBAL_MINUS_INF;
        const int beg = i > w ? i - w : 0;
        const int end = i + w + 1 < qlen ? i + w + 1 : qlen;
        int h1 = beg == 0 ? -(o_del + e_del * (i + 1)) : GLOBAL_MINUS_INF;
        const uint32_t tb = target[i];
        uint32_t *zi = z + (long long)i * pitch;
        uint32_t zw = 0;
        if (end > beg) cells += end - beg;
        int slot = beg % M;
        for (int j = beg; j < end; ++j) {
            int2 *p = &eh[(long long)slot * eh_stride];
            const int2 cur = *p;
            int m = cur.x, e = cur.y;
            m += global2_score(c, tb, query[j]);
            uint32_t d = m >= e ? 0u : 1u;
            int h = m >= e ? m : e;
            d = h >= f ? d : 2u;
            h = h >= f ? h : f;
            int t = m - oe_del;
            e -= e_del;
            d |= e > t ? 1u << 2 : 0u;
            e = e > t ? e : t;
            int2 nv; nv.x = h1; nv.y = e;
            *p = nv;
            h1 = h;
            t = m - oe_ins;
            f -= e_ins;
            d |= f > t ? 2u << 4 : 0u;
            f = f > t ? f : t;
            const int jj = j - beg;
            zw |= d << (8 * (jj & 3));
            if ((jj & 3) == 3) { zi[jj >> 2] = zw; zw = 0; }
            slot = slot + 1 == M ? 0 : slot + 1;
        }
        if (end > beg && ((end - beg) & 3)) zi[(end - beg - 1) >> 2] = zw;
        {   /* eh[end] = { h1, -inf } */
            int2 nv; nv.x = h1; nv.y = GLOBAL_MINUS_INF;
            eh[(long long)(end % M) * eh_stride] = nv;
            if (end == qlen) hq = h1;
        }
    }
    cells_out = cells;
    return hq;
}

BSW_HD uint32_t global2_dir(const uint32_t *z, int pitch, int i, int jj)
{
    return (z[(long long)i * pitch + (jj >> 2)] >> (8 * (jj & 3))) & 0xffu;
}

/* Backtrack from the last cell.  WRITE = false: returns the number of ops.  WRITE = true: also stores them, op k of
 * the walk (last op of the CIGAR first) at out[n_cigar - 1 - k]. */
template <bool WRITE>
BSW_HD int global2_backtrack(int tlen, int qlen, int w, const uint32_t *z, uint32_t *out, int n_cigar)
{
    const int pitch = global2_pitch(qlen, w);
    int n = 0, which = 0;
    uint32_t cur = 0;                          /* the op being extended: len << 4 | op; valid when n > 0 */
    int i = tlen - 1, k = (i + w + 1 < qlen ? i + w + 1 : qlen) - 1;
    auto push = [&](uint32_t op, uint32_t len) {
        if (n == 0 || op != (cur & 0xfu)) {
            if (WRITE && n > 0) out[n_cigar - n] = cur;
            cur = len << 4 | op; ++n;
        } else cur += len << 4;
    };
    while (i >= 0 && k >= 0) {
        const uint32_t d = global2_dir(z, pitch, i, k - (i > w ? i - w : 0));
        which = (int)((d >> (which << 1)) & 3u);
        if (which == 0) { push(0u, 1u); --i; --k; }
        else if (which == 1) { push(2u, 1u); --i; }
        else { push(1u, 1u); --k; }
    }
    if (i >= 0) push(2u, (uint32_t)(i + 1));
    if (k >= 0) push(1u, (uint32_t)(k + 1));
    if (WRITE && n > 0) out[n_cigar - n] = cur;
    return n;
}

} /* namespace bswk */
