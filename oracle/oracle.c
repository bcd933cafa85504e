/*
 * oracle/oracle.c -- scalar CPU restatement of the hot path (TEST INFRASTRUCTURE,
 * see oracle.h: "parity unpinned"; never linked into the product library).
 *
 * Follows, statement by statement:
 *   BSW   : bwa 0.7.17 ksw.c ksw_extend2  == bwa-mem2 v2.2.1 src/bandedSWA.cpp
 *           scalarBandedSWA / smithWaterman{128,256,512}_{8,16} semantics
 *           (SURVEY.md section 8 rows a3-a6, Appendix A.1).
 *   kswv  : bwa 0.7.17 ksw.c ksw_u8 / ksw_i16 / ksw_align2 == bwa-mem2 v2.2.1
 *           src/kswv.cpp kswv512_u8 / kswv512_16 semantics, evaluated in exact
 *           arithmetic with the 8-bit overflow rule.
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

/* ------------------------------------------------------------------ BSW */

typedef struct { int32_t h, e; } eh_t;

/* SURVEY.md Appendix A.1; bwa ksw.c ksw_extend2 */
int ora_bsw_extend(int qlen, const uint8_t *query, int tlen, const uint8_t *target,
                   const ora_bsw_params_t *p, int w, int h0,
                   int *_qle, int *_tle, int *_gtle, int *_gscore, int *_max_off,
                   int64_t *_cells)
{
    const int m = 5;
    const int8_t *mat = p->mat;
    const int o_del = p->o_del, e_del = p->e_del, o_ins = p->o_ins, e_ins = p->e_ins;
    const int zdrop = p->zdrop, end_bonus = p->end_bonus;
    int i, j, k, oe_del = o_del + e_del, oe_ins = o_ins + e_ins;
    int beg, end, max, max_i, max_j, max_ins, max_del, max_ie, gscore, max_off;
    int64_t cells = 0;
    eh_t *eh = (eh_t *)calloc((size_t)qlen + 1, sizeof(eh_t));

    /* first row (A.1: H[0]=h0; H[1]=h0>oe_ins?h0-oe_ins:0; then -e_ins while >e_ins) */
    eh[0].h = h0;
    if (qlen >= 1) eh[1].h = h0 > oe_ins ? h0 - oe_ins : 0;
    for (j = 2; j <= qlen && eh[j - 1].h > e_ins; ++j) eh[j].h = eh[j - 1].h - e_ins;
    /* band adjustment */
    k = m * m;
    for (i = 0, max = 0; i < k; ++i) max = max > mat[i] ? max : mat[i];
    max_ins = (int)((double)(qlen * max + end_bonus - o_ins) / e_ins + 1.);
    max_ins = max_ins > 1 ? max_ins : 1;
    w = w < max_ins ? w : max_ins;
    max_del = (int)((double)(qlen * max + end_bonus - o_del) / e_del + 1.);
    max_del = max_del > 1 ? max_del : 1;
    w = w < max_del ? w : max_del;
    /* DP loop */
    max = h0; max_i = max_j = -1; max_ie = -1; gscore = -1; max_off = 0;
    beg = 0; end = qlen;
    for (i = 0; i < tlen; ++i) {
        int t, f = 0, h1, mrow = 0, mj = -1;
        const int8_t *q = &mat[target[i] * m];
        if (beg < i - w) beg = i - w;
        if (end > i + w + 1) end = i + w + 1;
        if (end > qlen) end = qlen;
        if (beg == 0) {
            h1 = h0 - (o_del + e_del * (i + 1));
            if (h1 < 0) h1 = 0;
        } else h1 = 0;
        if (end > beg) cells += end - beg;
        for (j = beg; j < end; ++j) {
            eh_t *c = &eh[j];
            int h, M = c->h, e = c->e;
            c->h = h1;
            M = M ? M + q[query[j]] : 0;
            h = M > e ? M : e;
            h = h > f ? h : f;
            h1 = h;
            mj = mrow > h ? mj : j;      /* ties -> last j */
            mrow = mrow > h ? mrow : h;
            t = M - oe_del; t = t > 0 ? t : 0;
            e -= e_del; e = e > t ? e : t;
            c->e = e;
            t = M - oe_ins; t = t > 0 ? t : 0;
            f -= e_ins; f = f > t ? f : t;
        }
        eh[end].h = h1; eh[end].e = 0;
        if (j == qlen) {
            max_ie = gscore > h1 ? max_ie : i;   /* ties -> later i */
            gscore = gscore > h1 ? gscore : h1;
        }
        if (mrow == 0) break;
        if (mrow > max) {
            int off = mj - i; if (off < 0) off = -off;
            max = mrow; max_i = i; max_j = mj;
            max_off = max_off > off ? max_off : off;
        } else if (zdrop > 0) {
            if (i - max_i > mj - max_j) {
                if (max - mrow - ((i - max_i) - (mj - max_j)) * e_del > zdrop) break;
            } else {
                if (max - mrow - ((mj - max_j) - (i - max_i)) * e_ins > zdrop) break;
            }
        }
        for (j = beg; j < end && eh[j].h == 0 && eh[j].e == 0; ++j) ;
        beg = j;
        for (j = end; j >= beg && eh[j].h == 0 && eh[j].e == 0; --j) ;
        end = j + 2 < qlen ? j + 2 : qlen;
    }
    free(eh);
    if (_qle) *_qle = max_j + 1;
    if (_tle) *_tle = max_i + 1;
    if (_gtle) *_gtle = max_ie + 1;
    if (_gscore) *_gscore = gscore;
    if (_max_off) *_max_off = max_off;
    if (_cells) *_cells = cells;
    return max;
}

/* ------------------------------------------------------------------ kswv */

/* one ksw_u8/ksw_i16 pass in exact arithmetic. */
void ora_ksw_pass(int qlen, const uint8_t *query, int tlen, const uint8_t *target,
                  const ora_bsw_params_t *p, int xtra, int is8,
                  ora_kswr_t *r, int64_t *_cells)
{
    const int8_t *mat = p->mat;
    const int e_del = p->e_del, e_ins = p->e_ins;
    const int oe_del = p->o_del + p->e_del, oe_ins = p->o_ins + p->e_ins;
    int i, j, k, shift = 127, mx = 0;
    int minsc = (xtra & ORA_KSW_XSUBO) ? (xtra & 0xffff) : 0x10000;
    int endsc = (xtra & ORA_KSW_XSTOP) ? (xtra & 0xffff) : 0x10000;
    int gmax = 0, te = -1, n_b = 0, m_b = 0;
    int64_t cells = 0;
    int64_t *b = 0;                       /* (imax<<32 | row) as in ksw_u8 */
    int32_t *H = (int32_t *)calloc((size_t)qlen + 1, sizeof(int32_t));   /* H[j+1] = H(i-1, j) */
    int32_t *E = (int32_t *)calloc((size_t)qlen + 1, sizeof(int32_t));
    int32_t *Hmax = (int32_t *)calloc((size_t)qlen + 1, sizeof(int32_t));

    for (k = 0; k < 25; ++k) {            /* ksw_qinit: shift = |min|, max = max entry */
        if (mat[k] < shift) shift = mat[k];
        if (mat[k] > mx) mx = mat[k];
    }
    shift = shift < 0 ? -shift : shift;   /* ksw_qinit: q->shift = 256 - (uint8)min  == |min| */

    r->score = 0; r->te = -1; r->qe = -1; r->score2 = -1; r->te2 = -1; r->tb = -1; r->qb = -1;

    for (i = 0; i < tlen; ++i) {
        int f = 0, imax = 0, hdiag = 0;   /* first column: H(i,-1) = 0 */
        const int8_t *srow = &mat[target[i] * 5];
        cells += qlen;
        for (j = 0; j < qlen; ++j) {
            int h = hdiag + srow[query[j]];
            int e = E[j], t;
            hdiag = H[j + 1];
            if (h < 0) h = 0;             /* saturating subtract of the bias */
            h = h > e ? h : e;
            h = h > f ? h : f;
            imax = imax > h ? imax : h;
            H[j + 1] = h;
            e -= e_del; if (e < 0) e = 0;
            t = h - oe_del; if (t < 0) t = 0;
            E[j] = e > t ? e : t;
            f -= e_ins; if (f < 0) f = 0;
            t = h - oe_ins; if (t < 0) t = 0;
            f = f > t ? f : t;
        }
        if (imax >= minsc) {
            if (n_b == 0 || (int32_t)b[n_b - 1] + 1 != i) {
                if (n_b == m_b) {
                    m_b = m_b ? m_b << 1 : 8;
                    b = (int64_t *)realloc(b, 8 * (size_t)m_b);
                }
                b[n_b++] = (int64_t)imax << 32 | i;
            } else if ((int)(b[n_b - 1] >> 32) < imax) b[n_b - 1] = (int64_t)imax << 32 | i;
        }
        if (imax > gmax) {
            gmax = imax; te = i;
            memcpy(Hmax, H + 1, (size_t)qlen * sizeof(int32_t));
            if ((is8 && gmax + shift >= 255) || gmax >= endsc) break;
        }
    }
    r->score = (is8 && gmax + shift >= 255) ? 255 : gmax;
    r->te = te;
    if (!(is8 && r->score == 255)) {
        int max = -1, low, high;
        for (j = 0; j < qlen; ++j)          /* smallest j attaining the row maximum */
            if (Hmax[j] > max) { max = Hmax[j]; r->qe = j; }
        if (b) {
            i = (r->score + mx - 1) / mx;
            low = te - i; high = te + i;
            for (i = 0; i < n_b; ++i) {
                int e = (int32_t)b[i];
                if ((e < low || e > high) && (int)(b[i] >> 32) > r->score2) {
                    r->score2 = (int)(b[i] >> 32); r->te2 = e;
                }
            }
        }
    }
    free(b); free(H); free(E); free(Hmax);
    if (_cells) *_cells = cells;
}

static void revseq(int l, uint8_t *s)
{
    int i;
    for (i = 0; i < l >> 1; ++i) { uint8_t t = s[i]; s[i] = s[l - 1 - i]; s[l - 1 - i] = t; }
}

/* bwa ksw.c ksw_align2 (forward, hand-off, reverse on the reversed prefixes) */
void ora_ksw_align2(int qlen, const uint8_t *query, int tlen, const uint8_t *target,
                    const ora_bsw_params_t *p, int xtra, ora_kswr_t *r, int64_t *_cells)
{
    int is8 = (xtra & ORA_KSW_XBYTE) ? 1 : 0;
    int64_t c0 = 0, c1 = 0;
    ora_kswr_t rr;
    ora_ksw_pass(qlen, query, tlen, target, p, xtra, is8, r, &c0);
    if (_cells) *_cells = c0;
    if ((xtra & ORA_KSW_XSTART) == 0 || ((xtra & ORA_KSW_XSUBO) && r->score < (xtra & 0xffff))) return;
    {
        int ql = r->qe + 1, tl = r->te + 1;
        uint8_t *q2 = (uint8_t *)malloc((size_t)(ql > 0 ? ql : 1));
        uint8_t *t2 = (uint8_t *)malloc((size_t)(tl > 0 ? tl : 1));
        if (ql > 0) memcpy(q2, query, (size_t)ql);
        if (tl > 0) memcpy(t2, target, (size_t)tl);
        revseq(ql, q2); revseq(tl, t2);
        ora_ksw_pass(ql, q2, tl, t2, p, ORA_KSW_XSTOP | r->score, is8, &rr, &c1);
        free(q2); free(t2);
        if (r->score == rr.score) { r->tb = r->te - rr.te; r->qb = r->qe - rr.qe; }
        if (_cells) *_cells = c0 + c1;
    }
}

/* ------------------------------------------------------------------ batch drivers */

typedef struct {
    int kind;                      /* 0 = bsw, 1 = kswv */
    ora_seqpair_t *pairs; const uint8_t *ref, *qer; int32_t n;
    const ora_bsw_params_t *p; int32_t w;
    ora_kswr_t *aln; int is8, phase;
    int64_t *cells_out;
    volatile int32_t next; int32_t chunk;
    int64_t total; pthread_mutex_t mu;
} job_t;

static void run_range(job_t *jb, int32_t lo, int32_t hi, int64_t *acc)
{
    int32_t i;
    for (i = lo; i < hi; ++i) {
        ora_seqpair_t *sp = &jb->pairs[i];
        int64_t c = 0;
        if (jb->kind == 0) {
            int qle, tle, gtle, gscore, max_off;
            sp->score = ora_bsw_extend(sp->len2, jb->qer + sp->idq, sp->len1, jb->ref + sp->idr,
                                       jb->p, jb->w, sp->h0, &qle, &tle, &gtle, &gscore, &max_off, &c);
            sp->qle = qle; sp->tle = tle; sp->gtle = gtle; sp->gscore = gscore; sp->max_off = max_off;
        } else {
            ora_kswr_t r;
            ora_ksw_pass(sp->len2, jb->qer + sp->idq, sp->len1, jb->ref + sp->idr,
                         jb->p, sp->h0, jb->is8, &r, &c);
            if (jb->phase == 0) jb->aln[i] = r;
            else {
                ora_kswr_t *a = &jb->aln[sp->regid];
                if (a->score == r.score) { a->tb = a->te - r.te; a->qb = a->qe - r.qe; }
            }
        }
        if (jb->cells_out) jb->cells_out[i] = c;
        *acc += c;
    }
}

static void *worker(void *arg)
{
    job_t *jb = (job_t *)arg;
    int64_t acc = 0;
    for (;;) {
        int32_t lo = __sync_fetch_and_add(&jb->next, jb->chunk);
        int32_t hi;
        if (lo >= jb->n) break;
        hi = lo + jb->chunk < jb->n ? lo + jb->chunk : jb->n;
        run_range(jb, lo, hi, &acc);
    }
    pthread_mutex_lock(&jb->mu); jb->total += acc; pthread_mutex_unlock(&jb->mu);
    return 0;
}

static int64_t run_job(job_t *jb, int nthreads)
{
    jb->next = 0; jb->chunk = 256; jb->total = 0;
    pthread_mutex_init(&jb->mu, 0);
    if (nthreads <= 1 || jb->n < 2 * jb->chunk) {
        int64_t acc = 0;
        run_range(jb, 0, jb->n, &acc);
        jb->total = acc;
    } else {
        pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)nthreads);
        int t;
        for (t = 0; t < nthreads; ++t) pthread_create(&th[t], 0, worker, jb);
        for (t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
        free(th);
    }
    pthread_mutex_destroy(&jb->mu);
    return jb->total;
}

int64_t ora_bsw_batch(ora_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer,
                      int32_t n, const ora_bsw_params_t *p, int32_t w, int nthreads,
                      int64_t *cells_out)
{
    job_t jb; memset(&jb, 0, sizeof jb);
    jb.kind = 0; jb.pairs = pairs; jb.ref = ref; jb.qer = qer; jb.n = n; jb.p = p; jb.w = w;
    jb.cells_out = cells_out;
    return run_job(&jb, nthreads);
}

int64_t ora_kswv_batch(const ora_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer,
                       ora_kswr_t *aln, int32_t n, const ora_bsw_params_t *p, int is8,
                       int phase, int nthreads, int64_t *cells_out)
{
    job_t jb; memset(&jb, 0, sizeof jb);
    jb.kind = 1; jb.pairs = (ora_seqpair_t *)pairs; jb.ref = ref; jb.qer = qer; jb.n = n; jb.p = p;
    jb.aln = aln; jb.is8 = is8; jb.phase = phase; jb.cells_out = cells_out;
    /* phase 1 may have several pairs per aln entry only by caller error; entries are disjoint */
    return run_job(&jb, nthreads);
}
