/*
 * oracle/global.c -- scalar CPU restatement of the CIGAR-producing global alignment that follows the extension
 * (TEST INFRASTRUCTURE, see oracle.h: "parity unpinned"; never linked into the product library).
 *
 * Follows bwa 0.7.17 ksw.c ksw_global2 == bwa-mem2 v2.2.1 src/ksw.cpp ksw_global2, called per alignment from
 * bwa_gen_cigar2 / mem_reg2aln: banded Needleman-Wunsch with affine gaps, target in the
 * outer loop, a backtrack byte per cell (h direction in bits 0-1, "E extends" in bit 2, "F extends" in bit 5 = 2 << 4),
 * backtrack from the last cell, CIGAR ops 0 = M, 1 = I (query only), 2 = D (target only), length << 4 | op.
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

#define MINUS_INF (-0x40000000)

typedef struct { int32_t h, e; } geh_t;

static uint32_t *push_cigar(int *n_cigar, int *m_cigar, uint32_t *cigar, int op, int len)
{
    if (*n_cigar == 0 || op != (int)(cigar[(*n_cigar) - 1] & 0xf)) {
        if (*n_cigar == *m_cigar) {
            *m_cigar = *m_cigar ? (*m_cigar) << 1 : 4;
            cigar = (uint32_t *)realloc(cigar, (size_t)(*m_cigar) << 2);
        }
        cigar[(*n_cigar)++] = (uint32_t)len << 4 | (uint32_t)op;
    } else cigar[(*n_cigar) - 1] += (uint32_t)len << 4;
    return cigar;
}

/* returns the score; *n_cigar_ / *cigar_ (malloc'ed, caller frees) when both are non-null */
int ora_ksw_global2(int qlen, const uint8_t *query, int tlen, const uint8_t *target, const ora_bsw_params_t *p, int w,
                    int *n_cigar_, uint32_t **cigar_, int64_t *cells_)
{
    const int m = 5;
    const int8_t *mat = p->mat;
    const int o_del = p->o_del, e_del = p->e_del, o_ins = p->o_ins, e_ins = p->e_ins;
    const int oe_del = o_del + e_del, oe_ins = o_ins + e_ins;
    int i, j, score, n_col;
    int64_t cells = 0;
    uint8_t *z;
    geh_t *eh;
    if (n_cigar_) *n_cigar_ = 0;
    n_col = qlen < 2 * w + 1 ? qlen : 2 * w + 1;
    z = (n_cigar_ && cigar_) ? (uint8_t *)malloc((size_t)(n_col > 0 ? n_col : 1) * (size_t)(tlen > 0 ? tlen : 1)) : 0;
    eh = (geh_t *)calloc((size_t)qlen + 1, sizeof(geh_t));
    /* first row */
    eh[0].h = 0; eh[0].e = MINUS_INF;
    for (j = 1; j <= qlen && j <= w; ++j) { eh[j].h = -(o_ins + e_ins * j); eh[j].e = MINUS_INF; }
    for (; j <= qlen; ++j) eh[j].h = eh[j].e = MINUS_INF;
    for (i = 0; i < tlen; ++i) {
        int32_t f = MINUS_INF, h1, beg, end, t;
        const int8_t *q = &mat[target[i] * m];
        uint8_t *zi = z ? &z[(size_t)i * (size_t)n_col] : 0;
        beg = i > w ? i - w : 0;
        end = i + w + 1 < qlen ? i + w + 1 : qlen;
        h1 = beg == 0 ? -(o_del + e_del * (i + 1)) : MINUS_INF;
        if (end > beg) cells += end - beg;
        for (j = beg; j < end; ++j) {
            geh_t *c = &eh[j];
            int32_t h, M = c->h, e = c->e;
            uint8_t d;
            c->h = h1;
            M += q[query[j]];
            d = M >= e ? 0 : 1;
            h = M >= e ? M : e;
            d = h >= f ? d : 2;
            h = h >= f ? h : f;
            h1 = h;
            t = M - oe_del;
            e -= e_del;
            d |= e > t ? 1 << 2 : 0;
            e = e > t ? e : t;
            c->e = e;
            t = M - oe_ins;
            f -= e_ins;
            d |= f > t ? 2 << 4 : 0;
            f = f > t ? f : t;
            if (zi) zi[j - beg] = d;
        }
        eh[end].h = h1; eh[end].e = MINUS_INF;
    }
    score = eh[qlen].h;
    if (z) {
        int n_cigar = 0, m_cigar = 0, which = 0, k;
        uint32_t *cigar = 0, tmp;
        i = tlen - 1; k = (i + w + 1 < qlen ? i + w + 1 : qlen) - 1;
        while (i >= 0 && k >= 0) {
            which = z[(size_t)i * (size_t)n_col + (size_t)(k - (i > w ? i - w : 0))] >> (which << 1) & 3;
            if (which == 0) { cigar = push_cigar(&n_cigar, &m_cigar, cigar, 0, 1); --i; --k; }
            else if (which == 1) { cigar = push_cigar(&n_cigar, &m_cigar, cigar, 2, 1); --i; }
            else { cigar = push_cigar(&n_cigar, &m_cigar, cigar, 1, 1); --k; }
        }
        if (i >= 0) cigar = push_cigar(&n_cigar, &m_cigar, cigar, 2, i + 1);
        if (k >= 0) cigar = push_cigar(&n_cigar, &m_cigar, cigar, 1, k + 1);
        for (i = 0; i < n_cigar >> 1; ++i) { tmp = cigar[i]; cigar[i] = cigar[n_cigar - 1 - i]; cigar[n_cigar - 1 - i] = tmp; }
        *n_cigar_ = n_cigar; *cigar_ = cigar;
    }
    free(eh); free(z);
    if (cells_) *cells_ = cells;
    return score;
}

/* ------------------------------------------------------------------ batch driver
 * jobs[i]: target = ref + idr (tlen bases), query = qer + idq (qlen bases), band w.  Outputs: score, n_cigar, and the
 * ops at cigar[cigar_off .. cigar_off + n_cigar) with cigar_off = sum of the n_cigar of the jobs before i (dense, job order). */
typedef struct {
    ora_global_job_t *jobs; const uint8_t *ref, *qer; int32_t n; const ora_bsw_params_t *p;
    uint32_t **tmp; volatile int32_t next; int64_t total; pthread_mutex_t mu;
} gjob_t;

static void *gworker(void *arg)
{
    gjob_t *jb = (gjob_t *)arg;
    int64_t acc = 0;
    for (;;) {
        int32_t lo = __sync_fetch_and_add(&jb->next, 64), hi, i;
        if (lo >= jb->n) break;
        hi = lo + 64 < jb->n ? lo + 64 : jb->n;
        for (i = lo; i < hi; ++i) {
            ora_global_job_t *g = &jb->jobs[i];
            int nc = 0; int64_t c = 0;
            g->score = ora_ksw_global2(g->qlen, jb->qer + g->idq, g->tlen, jb->ref + g->idr, jb->p, g->w, &nc, &jb->tmp[i], &c);
            g->n_cigar = nc;
            acc += c;
        }
    }
    pthread_mutex_lock(&jb->mu); jb->total += acc; pthread_mutex_unlock(&jb->mu);
    return 0;
}

/* cigar: room for cigar_cap ops; returns the DP cells, or -1 when the ops do not fit (n_cigar are still filled in) */
int64_t ora_global2_batch(ora_global_job_t *jobs, const uint8_t *ref, const uint8_t *qer, int32_t n,
                          const ora_bsw_params_t *p, uint32_t *cigar, int64_t cigar_cap, int nthreads)
{
    gjob_t jb; memset(&jb, 0, sizeof jb);
    int32_t i; int t; int64_t off = 0; int fits = 1;
    jb.jobs = jobs; jb.ref = ref; jb.qer = qer; jb.n = n; jb.p = p;
    jb.tmp = (uint32_t **)calloc((size_t)(n > 0 ? n : 1), sizeof(uint32_t *));
    pthread_mutex_init(&jb.mu, 0);
    if (nthreads <= 1) gworker(&jb);
    else {
        pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)nthreads);
        for (t = 0; t < nthreads; ++t) pthread_create(&th[t], 0, gworker, &jb);
        for (t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
        free(th);
    }
    pthread_mutex_destroy(&jb.mu);
    for (i = 0; i < n; ++i) {
        jobs[i].cigar_off = (int32_t)off;
        if (off + jobs[i].n_cigar <= cigar_cap) memcpy(cigar + off, jb.tmp[i], (size_t)jobs[i].n_cigar * 4);
        else fits = 0;
        off += jobs[i].n_cigar;
        free(jb.tmp[i]);
    }
    free(jb.tmp);
    return fits ? jb.total : -1;
}
