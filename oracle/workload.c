/*
 * oracle/workload.c -- deterministic synthetic inputs for the BASELINE.json
 * configs (SURVEY.md section 8d "Synthetic inputs"), shared by tests/ and
 * bench.py through ctypes.  TEST INFRASTRUCTURE (see oracle.h).
 *
 * RNG: xoshiro256** seeded through splitmix64 from (seed, read index), so the
 * output is independent of threading and of how a workload is sliced.
 *
 * BSW pairs follow the shape mem_chain2aln_across_reads_V2 produces
 * (SURVEY.md section 3.1 / row a9): one seed per read, a left pair (both
 * sequences reversed) and a right pair; len1 = len2 + cal_max_gap(len2).
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

typedef struct { uint64_t s[4]; } rng_t;

static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static inline uint64_t splitmix(uint64_t *x)
{
    uint64_t z = (*x += 0x9e3779b97f4a7c15ULL);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
static inline void rng_seed(rng_t *r, uint64_t seed, uint64_t stream)
{
    uint64_t x = seed * 0x2545f4914f6cdd1dULL + stream * 0x9e3779b97f4a7c15ULL + 0x1234567ULL;
    r->s[0] = splitmix(&x); r->s[1] = splitmix(&x); r->s[2] = splitmix(&x); r->s[3] = splitmix(&x);
}
static inline uint64_t rng_next(rng_t *r)
{
    uint64_t *s = r->s, result = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
    return result;
}
static inline uint32_t rng_below(rng_t *r, uint32_t n) { return (uint32_t)((rng_next(r) >> 32) * (uint64_t)n >> 32); }
static inline double rng_unif(rng_t *r) { return (double)(rng_next(r) >> 11) * (1.0 / 9007199254740992.0); }
static inline uint8_t rng_base(rng_t *r, double p_n) { return rng_unif(r) < p_n ? 4 : (uint8_t)rng_below(r, 4); }

typedef struct {
    uint64_t seed;
    int32_t n_reads, read_len;
    int32_t seed_min, seed_max;      /* seed length ~ U[seed_min, seed_max] */
    double  div;                     /* divergence rate; sub:ins:del = 8:1:1 */
    double  p_n;                     /* probability of an N (code 4) per base */
    int32_t a, o_del, e_del, o_ins, e_ins, w;
} wl_bsw_cfg_t;

/* bwa-mem cal_max_gap */
static int cal_max_gap(const wl_bsw_cfg_t *c, int qlen)
{
    int l_del = (int)((double)(qlen * c->a - c->o_del) / c->e_del + 1.);
    int l_ins = (int)((double)(qlen * c->a - c->o_ins) / c->e_ins + 1.);
    int l = l_del > l_ins ? l_del : l_ins;
    l = l > 1 ? l : 1;
    return l < c->w << 1 ? l : c->w << 1;
}

static void read_shape(const wl_bsw_cfg_t *c, int64_t rid, rng_t *r, int *slen, int *qbeg)
{
    rng_seed(r, c->seed, (uint64_t)rid);
    int smax = c->seed_max < c->read_len ? c->seed_max : c->read_len;
    int smin = c->seed_min < smax ? c->seed_min : smax;
    *slen = smin + (int)rng_below(r, (uint32_t)(smax - smin + 1));
    *qbeg = (int)rng_below(r, (uint32_t)(c->read_len - *slen + 1));
}

/* pass 1: sizes.  Per read r: pair slots 2r (left) and 2r+1 (right); a side with
 * len2 == 0 is not emitted (len2 recorded as -1). */
void wl_bsw_sizes(const wl_bsw_cfg_t *c, int32_t *len2_lr, int32_t *len1_lr)
{
    rng_t r; int64_t i;
    for (i = 0; i < c->n_reads; ++i) {
        int s, qb, l, rr;
        read_shape(c, i, &r, &s, &qb);
        l = qb; rr = c->read_len - qb - s;
        len2_lr[2 * i] = l > 0 ? l : -1;       len1_lr[2 * i] = l > 0 ? l + cal_max_gap(c, l) : 0;
        len2_lr[2 * i + 1] = rr > 0 ? rr : -1; len1_lr[2 * i + 1] = rr > 0 ? rr + cal_max_gap(c, rr) : 0;
    }
}

/* target = query mutated at rate div (8:1:1), then padded with random bases to tlen */
static void mutate(rng_t *r, const uint8_t *q, int qlen, uint8_t *t, int tlen, double div, double p_n)
{
    int i = 0, k = 0;
    while (k < tlen && i < qlen) {
        if (rng_unif(r) < div) {
            uint32_t kind = rng_below(r, 10);
            if (kind < 8) {                         /* substitution */
                uint8_t b = q[i] < 4 ? (uint8_t)((q[i] + 1 + rng_below(r, 3)) & 3) : (uint8_t)rng_below(r, 4);
                t[k++] = b; ++i;
            } else if (kind == 8) {                 /* extra base in the target */
                t[k++] = rng_base(r, p_n);
            } else {                                /* base missing from the target */
                ++i;
            }
        } else t[k++] = q[i++];
    }
    while (k < tlen) t[k++] = rng_base(r, p_n);
}

typedef struct {
    const wl_bsw_cfg_t *c;
    ora_seqpair_t *left, *right;        /* compacted arrays */
    const int64_t *lslot, *rslot;       /* per read: index in left/right or -1 */
    const int64_t *loff_r, *loff_q, *roff_r, *roff_q;  /* byte offsets per emitted pair */
    uint8_t *lref, *lqer, *rref, *rqer;
    int64_t lo, hi;
} fill_t;

static void fill_range(const fill_t *f)
{
    const wl_bsw_cfg_t *c = f->c;
    uint8_t *read = (uint8_t *)malloc((size_t)c->read_len + 8);
    rng_t r; int64_t i; int k;
    for (i = f->lo; i < f->hi; ++i) {
        int s, qb, l, rr;
        read_shape(c, i, &r, &s, &qb);
        for (k = 0; k < c->read_len; ++k) read[k] = rng_base(&r, c->p_n);
        l = qb; rr = c->read_len - qb - s;
        if (l > 0) {
            int64_t ix = f->lslot[i];
            ora_seqpair_t *sp = &f->left[ix];
            uint8_t *q = f->lqer + f->loff_q[ix], *t = f->lref + f->loff_r[ix];
            int tl = l + cal_max_gap(c, l);
            for (k = 0; k < l; ++k) q[k] = read[qb - 1 - k];        /* reversed prefix */
            mutate(&r, q, l, t, tl, c->div, c->p_n);
            memset(sp, 0, sizeof *sp);
            sp->idr = (int32_t)f->loff_r[ix]; sp->idq = (int32_t)f->loff_q[ix];
            sp->id = (int32_t)ix; sp->len1 = tl; sp->len2 = l; sp->h0 = s * c->a;
            sp->seqid = (int32_t)i; sp->regid = 0;
        }
        if (rr > 0) {
            int64_t ix = f->rslot[i];
            ora_seqpair_t *sp = &f->right[ix];
            uint8_t *q = f->rqer + f->roff_q[ix], *t = f->rref + f->roff_r[ix];
            int tl = rr + cal_max_gap(c, rr);
            memcpy(q, read + qb + s, (size_t)rr);
            mutate(&r, q, rr, t, tl, c->div, c->p_n);
            memset(sp, 0, sizeof *sp);
            sp->idr = (int32_t)f->roff_r[ix]; sp->idq = (int32_t)f->roff_q[ix];
            sp->id = (int32_t)ix; sp->len1 = tl; sp->len2 = rr; sp->h0 = s * c->a;
            sp->seqid = (int32_t)i; sp->regid = (int32_t)(l > 0 ? f->lslot[i] : -1); /* left partner */
        }
    }
    free(read);
}

static void *fill_thread(void *a) { fill_range((const fill_t *)a); return 0; }

/* pass 2: fill.  Offsets/slots are computed by the caller from wl_bsw_sizes
 * (numpy cumsum).  right[k].regid = index of the left pair of the same read
 * (or -1): the caller sets right[k].h0 = left[regid].score (SURVEY.md row a9:
 * "right h0 = left score"). */
void wl_bsw_fill(const wl_bsw_cfg_t *c, ora_seqpair_t *left, ora_seqpair_t *right,
                 const int64_t *lslot, const int64_t *rslot,
                 const int64_t *loff_r, const int64_t *loff_q,
                 const int64_t *roff_r, const int64_t *roff_q,
                 uint8_t *lref, uint8_t *lqer, uint8_t *rref, uint8_t *rqer, int nthreads)
{
    fill_t base = { c, left, right, lslot, rslot, loff_r, loff_q, roff_r, roff_q, lref, lqer, rref, rqer, 0, 0 };
    if (nthreads <= 1) { base.lo = 0; base.hi = c->n_reads; fill_range(&base); return; }
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)nthreads);
    fill_t *fs = (fill_t *)malloc(sizeof(fill_t) * (size_t)nthreads);
    int t;
    for (t = 0; t < nthreads; ++t) {
        fs[t] = base;
        fs[t].lo = (int64_t)c->n_reads * t / nthreads;
        fs[t].hi = (int64_t)c->n_reads * (t + 1) / nthreads;
        pthread_create(&th[t], 0, fill_thread, &fs[t]);
    }
    for (t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
    free(th); free(fs);
}

/* ---------------------------------------------------------------- kswv (C4) */

typedef struct {
    uint64_t seed;
    int32_t n_pairs, qlen, tmin, tmax;
    double  p_contain;               /* fraction of targets that contain the query */
    double  div, indel_frac;         /* mutation rate inside the contained copy */
    double  p_n;
    int32_t xtra;
} wl_kswv_cfg_t;

/* fixed strides: ref slot = tmax bytes, qer slot = qlen bytes per pair */
void wl_kswv_fill(const wl_kswv_cfg_t *c, ora_seqpair_t *pairs, uint8_t *ref, uint8_t *qer)
{
    rng_t r; int64_t i; int k;
    for (i = 0; i < c->n_pairs; ++i) {
        ora_seqpair_t *sp = &pairs[i];
        uint8_t *q = qer + (int64_t)i * c->qlen, *t = ref + (int64_t)i * c->tmax;
        int tl;
        rng_seed(&r, c->seed, (uint64_t)i);
        tl = c->tmin + (int)rng_below(&r, (uint32_t)(c->tmax - c->tmin + 1));
        for (k = 0; k < c->qlen; ++k) q[k] = rng_base(&r, c->p_n);
        for (k = 0; k < tl; ++k) t[k] = rng_base(&r, c->p_n);
        if (rng_unif(&r) < c->p_contain && tl >= c->qlen) {
            int off = (int)rng_below(&r, (uint32_t)(tl - c->qlen + 1)), j = 0, o = off;
            while (j < c->qlen && o < tl) {
                if (rng_unif(&r) < c->div) {
                    double u = rng_unif(&r);
                    if (u >= c->indel_frac) { t[o++] = q[j] < 4 ? (uint8_t)((q[j] + 1 + rng_below(&r, 3)) & 3) : 0; ++j; }
                    else if (u < c->indel_frac * 0.5) { t[o++] = rng_base(&r, c->p_n); }
                    else ++j;
                } else t[o++] = q[j++];
            }
        }
        memset(sp, 0, sizeof *sp);
        sp->idr = (int32_t)((int64_t)i * c->tmax); sp->idq = (int32_t)((int64_t)i * c->qlen);
        sp->id = (int32_t)i; sp->len1 = tl; sp->len2 = c->qlen; sp->h0 = c->xtra;
        sp->seqid = (int32_t)i; sp->regid = (int32_t)i;
    }
}

/* uniform random pairs of arbitrary lengths (edge-case fuzzing) */
void wl_random_pairs(uint64_t seed, int32_t n, int32_t qmin, int32_t qmax, int32_t tmin, int32_t tmax,
                     int32_t h0min, int32_t h0max, double sim, double p_n,
                     ora_seqpair_t *pairs, uint8_t *ref, uint8_t *qer)
{
    rng_t r; int64_t i; int k;
    for (i = 0; i < n; ++i) {
        ora_seqpair_t *sp = &pairs[i];
        uint8_t *q = qer + (int64_t)i * (qmax > 0 ? qmax : 1), *t = ref + (int64_t)i * (tmax > 0 ? tmax : 1);
        int ql, tl;
        rng_seed(&r, seed, (uint64_t)i);
        ql = qmin + (int)rng_below(&r, (uint32_t)(qmax - qmin + 1));
        tl = tmin + (int)rng_below(&r, (uint32_t)(tmax - tmin + 1));
        for (k = 0; k < ql; ++k) q[k] = rng_base(&r, p_n);
        for (k = 0; k < tl; ++k)
            t[k] = (k < ql && rng_unif(&r) < sim) ? q[k] : rng_base(&r, p_n);
        memset(sp, 0, sizeof *sp);
        sp->idr = (int32_t)(i * (tmax > 0 ? tmax : 1)); sp->idq = (int32_t)(i * (qmax > 0 ? qmax : 1));
        sp->id = (int32_t)i; sp->len1 = tl; sp->len2 = ql;
        sp->h0 = h0min + (int)rng_below(&r, (uint32_t)(h0max - h0min + 1));
        sp->seqid = (int32_t)i; sp->regid = (int32_t)i;
    }
}

/* ------------------------------------------------------------------ CIGAR jobs
 * One job per final alignment, shaped like bwa_gen_cigar2's calls from mem_reg2aln: query = the aligned part of a
 * read (qmin..qmax bases), target = the query mutated at `div` (8:1:1 substitutions : insertions : deletions), its
 * length whatever the mutations leave (both ends aligned: no padding), band by bwa_gen_cigar2's rule from the two
 * lengths and opt->w.  Fixed strides: qer slot = qmax bytes, ref slot = 2 * qmax + 8 bytes per job. */
void wl_global_fill(uint64_t seed, int32_t n, int32_t qmin, int32_t qmax, double div, double p_n, int32_t w_opt,
                    int32_t a, int32_t o_del, int32_t e_del, int32_t o_ins, int32_t e_ins,
                    ora_global_job_t *jobs, uint8_t *ref, uint8_t *qer)
{
    rng_t r; int64_t i; int k;
    const int64_t tstride = 2 * (int64_t)qmax + 8;
    for (i = 0; i < n; ++i) {
        ora_global_job_t *g = &jobs[i];
        uint8_t *q = qer + i * (int64_t)qmax, *t = ref + i * tstride;
        int ql, tl = 0, qi = 0, w, max_ins, max_del, max_gap, min_w, dl;
        rng_seed(&r, seed, (uint64_t)i);
        ql = qmin + (int)rng_below(&r, (uint32_t)(qmax - qmin + 1));
        for (k = 0; k < ql; ++k) q[k] = rng_base(&r, p_n);
        while (qi < ql && tl < tstride - 1) {
            if (rng_unif(&r) < div) {
                uint32_t kind = rng_below(&r, 10);
                if (kind < 8) { t[tl++] = q[qi] < 4 ? (uint8_t)((q[qi] + 1 + rng_below(&r, 3)) & 3) : (uint8_t)rng_below(&r, 4); ++qi; }
                else if (kind == 8) { int gl = 1 + (int)rng_below(&r, 4); while (gl-- > 0 && tl < tstride - 1) t[tl++] = rng_base(&r, p_n); }
                else qi += 1 + (int)rng_below(&r, 4);
            } else t[tl++] = q[qi++];
        }
        /* bwa_gen_cigar2's band */
        max_ins = (int)((double)(((ql + 1) >> 1) * a - o_ins) / e_ins + 1.);
        max_del = (int)((double)(((ql + 1) >> 1) * a - o_del) / e_del + 1.);
        max_gap = max_ins > max_del ? max_ins : max_del;
        max_gap = max_gap > 1 ? max_gap : 1;
        dl = tl > ql ? tl - ql : ql - tl;
        w = (max_gap + dl + 1) >> 1;
        w = w < w_opt ? w : w_opt;
        min_w = dl + 3;
        w = w > min_w ? w : min_w;
        g->idr = (int32_t)(i * tstride); g->idq = (int32_t)(i * (int64_t)qmax);
        g->tlen = tl; g->qlen = ql; g->w = w; g->score = 0; g->n_cigar = 0; g->cigar_off = 0;
    }
}
