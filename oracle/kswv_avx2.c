/*
 * oracle/kswv_avx2.c -- inter-sequence SIMD CPU port of the kswv path (TEST / BASELINE INFRASTRUCTURE, see oracle.h;
 * never linked into the product library).
 *
 * What it stands for: the reference's x86 kswv kernels (bwa-mem2 v2.2.1 src/kswv.cpp) run
 * one pair per SIMD lane -- upstream compiles them for AVX512BW only; this is the same organisation on the 16-bit lanes
 * of an AVX2 register, 16 pairs side by side, pairs sorted by length.  Those sources are not in /root/reference
 * (docs-only snapshot), so this is a PORT written from the algorithm (oracle.c ora_ksw_pass == ksw_u8 / ksw_i16 in
 * exact arithmetic), not the reference's code: bench.py reports it as cpu_baseline kind "port-avx2".  It is bit-exact
 * against ora_ksw_pass on every field and on the cell count (tests/test_oracle.py).
 *
 * Lane l of a group owns one pair.  H[j] / E[j] are vectors over the 16 pairs; a row sweeps the columns of the longest
 * query of the group, lanes whose query is shorter are masked per column; every lane carries its own (gmax, te, qe), stop
 * row and live flag; the row maxima go to a [row][lane] array from which ksw_u8's `b` list is replayed per lane once te
 * and score are final.  Pairs whose score bound does not fit 15 bits, or a scoring matrix that is not the
 * (a, -b, ambiguous) shape of bwa_fill_scmat, go to the scalar function.
 */
#include "oracle.h"
#include <immintrin.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

#define LANES 16
typedef __m256i v16;

static inline v16 vset(int x) { return _mm256_set1_epi16((short)x); }
static inline v16 vsel(v16 a, v16 b, v16 m) { return _mm256_blendv_epi8(a, b, m); }      /* m ? b : a (m = 0 / -1 lanes) */
static inline v16 vmax(v16 a, v16 b) { return _mm256_max_epi16(a, b); }
static inline v16 vgt(v16 a, v16 b) { return _mm256_cmpgt_epi16(a, b); }
static inline v16 veq(v16 a, v16 b) { return _mm256_cmpeq_epi16(a, b); }
static inline v16 vand(v16 a, v16 b) { return _mm256_and_si256(a, b); }
static inline v16 vandn(v16 m, v16 a) { return _mm256_andnot_si256(m, a); }               /* ~m & a */
static inline v16 vor(v16 a, v16 b) { return _mm256_or_si256(a, b); }
static inline v16 vadd(v16 a, v16 b) { return _mm256_add_epi16(a, b); }
static inline v16 vsub(v16 a, v16 b) { return _mm256_sub_epi16(a, b); }
static inline int vany(v16 m) { return !_mm256_testz_si256(m, m); }

typedef struct {
    v16 *H, *E, *Q, *QN, *MQ;  /* per column: H(i-1, j), E(i, j), query base, "query base is N", "column inside the lane's query" */
    v16 *RM;                   /* per row: row maxima of the 16 lanes */
    int16_t *T;                /* per row: target base of every lane */
    int cap_q, cap_t;
} kws_t;

static void kws_reserve(kws_t *s, int qmax, int tmax)
{
    if (qmax + 2 > s->cap_q) {
        free(s->H); free(s->E); free(s->Q); free(s->QN); free(s->MQ);
        s->cap_q = qmax + 64;
        s->H = (v16 *)aligned_alloc(32, sizeof(v16) * (size_t)s->cap_q);
        s->E = (v16 *)aligned_alloc(32, sizeof(v16) * (size_t)s->cap_q);
        s->Q = (v16 *)aligned_alloc(32, sizeof(v16) * (size_t)s->cap_q);
        s->QN = (v16 *)aligned_alloc(32, sizeof(v16) * (size_t)s->cap_q);
        s->MQ = (v16 *)aligned_alloc(32, sizeof(v16) * (size_t)s->cap_q);
    }
    if (tmax + 2 > s->cap_t) {
        free(s->RM); free(s->T);
        s->cap_t = tmax + 64;
        s->RM = (v16 *)aligned_alloc(32, sizeof(v16) * (size_t)s->cap_t);
        s->T = (int16_t *)aligned_alloc(32, sizeof(int16_t) * LANES * (size_t)s->cap_t);
    }
}
static void kws_free(kws_t *s) { free(s->H); free(s->E); free(s->Q); free(s->QN); free(s->MQ); free(s->RM); free(s->T); }

/* one group of up to 16 pairs (ids[0..cnt)); results to res[0..cnt), cells to cells[0..cnt) */
static void kgroup16(const ora_seqpair_t *pairs, const int32_t *ids, int cnt, const uint8_t *ref, const uint8_t *qer,
                     const ora_bsw_params_t *p, int is8, int a, int b, int amb, int shift, int mx, ora_kswr_t *res, int64_t *cells, kws_t *ws)
{
    int16_t qlen[LANES], tlen[LANES], minsc[LANES], endsc[LANES];
    int qmax = 0, tmax = 0;
    for (int l = 0; l < LANES; ++l) {
        qlen[l] = tlen[l] = 0; minsc[l] = 0x7fff; endsc[l] = 0x7fff;
        if (l < cnt) {
            const ora_seqpair_t *sp = &pairs[ids[l]];
            const int xtra = sp->h0;
            qlen[l] = (int16_t)sp->len2; tlen[l] = (int16_t)sp->len1;
            const int ms = (xtra & ORA_KSW_XSUBO) ? (xtra & 0xffff) : 0x10000, es = (xtra & ORA_KSW_XSTOP) ? (xtra & 0xffff) : 0x10000;
            minsc[l] = (int16_t)(ms < 0x7fff ? ms : 0x7fff); endsc[l] = (int16_t)(es < 0x7fff ? es : 0x7fff);
            if (sp->len2 > qmax) qmax = sp->len2;
            if (sp->len1 > tmax) tmax = sp->len1;
        }
    }
    kws_reserve(ws, qmax, tmax);
    {   /* lane-interleaved sequences */
        int16_t *Q = (int16_t *)ws->Q, *QN = (int16_t *)ws->QN, *MQ = (int16_t *)ws->MQ;
        for (int j = 0; j < qmax; ++j)
            for (int l = 0; l < LANES; ++l) {
                int q = 4, in = 0;
                if (l < cnt && j < qlen[l]) { q = qer[pairs[ids[l]].idq + j]; q = q > 4 ? 4 : q; in = 1; }
                Q[j * LANES + l] = (int16_t)q; QN[j * LANES + l] = (int16_t)(q == 4 ? -1 : 0); MQ[j * LANES + l] = (int16_t)(in ? -1 : 0);
            }
        for (int i = 0; i < tmax; ++i)
            for (int l = 0; l < LANES; ++l) {
                int t = 4;
                if (l < cnt && i < tlen[l]) { t = ref[pairs[ids[l]].idr + i]; t = t > 4 ? 4 : t; }
                ws->T[i * LANES + l] = (int16_t)t;
            }
    }
    const v16 zero = _mm256_setzero_si256(), va = vset(a), vnb = vset(-b), vamb = vset(amb), four = vset(4);
    const v16 ved = vset(p->e_del), vei = vset(p->e_ins), voed = vset(p->o_del + p->e_del), voei = vset(p->o_ins + p->e_ins);
    const v16 vtlen = _mm256_loadu_si256((const __m256i *)tlen), vendm1 = vsub(_mm256_loadu_si256((const __m256i *)endsc), vset(1));
    const v16 vovf = vset(254 - shift);                       /* 8-bit rule: gmax + shift >= 255 <=> gmax > 254 - shift */
    for (int j = 0; j < qmax; ++j) { ws->H[j] = zero; ws->E[j] = zero; }
    v16 gmax = zero, te = vset(-1), qe = vset(-1), stopped = zero, rows = vtlen;
    int i;
    for (i = 0; i < tmax; ++i) {
        const v16 tv = _mm256_load_si256((const __m256i *)(ws->T + i * LANES));
        const v16 tn = veq(tv, four);
        v16 f = zero, hd = zero, imax = zero, jmax = zero;
        for (int j = 0; j < qmax; ++j) {
            const v16 s = vsel(vsel(vnb, va, veq(tv, ws->Q[j])), vamb, vor(tn, ws->QN[j]));
            v16 h = vmax(vmax(vadd(hd, s), ws->E[j]), vmax(f, zero));
            h = vand(h, ws->MQ[j]);
            hd = ws->H[j]; ws->H[j] = h;
            const v16 gt = vgt(h, imax);
            imax = vmax(imax, h); jmax = vsel(jmax, vset(j), gt);
            ws->E[j] = vmax(vmax(vsub(ws->E[j], ved), vsub(h, voed)), zero);
            f = vmax(vmax(vsub(f, vei), vsub(h, voei)), zero);
        }
        ws->RM[i] = imax;
        const v16 live = vandn(stopped, vgt(vtlen, vset(i)));
        const v16 better = vand(vgt(imax, gmax), live);
        gmax = vsel(gmax, imax, better); te = vsel(te, vset(i), better); qe = vsel(qe, jmax, better);
        v16 stop = vgt(gmax, vendm1);
        if (is8) stop = vor(stop, vgt(gmax, vovf));
        stop = vand(stop, better);
        rows = vsel(rows, vset(i + 1), stop);
        stopped = vor(stopped, stop);
        if (!vany(vandn(stopped, vgt(vtlen, vset(i + 1))))) break;
    }
    int16_t g16[LANES], te16[LANES], qe16[LANES], rw16[LANES];
    _mm256_storeu_si256((__m256i *)g16, gmax); _mm256_storeu_si256((__m256i *)te16, te);
    _mm256_storeu_si256((__m256i *)qe16, qe); _mm256_storeu_si256((__m256i *)rw16, rows);
    const int16_t *RM = (const int16_t *)ws->RM;
    for (int l = 0; l < cnt; ++l) {
        ora_kswr_t *r = &res[l];
        const int g = g16[l], ovf = is8 && g + shift >= 255;
        r->score = ovf ? 255 : g; r->te = te16[l]; r->qe = -1; r->score2 = -1; r->te2 = -1; r->tb = -1; r->qb = -1;
        cells[l] = (int64_t)rw16[l] * qlen[l];
        if (ovf) continue;
        r->qe = qlen[l] > 0 ? (g > 0 ? qe16[l] : 0) : -1;
        if (minsc[l] < 0x7fff || (pairs[ids[l]].h0 & ORA_KSW_XSUBO)) {
            /* ksw_u8's `b` list over the stored row maxima */
            const int ms = (pairs[ids[l]].h0 & ORA_KSW_XSUBO) ? (pairs[ids[l]].h0 & 0xffff) : 0x10000;
            const int d = (r->score + mx - 1) / mx, low = r->te - d, high = r->te + d;
            int n = 0, sc = 0, row = 0;
            for (int rr = 0; rr < rw16[l]; ++rr) {
                const int im = RM[rr * LANES + l];
                if (im < ms) continue;
                if (n == 0 || row + 1 != rr) {
                    if (n && (row < low || row > high) && sc > r->score2) { r->score2 = sc; r->te2 = row; }
                    n = 1; sc = im; row = rr;
                } else if (sc < im) { sc = im; row = rr; }
            }
            if (n && (row < low || row > high) && sc > r->score2) { r->score2 = sc; r->te2 = row; }
        }
    }
}

/* ------------------------------------------------------------------ batch driver */

typedef struct {
    const ora_seqpair_t *pairs; const uint8_t *ref, *qer; const ora_bsw_params_t *p; ora_kswr_t *aln;
    int is8, phase; int64_t *cells_out;
    const int32_t *order; int32_t n_vec;            /* sorted ids of the vector-eligible pairs */
    const int32_t *rest; int32_t n_rest;            /* pairs for the scalar function */
    int a, b, amb, shift, mx;
    volatile int32_t next_group, next_rest;
    int64_t total; pthread_mutex_t mu;
} kjob_t;

/* phase 0: aln[i] <- pass result; phase 1: refine aln[regid] (oracle.h, ora_kswv_batch) */
static inline void kstore(const kjob_t *jb, int32_t i, const ora_kswr_t *r)
{
    if (jb->phase == 0) jb->aln[i] = *r;
    else {
        ora_kswr_t *a = &jb->aln[jb->pairs[i].regid];
        if (a->score == r->score) { a->tb = a->te - r->te; a->qb = a->qe - r->qe; }
    }
}

static void *kworker(void *arg)
{
    kjob_t *jb = (kjob_t *)arg;
    kws_t ws; memset(&ws, 0, sizeof ws);
    int64_t acc = 0;
    const int32_t ngroups = (jb->n_vec + LANES - 1) / LANES;
    for (;;) {
        const int32_t g0 = __sync_fetch_and_add(&jb->next_group, 8);
        if (g0 >= ngroups) break;
        for (int32_t g = g0; g < g0 + 8 && g < ngroups; ++g) {
            const int32_t first = g * LANES, cnt = jb->n_vec - first < LANES ? jb->n_vec - first : LANES;
            ora_kswr_t res[LANES]; int64_t cl[LANES];
            kgroup16(jb->pairs, jb->order + first, cnt, jb->ref, jb->qer, jb->p, jb->is8, jb->a, jb->b, jb->amb, jb->shift, jb->mx, res, cl, &ws);
            for (int l = 0; l < cnt; ++l) {
                const int32_t i = jb->order[first + l];
                kstore(jb, i, &res[l]);
                if (jb->cells_out) jb->cells_out[i] = cl[l];
                acc += cl[l];
            }
        }
    }
    for (;;) {
        const int32_t r0 = __sync_fetch_and_add(&jb->next_rest, 16);
        if (r0 >= jb->n_rest) break;
        for (int32_t r = r0; r < r0 + 16 && r < jb->n_rest; ++r) {
            const int32_t i = jb->rest[r];
            const ora_seqpair_t *sp = &jb->pairs[i];
            ora_kswr_t rr; int64_t c = 0;
            ora_ksw_pass(sp->len2, jb->qer + sp->idq, sp->len1, jb->ref + sp->idr, jb->p, sp->h0, jb->is8, &rr, &c);
            kstore(jb, i, &rr);
            if (jb->cells_out) jb->cells_out[i] = c;
            acc += c;
        }
    }
    kws_free(&ws);
    pthread_mutex_lock(&jb->mu); jb->total += acc; pthread_mutex_unlock(&jb->mu);
    return 0;
}

/* LSD radix sort of ids by a 24-bit key, descending */
static void ksort_desc(int32_t *ids, const uint32_t *key, int32_t n, int32_t *tmp)
{
    for (int pass = 0; pass < 3; ++pass) {
        size_t cnt[257]; memset(cnt, 0, sizeof cnt);
        const int sh = pass * 8;
        for (int32_t i = 0; i < n; ++i) cnt[255 - ((key[ids[i]] >> sh) & 255) + 1]++;
        for (int b = 0; b < 256; ++b) cnt[b + 1] += cnt[b];
        for (int32_t i = 0; i < n; ++i) tmp[cnt[255 - ((key[ids[i]] >> sh) & 255)]++] = ids[i];
        memcpy(ids, tmp, sizeof(int32_t) * (size_t)n);
    }
}

/* same contract as ora_kswv_batch (oracle.h); identical results and cell counts.  Phase 1: the aln entries the pairs name
 * must be distinct (they are: one reverse pass per hit). */
int64_t ora_kswv_batch_avx2(const ora_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, ora_kswr_t *aln, int32_t n,
                            const ora_bsw_params_t *p, int is8, int phase, int nthreads, int64_t *cells_out)
{
    kjob_t jb; memset(&jb, 0, sizeof jb);
    int ok = 1, a = p->mat[0], b = -p->mat[1], amb = p->mat[4], shift = 127, mx = 0;
    for (int t = 0; t < 5 && ok; ++t)
        for (int q = 0; q < 5; ++q) {
            const int v = p->mat[t * 5 + q], want = (t == 4 || q == 4) ? amb : (t == q ? a : -b);
            if (v != want) { ok = 0; break; }
        }
    for (int k = 0; k < 25; ++k) { if (p->mat[k] < shift) shift = p->mat[k]; if (p->mat[k] > mx) mx = p->mat[k]; }
    shift = shift < 0 ? -shift : shift;
    if (a <= 0 || b < 0 || amb > 0 || mx <= 0 || p->e_del <= 0 || p->e_ins <= 0 || p->o_del + p->e_del > 16000 || p->o_ins + p->e_ins > 16000) ok = 0;
    int32_t *order = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1) * 3);
    int32_t *rest = order + n, *tmp = rest + n;
    uint32_t *key = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
    int32_t nv = 0, nr = 0;
    for (int32_t i = 0; i < n; ++i) {
        const ora_seqpair_t *sp = &pairs[i];
        const int mnl = sp->len1 < sp->len2 ? sp->len1 : sp->len2;
        if (ok && sp->len1 >= 0 && sp->len2 >= 0 && sp->len1 < 16000 && sp->len2 < 4000 && (long long)mnl * a < 16000) {
            order[nv++] = i;
            const uint32_t q = (uint32_t)(sp->len2 > 4095 ? 4095 : sp->len2), t = (uint32_t)((sp->len1 >> 2) > 4095 ? 4095 : (sp->len1 >> 2));
            key[i] = (q << 12) | t;
        } else rest[nr++] = i;
    }
    ksort_desc(order, key, nv, tmp);
    jb.pairs = pairs; jb.ref = ref; jb.qer = qer; jb.p = p; jb.aln = aln; jb.is8 = is8; jb.phase = phase; jb.cells_out = cells_out;
    jb.order = order; jb.n_vec = nv; jb.rest = rest; jb.n_rest = nr;
    jb.a = a; jb.b = b; jb.amb = amb; jb.shift = shift; jb.mx = mx;
    pthread_mutex_init(&jb.mu, 0);
    if (nthreads <= 1) kworker(&jb);
    else {
        pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)nthreads);
        for (int t = 0; t < nthreads; ++t) pthread_create(&th[t], 0, kworker, &jb);
        for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
        free(th);
    }
    pthread_mutex_destroy(&jb.mu);
    free(order); free(key);
    return jb.total;
}
