/*
 * oracle/bsw_avx2.c -- inter-sequence SIMD CPU port of the BSW path (TEST / BASELINE INFRASTRUCTURE, see oracle.h;
 * never linked into the product library).
 *
 * What it stands for: the reference's x86 kernels BandedPairWiseSW::smithWaterman256_16 (bwa-mem2 v2.2.1
 * src/bandedSWA.cpp) run 16 pairs side by side in the 16-bit lanes of an AVX2
 * register, pairs sorted by length, every lane carrying its own band / beg / end / z-drop state.  Those sources are
 * not in /root/reference (docs-only snapshot), so this is a PORT of that organisation written from the algorithm
 * (oracle.c ora_bsw_extend == ksw_extend2), not the reference's code: bench.py reports it as cpu_baseline kind
 * "port-avx2".  It is bit-exact against ora_bsw_extend on every field (tests/test_oracle.py), including the exact
 * beg / end trimming and the stale rolling-array entries, so its cell count equals the oracle's.
 *
 * Lane l of a group owns one pair.  H[j] / E[j] are vectors over the 16 pairs (the oracle's eh[j].h / .e); a row
 * sweeps the columns [min beg, max end] of the group with per-lane "in range" masks; nothing outside a lane's
 * [beg, end) is modified.  Pairs whose score bound does not fit 15 bits, or a scoring matrix that is not the
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
static inline v16 vmin(v16 a, v16 b) { return _mm256_min_epi16(a, b); }
static inline v16 vgt(v16 a, v16 b) { return _mm256_cmpgt_epi16(a, b); }
static inline v16 veq(v16 a, v16 b) { return _mm256_cmpeq_epi16(a, b); }
static inline v16 vand(v16 a, v16 b) { return _mm256_and_si256(a, b); }
static inline v16 vandn(v16 m, v16 a) { return _mm256_andnot_si256(m, a); }               /* ~m & a */
static inline v16 vor(v16 a, v16 b) { return _mm256_or_si256(a, b); }
static inline v16 vadd(v16 a, v16 b) { return _mm256_add_epi16(a, b); }
static inline v16 vsub(v16 a, v16 b) { return _mm256_sub_epi16(a, b); }
static inline int vany(v16 m) { return !_mm256_testz_si256(m, m); }

static inline int hmin16(v16 v)
{
    __m128i m = _mm_min_epi16(_mm256_castsi256_si128(v), _mm256_extracti128_si256(v, 1));
    m = _mm_min_epi16(m, _mm_shuffle_epi32(m, 0x4e)); m = _mm_min_epi16(m, _mm_shuffle_epi32(m, 0xb1));
    m = _mm_min_epi16(m, _mm_shufflelo_epi16(m, 0xb1));
    return (short)_mm_extract_epi16(m, 0);
}
static inline int hmax16(v16 v)
{
    __m128i m = _mm_max_epi16(_mm256_castsi256_si128(v), _mm256_extracti128_si256(v, 1));
    m = _mm_max_epi16(m, _mm_shuffle_epi32(m, 0x4e)); m = _mm_max_epi16(m, _mm_shuffle_epi32(m, 0xb1));
    m = _mm_max_epi16(m, _mm_shufflelo_epi16(m, 0xb1));
    return (short)_mm_extract_epi16(m, 0);
}

typedef struct {
    v16 *H, *E, *Q;            /* per column: eh[j].h, eh[j].e, query base (16 lanes each) */
    int16_t *T;                /* per row: target base of every lane */
    int cap_q, cap_t;
} ws_t;

static void ws_reserve(ws_t *s, int qmax, int tmax)
{
    if (qmax + 2 > s->cap_q) {
        free(s->H); free(s->E); free(s->Q);
        s->cap_q = qmax + 64;
        s->H = (v16 *)aligned_alloc(32, sizeof(v16) * (size_t)s->cap_q);
        s->E = (v16 *)aligned_alloc(32, sizeof(v16) * (size_t)s->cap_q);
        s->Q = (v16 *)aligned_alloc(32, sizeof(v16) * (size_t)s->cap_q);
    }
    if (tmax + 1 > s->cap_t) {
        free(s->T);
        s->cap_t = tmax + 64;
        s->T = (int16_t *)aligned_alloc(32, sizeof(int16_t) * LANES * (size_t)s->cap_t);
    }
}
static void ws_free(ws_t *s) { free(s->H); free(s->E); free(s->Q); free(s->T); memset(s, 0, sizeof *s); }

static int band_of(const ora_bsw_params_t *p, int qlen, int w, int maxsc)
{
    int max_ins = (int)((double)(qlen * maxsc + p->end_bonus - p->o_ins) / p->e_ins + 1.);
    int max_del = (int)((double)(qlen * maxsc + p->end_bonus - p->o_del) / p->e_del + 1.);
    max_ins = max_ins > 1 ? max_ins : 1; w = w < max_ins ? w : max_ins;
    max_del = max_del > 1 ? max_del : 1; w = w < max_del ? w : max_del;
    return w;
}

/* one group of up to 16 pairs (ids[0..cnt)); returns the DP cells */
static int64_t group16(ora_seqpair_t *pairs, const int32_t *ids, int cnt, const uint8_t *ref, const uint8_t *qer,
                       const ora_bsw_params_t *p, int w, int sc_a, int sc_b, int sc_amb, ws_t *ws)
{
    int16_t qlen[LANES], tlen[LANES], h0[LANES], wb[LANES];
    int l, i, j, qmax = 0, tmax = 0;
    const int oe_del = p->o_del + p->e_del, oe_ins = p->o_ins + p->e_ins;
    for (l = 0; l < LANES; ++l) {
        if (l < cnt) {
            const ora_seqpair_t *sp = &pairs[ids[l]];
            qlen[l] = (int16_t)sp->len2; tlen[l] = (int16_t)sp->len1; h0[l] = (int16_t)sp->h0;
            wb[l] = (int16_t)band_of(p, sp->len2, w, sc_a > sc_amb ? sc_a : sc_amb);
        } else { qlen[l] = tlen[l] = 0; h0[l] = 0; wb[l] = 0; }
        qmax = qmax > qlen[l] ? qmax : qlen[l]; tmax = tmax > tlen[l] ? tmax : tlen[l];
    }
    ws_reserve(ws, qmax, tmax);
    /* sequences, lane-interleaved (the reference's seqBufRef / seqBufQer SoA transposition) */
    for (j = 0; j < qmax; ++j) {
        int16_t *q = (int16_t *)&ws->Q[j];
        for (l = 0; l < LANES; ++l) q[l] = j < qlen[l] ? (int16_t)qer[pairs[ids[l]].idq + j] : 4;
    }
    for (i = 0; i < tmax; ++i)
        for (l = 0; l < LANES; ++l) ws->T[i * LANES + l] = i < tlen[l] ? (int16_t)ref[pairs[ids[l]].idr + i] : 4;
    v16 *H = ws->H, *E = ws->E;
    const v16 zero = _mm256_setzero_si256(), ones = _mm256_set1_epi16(-1);
    const v16 vqlen = _mm256_loadu_si256((const __m256i *)qlen), vtlen = _mm256_loadu_si256((const __m256i *)tlen);
    const v16 vh0 = _mm256_loadu_si256((const __m256i *)h0), vw = _mm256_loadu_si256((const __m256i *)wb);
    const v16 v_oe_del = vset(oe_del), v_oe_ins = vset(oe_ins), v_e_del = vset(p->e_del), v_e_ins = vset(p->e_ins);
    const v16 v_a = vset(sc_a), v_nb = vset(-sc_b), v_amb = vset(sc_amb), v3 = vset(3);
    /* first row */
    {
        v16 prev = vh0;
        H[0] = vh0; E[0] = zero;
        for (j = 1; j <= qmax + 1; ++j) {
            v16 J = vset(j), cur;
            if (j == 1) cur = vand(vsub(vh0, v_oe_ins), vgt(vh0, v_oe_ins));
            else cur = vand(vsub(prev, v_e_ins), vgt(prev, v_e_ins));
            cur = vandn(vgt(J, vqlen), cur);                    /* nothing right of the query */
            H[j] = cur; E[j] = zero; prev = cur;
        }
    }
    v16 beg = zero, end = vqlen, vmaxs = vh0, max_i = ones, max_j = ones, max_ie = ones, gscore = ones, max_off = zero;
    v16 act = vgt(vtlen, zero);
    {   /* lanes beyond cnt are dead */
        int16_t a[LANES]; for (l = 0; l < LANES; ++l) a[l] = l < cnt ? -1 : 0;
        act = vand(act, _mm256_loadu_si256((const __m256i *)a));
    }
    int64_t cells = 0;
    for (i = 0; i < tmax && vany(act); ++i) {
        const v16 I = vset(i), Trow = _mm256_loadu_si256((const __m256i *)&ws->T[i * LANES]);
        act = vand(act, vgt(vtlen, I));
        if (!vany(act)) break;
        /* band: beg = max(beg, i - w); end = min(end, i + w + 1, qlen) */
        beg = vsel(beg, vmax(beg, vsub(I, vw)), act);
        end = vsel(end, vmin(vmin(end, vadd(vadd(I, vw), vset(1))), vqlen), act);
        int dec = p->o_del + p->e_del * (i + 1); if (dec > 32767) dec = 32767;
        v16 h1 = vand(vmax(_mm256_subs_epi16(vh0, vset(dec)), zero), veq(beg, zero));
        v16 f = zero, mrow = zero, mj = ones;
        const v16 nonempty = vand(act, vgt(end, beg));
        {   /* cells += end - beg over the live lanes */
            v16 d = vand(vsub(end, beg), nonempty);
            __m256i s32 = _mm256_madd_epi16(d, vset(1));
            __m128i s = _mm_add_epi32(_mm256_castsi256_si128(s32), _mm256_extracti128_si256(s32, 1));
            s = _mm_add_epi32(s, _mm_shuffle_epi32(s, 0x4e)); s = _mm_add_epi32(s, _mm_shuffle_epi32(s, 0xb1));
            cells += _mm_cvtsi128_si32(s);
        }
        v16 found = zero, nb = end, lastnz = vsub(end, vset(1));
        if (vany(nonempty)) {
            const int jlo = hmin16(vsel(vset(32767), beg, nonempty)), jhi = hmax16(vsel(vset(-1), end, nonempty));
            for (j = jlo; j <= jhi; ++j) {
                const v16 J = vset(j);
                const v16 inr = vand(vandn(vgt(beg, J), vgt(end, J)), nonempty);     /* beg <= j < end */
                const v16 ate = vand(veq(end, J), nonempty);
                const v16 Hj = H[j], Ej = E[j], Qj = ws->Q[j < qmax ? j : 0];
                /* score of (target base, query base): a / -b / ambiguous */
                v16 s = vsel(v_nb, v_a, veq(Trow, Qj));
                s = vsel(s, v_amb, vgt(vmax(Trow, Qj), v3));
                const v16 M = vand(vadd(Hj, s), vgt(Hj, zero));                       /* M = M ? M + s : 0 */
                const v16 h = vmax(vmax(M, Ej), f);
                const v16 newH = vsel(Hj, h1, vor(inr, ate));                           /* eh[j].h = h1 (also eh[end].h) */
                h1 = vsel(h1, h, inr);
                mj = vsel(mj, J, vandn(vgt(mrow, h), inr));                              /* ties -> last j */
                mrow = vsel(mrow, vmax(mrow, h), inr);
                const v16 e = vmax(vsub(Ej, v_e_del), vmax(vsub(M, v_oe_del), zero));
                const v16 newE = vandn(ate, vsel(Ej, e, inr));                           /* eh[end].e = 0 */
                f = vsel(f, vmax(vsub(f, v_e_ins), vmax(vsub(M, v_oe_ins), zero)), inr);
                H[j] = newH; E[j] = newE;
                /* trimming: first / last non-zero entry among the columns this row wrote */
                const v16 nz = vandn(veq(vor(newH, newE), zero), vor(inr, ate));
                nb = vsel(nb, J, vandn(found, vand(nz, inr)));
                found = vor(found, vand(nz, inr));
                lastnz = vsel(lastnz, J, nz);
            }
        }
        /* an empty row ends its lane: j == beg after the oracle's loop */
        const v16 jfin = vsel(beg, end, vgt(end, beg));
        const v16 gsc = vand(act, veq(jfin, vqlen));
        max_ie = vsel(max_ie, I, vandn(vgt(gscore, h1), gsc));
        gscore = vsel(gscore, vmax(gscore, h1), gsc);
        v16 brk = vand(act, veq(mrow, zero));
        const v16 gt = vandn(brk, vand(act, vgt(mrow, vmaxs)));
        {
            v16 off = _mm256_abs_epi16(vsub(mj, I));
            max_off = vsel(max_off, vmax(max_off, off), gt);
        }
        if (p->zdrop > 0) {
            /* lanes that did not improve: z-drop test with the diagonal offset (exact in 32 bits when it matters: a product
             * that leaves 15 bits makes the left side negative, i.e. no break) */
            const v16 zl = vandn(vor(brk, gt), act);
            if (vany(zl)) {
                const v16 di = vsub(I, max_i), dj = vsub(mj, max_j);
                const v16 c1 = vgt(di, dj);
                const v16 d = _mm256_abs_epi16(vsub(di, dj));
                const v16 ee = vsel(v_e_ins, v_e_del, c1);
                const v16 lo = _mm256_mullo_epi16(d, ee), hi = _mm256_mulhi_epi16(d, ee);
                const v16 ovf = vor(vgt(hi, zero), vgt(zero, lo));
                const v16 lhs = _mm256_subs_epi16(vsub(vmaxs, mrow), lo);
                brk = vor(brk, vand(zl, vandn(ovf, vgt(lhs, vset(p->zdrop)))));
            }
        }
        vmaxs = vsel(vmaxs, mrow, gt); max_i = vsel(max_i, I, gt); max_j = vsel(max_j, mj, gt);
        act = vandn(brk, act);
        /* trimming for the lanes that go on */
        beg = vsel(beg, vsel(end, nb, found), act);
        end = vsel(end, vmin(vadd(lastnz, vset(2)), vqlen), act);
    }
    {
        int16_t o_max[LANES], o_i[LANES], o_j[LANES], o_ie[LANES], o_g[LANES], o_off[LANES];
        _mm256_storeu_si256((__m256i *)o_max, vmaxs); _mm256_storeu_si256((__m256i *)o_i, max_i);
        _mm256_storeu_si256((__m256i *)o_j, max_j); _mm256_storeu_si256((__m256i *)o_ie, max_ie);
        _mm256_storeu_si256((__m256i *)o_g, gscore); _mm256_storeu_si256((__m256i *)o_off, max_off);
        for (l = 0; l < cnt; ++l) {
            ora_seqpair_t *sp = &pairs[ids[l]];
            sp->score = o_max[l]; sp->tle = o_i[l] + 1; sp->qle = o_j[l] + 1; sp->gtle = o_ie[l] + 1;
            sp->gscore = o_g[l]; sp->max_off = o_off[l];
        }
    }
    return cells;
}

/* ------------------------------------------------------------------ batch driver */

typedef struct {
    ora_seqpair_t *pairs; const uint8_t *ref, *qer; const ora_bsw_params_t *p; int32_t w;
    const int32_t *order; int32_t n_vec;            /* sorted ids of the vector-eligible pairs */
    const int32_t *rest; int32_t n_rest;            /* pairs for the scalar function */
    int sc_a, sc_b, sc_amb;
    volatile int32_t next_group, next_rest;
    int64_t total; pthread_mutex_t mu;
} vjob_t;

static void *vworker(void *arg)
{
    vjob_t *jb = (vjob_t *)arg;
    ws_t ws; memset(&ws, 0, sizeof ws);
    int64_t acc = 0;
    const int32_t ngroups = (jb->n_vec + LANES - 1) / LANES;
    for (;;) {
        const int32_t g0 = __sync_fetch_and_add(&jb->next_group, 16);
        if (g0 >= ngroups) break;
        for (int32_t g = g0; g < g0 + 16 && g < ngroups; ++g) {
            const int32_t first = g * LANES, cnt = jb->n_vec - first < LANES ? jb->n_vec - first : LANES;
            acc += group16(jb->pairs, jb->order + first, cnt, jb->ref, jb->qer, jb->p, jb->w, jb->sc_a, jb->sc_b, jb->sc_amb, &ws);
        }
    }
    for (;;) {
        const int32_t r0 = __sync_fetch_and_add(&jb->next_rest, 64);
        if (r0 >= jb->n_rest) break;
        for (int32_t r = r0; r < r0 + 64 && r < jb->n_rest; ++r) {
            ora_seqpair_t *sp = &jb->pairs[jb->rest[r]];
            int qle, tle, gtle, gscore, max_off; int64_t c = 0;
            sp->score = ora_bsw_extend(sp->len2, jb->qer + sp->idq, sp->len1, jb->ref + sp->idr, jb->p, jb->w, sp->h0,
                                       &qle, &tle, &gtle, &gscore, &max_off, &c);
            sp->qle = qle; sp->tle = tle; sp->gtle = gtle; sp->gscore = gscore; sp->max_off = max_off;
            acc += c;
        }
    }
    ws_free(&ws);
    pthread_mutex_lock(&jb->mu); jb->total += acc; pthread_mutex_unlock(&jb->mu);
    return 0;
}

/* LSD radix sort of ids by key (16 bits), descending */
static void sort_by_key_desc(int32_t *ids, const uint32_t *key, int32_t n, int32_t *tmp)
{
    for (int pass = 0; pass < 2; ++pass) {
        size_t cnt[257]; memset(cnt, 0, sizeof cnt);
        const int sh = pass * 8;
        for (int32_t i = 0; i < n; ++i) cnt[255 - ((key[ids[i]] >> sh) & 255) + 1]++;
        for (int b = 0; b < 256; ++b) cnt[b + 1] += cnt[b];
        for (int32_t i = 0; i < n; ++i) tmp[cnt[255 - ((key[ids[i]] >> sh) & 255)]++] = ids[i];
        memcpy(ids, tmp, sizeof(int32_t) * (size_t)n);
    }
}

/* same contract as ora_bsw_batch (oracle.h); the six result fields are identical to it */
int64_t ora_bsw_batch_avx2(ora_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, int32_t n,
                           const ora_bsw_params_t *p, int32_t w, int nthreads)
{
    vjob_t jb; memset(&jb, 0, sizeof jb);
    /* the matrix must be bwa_fill_scmat's: a on the diagonal, one mismatch value, one value in row / column 4 */
    int ok = 1, a = p->mat[0], b = -p->mat[1], amb = p->mat[4];
    for (int t = 0; t < 5 && ok; ++t)
        for (int q = 0; q < 5; ++q) {
            const int v = p->mat[t * 5 + q], want = (t == 4 || q == 4) ? amb : (t == q ? a : -b);
            if (v != want) { ok = 0; break; }
        }
    if (a <= 0 || b < 0 || p->e_del <= 0 || p->e_ins <= 0 || p->e_del > 127 || p->e_ins > 127) ok = 0;
    int32_t *order = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1) * 3);
    int32_t *rest = order + n, *tmp = rest + n;
    uint32_t *key = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
    int32_t nv = 0, nr = 0;
    for (int32_t i = 0; i < n; ++i) {
        const ora_seqpair_t *sp = &pairs[i];
        const int mn = sp->len1 < sp->len2 ? sp->len1 : sp->len2;
        const long long bound = (long long)sp->h0 + (long long)mn * a;
        if (ok && sp->len1 >= 0 && sp->len2 >= 0 && sp->len1 < 32000 && sp->len2 < 32000 && sp->h0 >= 0 && bound < 32000
            && p->o_del + p->e_del < 32000 && p->o_ins + p->e_ins < 32000) {
            order[nv++] = i;
            /* query length major, target length minor: lanes of a group end and trim alike */
            const uint32_t q = (uint32_t)(sp->len2 > 255 ? 255 : sp->len2), t = (uint32_t)((sp->len1 >> 1) > 255 ? 255 : (sp->len1 >> 1));
            key[i] = (q << 8) | t;
        } else rest[nr++] = i;
    }
    sort_by_key_desc(order, key, nv, tmp);
    jb.pairs = pairs; jb.ref = ref; jb.qer = qer; jb.p = p; jb.w = w; jb.order = order; jb.n_vec = nv; jb.rest = rest; jb.n_rest = nr;
    jb.sc_a = a; jb.sc_b = b; jb.sc_amb = amb;
    pthread_mutex_init(&jb.mu, 0);
    if (nthreads <= 1) vworker(&jb);
    else {
        pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)nthreads);
        for (int t = 0; t < nthreads; ++t) pthread_create(&th[t], 0, vworker, &jb);
        for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
        free(th);
    }
    pthread_mutex_destroy(&jb.mu);
    free(order); free(key);
    return jb.total;
}
