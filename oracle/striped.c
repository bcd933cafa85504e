/*
 * oracle/striped.c -- lane-by-lane emulation of bwa 0.7.17 ksw.c ksw_u8 / ksw_i16 as written: Farrar's striped query
 * layout (slen segments x p lanes, query column c in segment c % slen, lane c / slen), zero-score pad columns up to
 * slen * p, the unsigned saturating arithmetic with bias (u8) / signed saturating (i16), E taken from H BEFORE the
 * lazy-F pass, and the lazy-F loop itself.  ANALYSIS TOOL ONLY (tools/striped_probe.py): it
 * answers "on how many pairs would a faithful striped kernel differ from the exact-row statement the oracle and the
 * CUDA kernels implement".  Not part of any parity claim and never linked into the product.
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

static inline int sat_u8(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }
static inline int sat_i16(int v) { return v < -32768 ? -32768 : (v > 32767 ? 32767 : v); }

/* one pass; is8: ksw_u8 (16 lanes, biased unsigned bytes), else ksw_i16 (8 lanes, signed words) */
void ora_ksw_pass_striped(int qlen, const uint8_t *query, int tlen, const uint8_t *target, const ora_bsw_params_t *p, int xtra,
                          int is8, ora_kswr_t *r)
{
    const int8_t *mat = p->mat;
    const int P = is8 ? 16 : 8;
    const int slen = (qlen + P - 1) / P, n = slen * P;
    const int e_del = p->e_del, e_ins = p->e_ins, oe_del = p->o_del + p->e_del, oe_ins = p->o_ins + p->e_ins;
    int i, j, k, l, shift = 127, mx = 0;
    const int minsc = (xtra & ORA_KSW_XSUBO) ? (xtra & 0xffff) : 0x10000;
    const int endsc = (xtra & ORA_KSW_XSTOP) ? (xtra & 0xffff) : 0x10000;
    int gmax = 0, te = -1, n_b = 0, m_b = 0;
    int64_t *b = 0;
    /* vectors as [segment][lane] */
    int *H0 = (int *)calloc((size_t)(n > 0 ? n : 1), sizeof(int)), *H1 = (int *)calloc((size_t)(n > 0 ? n : 1), sizeof(int));
    int *E = (int *)calloc((size_t)(n > 0 ? n : 1), sizeof(int)), *Hmax = (int *)calloc((size_t)(n > 0 ? n : 1), sizeof(int));
    int *qp = (int *)malloc(sizeof(int) * 5 * (size_t)(n > 0 ? n : 1));
    for (k = 0; k < 25; ++k) { if (mat[k] < shift) shift = mat[k]; if (mat[k] > mx) mx = mat[k]; }
    shift = shift < 0 ? -shift : shift;
    if (!is8) shift = 0;
    /* ksw_qinit: profile entry of pad columns is 0 (+ shift in the byte version) */
    for (k = 0; k < 5; ++k)
        for (j = 0; j < slen; ++j)
            for (l = 0; l < P; ++l) {
                const int c = j + l * slen;
                qp[(k * slen + j) * P + l] = (c >= qlen ? 0 : mat[k * 5 + query[c]]) + shift;
            }
    r->score = 0; r->te = -1; r->qe = -1; r->score2 = -1; r->te2 = -1; r->tb = -1; r->qb = -1;
    for (i = 0; i < tlen && slen > 0; ++i) {
        int f[16], h[16], maxv[16], imax;
        const int *S = qp + (size_t)target[i] * slen * P;
        int *tmp;
        for (l = 0; l < P; ++l) { f[l] = 0; maxv[l] = 0; }
        /* h = H0[slen - 1] shifted by one lane */
        for (l = P - 1; l >= 1; --l) h[l] = H0[(slen - 1) * P + l - 1];
        h[0] = 0;
        for (j = 0; j < slen; ++j)
            for (l = 0; l < P; ++l) {
                int hv, e, t;
                if (is8) { hv = sat_u8(h[l] + S[j * P + l]); hv = sat_u8(hv - shift); }
                else hv = sat_i16(h[l] + S[j * P + l]);
                e = E[j * P + l];
                hv = hv > e ? hv : e;
                hv = hv > f[l] ? hv : f[l];
                maxv[l] = maxv[l] > hv ? maxv[l] : hv;
                H1[j * P + l] = hv;
                if (is8) { e = sat_u8(e - e_del); t = sat_u8(hv - oe_del); }
                else { e = sat_i16(e - e_del); t = sat_i16(hv - oe_del); if (e < 0) e = e; }
                e = e > t ? e : t;
                if (!is8 && e < 0) e = 0;                   /* ksw_i16 uses _mm_subs_epu16: unsigned saturation at 0 */
                E[j * P + l] = e;
                if (is8) { f[l] = sat_u8(f[l] - e_ins); t = sat_u8(hv - oe_ins); }
                else { f[l] = f[l] - e_ins; if (f[l] < 0) f[l] = 0; t = hv - oe_ins; if (t < 0) t = 0; }
                f[l] = f[l] > t ? f[l] : t;
                h[l] = H0[j * P + l];
            }
        /* lazy F */
        for (k = 0; k < P; ++k) {
            int done = 0;
            for (l = P - 1; l >= 1; --l) f[l] = f[l - 1];
            f[0] = 0;
            for (j = 0; j < slen; ++j) {
                int all = 1;
                for (l = 0; l < P; ++l) {
                    int hv = H1[j * P + l], t;
                    hv = hv > f[l] ? hv : f[l];
                    H1[j * P + l] = hv;
                    t = hv - oe_ins; if (t < 0) t = 0;
                    f[l] = f[l] - e_ins; if (f[l] < 0) f[l] = 0;
                    /* u8: cmp = (f - h' saturating == 0) in every lane; i16: !(f > h') in every lane */
                    if (f[l] > t) all = 0;
                }
                if (all) { done = 1; break; }
            }
            if (done) break;
        }
        imax = 0;
        for (l = 0; l < P; ++l) imax = imax > maxv[l] ? imax : maxv[l];
        if (imax >= minsc) {
            if (n_b == 0 || (int32_t)b[n_b - 1] + 1 != i) {
                if (n_b == m_b) { m_b = m_b ? m_b << 1 : 8; b = (int64_t *)realloc(b, 8 * (size_t)m_b); }
                b[n_b++] = (int64_t)imax << 32 | i;
            } else if ((int)(b[n_b - 1] >> 32) < imax) b[n_b - 1] = (int64_t)imax << 32 | i;
        }
        if (imax > gmax) {
            gmax = imax; te = i;
            memcpy(Hmax, H1, sizeof(int) * (size_t)n);
            if ((is8 && gmax + shift >= 255) || gmax >= endsc) break;
        }
        tmp = H1; H1 = H0; H0 = tmp;
    }
    r->score = is8 ? (gmax + shift < 255 ? gmax : 255) : gmax;
    r->te = te;
    if (!(is8 && r->score == 255)) {
        int max = -1, low, high;
        /* scan in memory order of the striped vector: index x -> column x / P + x % P * slen; smallest column wins ties */
        for (i = 0; i < n; ++i) {
            const int v = Hmax[i], c = i / P + i % P * slen;
            if (v > max) { max = v; r->qe = c; }
            else if (v == max && c < r->qe) r->qe = c;
        }
        if (b) {
            i = (r->score + mx - 1) / mx;
            low = te - i; high = te + i;
            for (i = 0; i < n_b; ++i) {
                const int e = (int32_t)b[i];
                if ((e < low || e > high) && (int)(b[i] >> 32) > r->score2) { r->score2 = (int)(b[i] >> 32); r->te2 = e; }
            }
        }
    }
    free(b); free(H0); free(H1); free(E); free(Hmax); free(qp);
}

/* batch: phase 0 of pairs[0..n) into aln (xtra = pairs[i].h0) */
void ora_kswv_batch_striped(const ora_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, ora_kswr_t *aln, int32_t n,
                            const ora_bsw_params_t *p, int is8)
{
    int32_t i;
    for (i = 0; i < n; ++i)
        ora_ksw_pass_striped(pairs[i].len2, qer + pairs[i].idq, pairs[i].len1, ref + pairs[i].idr, p, pairs[i].h0, is8, &aln[i]);
}
