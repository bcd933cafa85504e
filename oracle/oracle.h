/*
 * oracle/oracle.h -- CPU restatement of the bwa-mem2 BSW / kswv hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under bwa-mem2-arm_b200/ may include,
 * link or execute this code; only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py use it, as the checker.
 *
 * PARITY UNPINNED: /root/reference ships documentation and shell scripts only
 * (SURVEY.md section 0); the C++ that implements the path lives in the
 * un-vendored fork scttfrdmn/bwa-mem2 @ arm-graviton-optimization (a fork of
 * bwa-mem2/bwa-mem2 v2.2.1) and is not on this machine.  This file restates
 * the published algorithm the reference's x86 kernels are documented to equal:
 * bwa 0.7.17 ksw.c ksw_extend2 (BSW) and ksw_u8/ksw_i16/ksw_align2 (kswv),
 * following SURVEY.md Appendix A.1 / A.2 line by line.  No golden vector of
 * the reference pins SeqPair outputs; the oracle is
 * pinned instead by closed-form known answers and by an independent
 * full-matrix Python transcription (tests/test_oracle.py).
 */
#ifndef BSW_ORACLE_H
#define BSW_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* reference: bwa-mem2 src/bandedSWA.h `typedef struct dnaSeqPair {...} SeqPair`
 * (field list SURVEY.md section 8 row a1; 14 x int32 = 56 bytes). */
typedef struct {
    int32_t idr, idq, id;
    int32_t len1, len2;
    int32_t h0;
    int32_t seqid, regid;
    int32_t score, tle, gtle, qle;
    int32_t gscore, max_off;
} ora_seqpair_t;

/* reference: bwa ksw.h `kswr_t` (7 x int32). */
typedef struct {
    int32_t score, te, qe, score2, te2, tb, qb;
} ora_kswr_t;

#define ORA_KSW_XBYTE  0x10000
#define ORA_KSW_XSTOP  0x20000
#define ORA_KSW_XSUBO  0x40000
#define ORA_KSW_XSTART 0x80000

typedef struct {
    int32_t o_del, e_del, o_ins, e_ins;
    int32_t zdrop, end_bonus;
    int8_t  mat[25];
} ora_bsw_params_t;

/* one pair: ksw_extend2 semantics. Returns score. */
int ora_bsw_extend(int qlen, const uint8_t *query, int tlen, const uint8_t *target,
                   const ora_bsw_params_t *p, int w, int h0,
                   int *qle, int *tle, int *gtle, int *gscore, int *max_off,
                   int64_t *cells);

/* batch driver with the reference call shape (getScores8/16 share it: the tier
 * only changes lane width in the reference, never the value).  Writes the six
 * result fields of pairs[0..n).  cells_out (nullable) gets per-pair DP cells;
 * returns the total.  nthreads<=1 -> single thread. */
int64_t ora_bsw_batch(ora_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer,
                      int32_t n, const ora_bsw_params_t *p, int32_t w, int nthreads,
                      int64_t *cells_out);

/* the same batch through the inter-sequence AVX2 port (bsw_avx2.c): 16 pairs per register in 16-bit lanes, sorted by
 * length, per-lane band / trimming / z-drop state -- the organisation of the reference's smithWaterman256_16.
 * Identical results; returns the total cells. */
int64_t ora_bsw_batch_avx2(ora_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer,
                           int32_t n, const ora_bsw_params_t *p, int32_t w, int nthreads);

/* one pair, one pass of ksw_u8 / ksw_i16 semantics in exact arithmetic
 * (SURVEY.md Appendix A.2). is8 selects the 8-bit overflow rule. */
void ora_ksw_pass(int qlen, const uint8_t *query, int tlen, const uint8_t *target,
                  const ora_bsw_params_t *p, int xtra, int is8,
                  ora_kswr_t *r, int64_t *cells);

/* batch driver with the kswv::getScores8/16 call shape.
 * phase 0: aln[i] <- forward pass of pairs[i] (xtra = pairs[i].h0).
 * phase 1: pairs[i] holds the reversed prefixes and h0 = KSW_XSTOP|score;
 *          the pass result rr updates aln[pairs[i].regid]: if rr.score equals
 *          its score, tb = te - rr.te and qb = qe - rr.qe. */
int64_t ora_kswv_batch(const ora_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer,
                       ora_kswr_t *aln, int32_t n, const ora_bsw_params_t *p, int is8,
                       int phase, int nthreads, int64_t *cells_out);

/* inter-sequence AVX2 port of the same (oracle/kswv_avx2.c: 16 pairs in the 16-bit lanes of a register, pairs sorted
 * by length).  Identical results and cell counts; the CPU baseline of the kswv bench. */
int64_t ora_kswv_batch_avx2(const ora_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer,
                            ora_kswr_t *aln, int32_t n, const ora_bsw_params_t *p, int is8,
                            int phase, int nthreads, int64_t *cells_out);

/* whole ksw_align2 for one pair (forward, hand-off rule, reverse), used by
 * tests to replay the caller protocol. */
void ora_ksw_align2(int qlen, const uint8_t *query, int tlen, const uint8_t *target,
                    const ora_bsw_params_t *p, int xtra, ora_kswr_t *r, int64_t *cells);

/* lane-by-lane emulation of bwa's striped ksw_u8 / ksw_i16 (oracle/striped.c): analysis tool, see its header */
void ora_ksw_pass_striped(int qlen, const uint8_t *query, int tlen, const uint8_t *target, const ora_bsw_params_t *p, int xtra,
                          int is8, ora_kswr_t *r);
void ora_kswv_batch_striped(const ora_seqpair_t *pairs, const uint8_t *ref, const uint8_t *qer, ora_kswr_t *aln, int32_t n,
                            const ora_bsw_params_t *p, int is8);

/* ---- CIGAR generation: bwa ksw.c ksw_global2, oracle/global.c ---- */
typedef struct {
    int32_t idr, idq;        /* target = ref + idr, query = qer + idq */
    int32_t tlen, qlen, w;   /* lengths and band (bwa_gen_cigar2 computes w per alignment) */
    int32_t score;           /* out */
    int32_t n_cigar;         /* out: number of ops */
    int32_t cigar_off;       /* out: first op in the dense cigar array (sum of the n_cigar before this job) */
} ora_global_job_t;

int ora_ksw_global2(int qlen, const uint8_t *query, int tlen, const uint8_t *target, const ora_bsw_params_t *p, int w,
                    int *n_cigar, uint32_t **cigar, int64_t *cells);
void wl_global_fill(uint64_t seed, int32_t n, int32_t qmin, int32_t qmax, double div, double p_n, int32_t w_opt,
                    int32_t a, int32_t o_del, int32_t e_del, int32_t o_ins, int32_t e_ins,
                    ora_global_job_t *jobs, uint8_t *ref, uint8_t *qer);
int64_t ora_global2_batch(ora_global_job_t *jobs, const uint8_t *ref, const uint8_t *qer, int32_t n,
                          const ora_bsw_params_t *p, uint32_t *cigar, int64_t cigar_cap, int nthreads);

#ifdef __cplusplus
}
#endif
#endif
