/*
 * include/bsw_cuda.h -- C ABI of the B200 (sm_100a) batched banded Smith-Waterman
 * seed-extension path and the kswv batched mate-rescue path.
 *
 * This is the drop-in boundary: every entry point keeps
 * the argument list of the bwa-mem2 method it replaces, with the object turned
 * into an opaque handle.  Plain pointers and sizes only.  INTEGRATION.md shows
 * the C++ shim classes (`BandedPairWiseSW`, `kswv`) a maintainer drops into
 * bwa-mem2 so that bwamem.cpp / bwamem_pair.cpp compile unchanged; the same
 * classes ship in bwa-mem2-arm_b200/bandedSWA_b200.h.
 *
 * Reference interfaces replaced ("R:" = /root/reference, which quotes the
 * signatures; the code itself is bwa-mem2 v2.2.1 src/bandedSWA.{h,cpp},
 * src/kswv.{h,cpp}):
 *   BandedPairWiseSW ctor   R:docs-archive/INTEGRATION_COMPLETE.md:61-63
 *   getScores8 / getScores16 R:docs-archive/WEEK1_WRAPPER_COMPLETE.md:263-268
 *                            R:docs-archive/WEEK2_STATUS.md:143-156,209-211
 *   kswv ctor / getScores8/16(…, kswr_t* aln, …, phase)
 *                            R:docs-archive/INTEGRATION_COMPLETE.md:57-58,69-71,95-97
 *                            R:docs-archive/WEEK2_STATUS.md:94
 *   SeqPair field use        R:PHASE2_COMPLETE.md:138-140, R:docs-archive/WEEK2_STATUS.md:49-53,159
 *   h0 carries `xtra` for kswv  R:docs-archive/WEEK1_BENCHMARK_RESULTS.md:97-103
 *
 * Error convention = the reference's (R:docs-archive/WEEK1_WRAPPER_COMPLETE.md:239-242):
 * void returns; on a CUDA failure or a contract violation the library prints a
 * message to stderr and calls exit(EXIT_FAILURE).  There is NO CPU fallback:
 * without a usable CUDA device every compute entry point fails loudly.
 */
#ifndef BSW_CUDA_H
#define BSW_CUDA_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* bwa-mem2 src/bandedSWA.h `typedef struct dnaSeqPair {...} SeqPair;`  14 x int32 = 56 B.
 * Inputs : idr, idq (byte offsets into seqBufRef / seqBufQer), len1 (target), len2 (query),
 *          h0 (BSW: initial score > 0;  kswv: the `xtra` flag word).
 * Outputs: score, tle, gtle, qle, gscore, max_off (BSW only).
 * id, seqid, regid are caller bookkeeping and are preserved; kswv phase 1 reads
 * regid as the index of the kswr_t entry the pair refines. */
typedef struct bsw_seqpair {
    int32_t idr, idq, id;
    int32_t len1, len2;
    int32_t h0;
    int32_t seqid, regid;
    int32_t score, tle, gtle, qle;
    int32_t gscore, max_off;
} bsw_seqpair_t;

/* bwa ksw.h `kswr_t`: 7 x int32 = 28 B */
typedef struct bsw_kswr {
    int32_t score;
    int32_t te, qe;
    int32_t score2, te2;
    int32_t tb, qb;
} bsw_kswr_t;

#define BSW_KSW_XBYTE  0x10000
#define BSW_KSW_XSTOP  0x20000
#define BSW_KSW_XSUBO  0x40000
#define BSW_KSW_XSTART 0x80000

/* tier limits of the reference (bandedSWA.h MAX_SEQ_LEN8 / MAX_SEQ_LEN16) */
#define BSW_MAX_SEQ_LEN8   128
#define BSW_MAX_SEQ_LEN16  32768

typedef struct bsw_handle  bsw_t;
typedef struct kswv_handle kswv_t;

/* ---- BandedPairWiseSW ------------------------------------------------------
 * replaces BandedPairWiseSW::BandedPairWiseSW(o_del,e_del,o_ins,e_ins,zdrop,
 *          end_bonus,mat,w_match,w_mismatch,numThreads) and its destructor.
 * w_mismatch may be given as the penalty (upstream passes opt->b) or already
 * negated (the fork's ARM call sites pass -opt->b); the ambiguous-base score is
 * the reference's DEFAULT_AMBIG = -1.  `mat` (5x5) is accepted for signature
 * compatibility; like the reference's vector kernels, scoring uses
 * w_match / w_mismatch / ambig.
 * The handle binds to the CUDA device(s) named by BSW_CUDA_DEVICES (comma list,
 * default: LOCAL_RANK if set, else device 0).  With several devices a large call
 * deals its chunks (BSW_CUDA_CHUNK_PAIRS pairs) round-robin to them and a small call
 * goes to the device whose engine has the fewest pairs outstanding.
 * Creating and destroying a handle is cheap (the reference builds its objects per
 * task and per worker thread): streams, pinned mailboxes and device scratch are
 * process-wide per device and pooled. */
bsw_t *bsw_create(int o_del, int e_del, int o_ins, int e_ins, int zdrop, int end_bonus,
                  const int8_t *mat, int8_t w_match, int8_t w_mismatch, int numThreads);
void   bsw_destroy(bsw_t *h);

/* replaces BandedPairWiseSW::getScores8 / getScores16 (same argument order).
 * Host pointers (pageable or pinned).  Writes score,tle,gtle,qle,gscore,max_off of
 * pairArray[0..numPairs) in place; synchronous; re-entrant across handles.
 * getScores8 enforces the reference's 8-bit contract (len1,len2 < 128); both tiers
 * compute in packed signed 16-bit lanes, which is value-identical on admitted
 * pairs (SURVEY.md A.1 note 2).
 * Calls below BSW_CUDA_DIRECT_MIN pairs (default 131072) -- the size the reference's
 * worker threads issue, R:docs-archive/WEEK1_BENCHMARK_RESULTS.md:88-95 -- are merged
 * with the concurrent calls of the other host threads into one device batch by a
 * per-device engine thread; the calling thread itself makes no CUDA call.  Larger
 * calls are cut into chunks and DMA'd straight from the caller's buffers. */
void bsw_getScores8 (bsw_t *h, bsw_seqpair_t *pairArray, const uint8_t *seqBufRef,
                     const uint8_t *seqBufQer, int32_t numPairs, uint16_t numThreads, int32_t w);
void bsw_getScores16(bsw_t *h, bsw_seqpair_t *pairArray, const uint8_t *seqBufRef,
                     const uint8_t *seqBufQer, int32_t numPairs, uint16_t numThreads, int32_t w);
/* replaces BandedPairWiseSW::scalarBandedSWAWrapper (pairs >= 32768 bases): same
 * device path, no length limit other than memory. */
void bsw_scalarBandedSWAWrapper(bsw_t *h, bsw_seqpair_t *pairArray, const uint8_t *seqBufRef,
                                const uint8_t *seqBufQer, int32_t numPairs, int nthreads, int32_t w);

/* ---- sequence arenas (optional) ----------------------------------------------
 * bwa-mem2 passes ONE pair of sequence buffers to the getScores8 and getScores16
 * calls of a batch and to both attempts of the band-retry loop (bwamem.cpp,
 * mem_chain2aln_across_reads_V2).  Declaring them lets the large-call path upload
 * them once: between bsw_arena_begin and bsw_arena_end every call on this handle
 * whose buffers lie inside [seqBufRef, +refBytes) / [seqBufQer, +qerBytes) reads the
 * device copy made by bsw_arena_begin (asynchronous upload) and sends only its
 * SeqPair records.  The caller must not modify the arenas in between.  Small
 * (staged) calls ignore the declaration. */
void bsw_arena_begin(bsw_t *h, const uint8_t *seqBufRef, size_t refBytes, const uint8_t *seqBufQer, size_t qerBytes);
void bsw_arena_end(bsw_t *h);
/* Deferred execution inside an arena bracket (opt-in, off by default).  mem_chain2aln_across_reads_V2 reads no result
 * between the getScores8 and the getScores16 call of a side (bwamem.cpp; it reads them once both tiers are back, to chain
 * h0 and to pick the band-retry pairs).  With on != 0 a LARGE call (direct route) between bsw_arena_begin and
 * bsw_arena_end only records its arguments and returns; bsw_arena_end runs all recorded calls of the bracket as ONE
 * interleaved chunk pipeline -- the call that mostly waits for PCIe (it brings the arenas up) and the one that mostly
 * waits for the GPU fill each other's gaps -- and returns when every result has landed in its SeqPair array.  The
 * arrays and buffers must stay valid and untouched until then.  Small (staged) calls are not deferred.  Results are
 * identical. */
void bsw_arena_defer(bsw_t *h, int on);

/* ---- resident reference (no reference counterpart) -----------
 * The host pipeline copies every target slice out of the 2-bit packed reference
 * into seqBufRef (bns_fetch_seq / bns_get_seq) only to send it to the callee.
 * bsw_ref_upload keeps the packed reference (bwa's .pac layout: base x in byte
 * x >> 2, bits ((~x & 3) << 1), l_pac bases) on every device of BSW_CUDA_DEVICES;
 * bsw_getScoresRef then takes, instead of target bases, one coordinate per pair in
 * bwa's doubled coordinate system (0 <= x < l_pac forward strand, l_pac <= x < 2 l_pac
 * reverse complement): rpos[i] >= 0 -> target[k] = base(rpos[i] + k) (right
 * extensions); rpos[i] < 0 -> target[k] = base(~rpos[i] - k) (the reversed slices of
 * left extensions).  pairArray[i].idr is ignored and preserved; everything else
 * as bsw_getScores16 (no 8-bit length contract). */
typedef struct bsw_ref bsw_ref_t;
bsw_ref_t *bsw_ref_upload(const uint8_t *pac, int64_t l_pac);
void       bsw_ref_free(bsw_ref_t *r);
void bsw_getScoresRef(bsw_t *h, bsw_seqpair_t *pairArray, const bsw_ref_t *ref, const int64_t *rpos,
                      const uint8_t *seqBufQer, int32_t numPairs, int32_t w);

/* ---- kswv ------------------------------------------------------------------
 * replaces kswv::kswv(o_del,e_del,o_ins,e_ins,w_match,w_mismatch,numThreads,
 *          maxRefLen,maxQerLen), ~kswv and kswv::getScores8/16(pairs,ref,qer,aln,
 *          numPairs,numThreads,phase).
 * phase 0: aln[i] <- {score, te, qe, score2, te2, tb=-1, qb=-1} of pairArray[i],
 *          xtra = pairArray[i].h0.
 * phase 1: pairArray[i] holds the reversed prefixes with h0 = KSW_XSTOP|score;
 *          updates aln[pairArray[i].regid].{tb,qb} when the reverse score matches. */
kswv_t *kswv_create(int o_del, int e_del, int o_ins, int e_ins, int8_t w_match, int8_t w_mismatch,
                    int numThreads, int32_t maxRefLen, int32_t maxQerLen);
void    kswv_destroy(kswv_t *h);
void kswv_getScores8 (kswv_t *h, bsw_seqpair_t *pairArray, const uint8_t *seqBufRef,
                      const uint8_t *seqBufQer, bsw_kswr_t *aln, int32_t numPairs,
                      uint16_t numThreads, int phase);
void kswv_getScores16(kswv_t *h, bsw_seqpair_t *pairArray, const uint8_t *seqBufRef,
                      const uint8_t *seqBufQer, bsw_kswr_t *aln, int32_t numPairs,
                      uint16_t numThreads, int phase);

/* ---- device-resident entry points (no reference counterpart) ----------------
 * Same computation on buffers that already live in the HBM of the handle's
 * first device; asynchronous on `stream` (a cudaStream_t, 0 = default stream).
 * Used by bench.py for the inputs-resident number and by pipelines that keep
 * read batches on the GPU.  The sequence buffers must be readable for 16 bytes
 * past the last base any pair names (the kernels fetch bases as aligned words).
 * The call blocks the host only for its 200-byte histogram; the DP kernels are
 * left in flight.  Calls issued on one stream are ordered by that stream; BSW
 * calls issued on DIFFERENT streams (from one host thread per handle) overlap on
 * the device -- each takes the next of the handle's four scratch slots, and a
 * slot is reused only after the call that last used it has finished.  Results
 * of a call are visible to work queued on its stream after the call returns. */
void bsw_dev_getScores(bsw_t *h, bsw_seqpair_t *d_pairArray, const uint8_t *d_seqBufRef,
                       const uint8_t *d_seqBufQer, int32_t numPairs, int32_t w, void *stream);

/* bsw_dev_getScores without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same computation and in-place results as bsw_dev_getScores on resident buffers of the handle's first
 * device, by the 32-bit kernel (one thread per pair) whose DP rows -- numPairs x (maxQueryLen + 1) x 8 bytes -- live in
 * a file-scope `static __device__` array of BSW_STATIC_SCRATCH_BYTES instead of a grow-on-demand buffer: every GPU
 * that loads the library holds those 8 MiB (+ 256 B) until the process exits.  The call makes no cudaMalloc / cudaFree
 * / cudaHostAlloc and uses no allocating library routine; the only set-up is one cudaGetSymbolAddress per device, made
 * by the first call on that device.
 *   limit:   numPairs * (maxQueryLen + 1) <= BSW_STATIC_MAX_ENTRIES (e.g. 6898 pairs of 151-base reads, 4096 of <= 255);
 *            maxQueryLen >= every pair's len2.  Larger requests are refused: nothing is launched, -2 is returned.
 *   returns: 0 = done; -1 = invalid argument; -2 = beyond the static scratch; -3 = a CUDA call failed;
 *            -4 = some pair has len2 > maxQueryLen or a negative length (those pairs get all six result words = -1,
 *            the others their results).  Every failure also prints one line to stderr.
 *   sharing: ALL calls on a device use the one array (the rows are rewritten from scratch by each call, the status word
 *            in its header is cleared by each call), so calls that overlap on the GPU -- different streams, different
 *            host threads or handles -- race.  Issue them on one stream, or order them yourself.
 *   sync:    queued on `stream`; the host blocks for the 4-byte status word (like bsw_dev_getScores for its histogram),
 *            so the results are complete when the call returns.  On a stream that is being captured into a CUDA graph the
 *            memset + kernel are recorded, nothing is waited for and 0 is returned (make one ordinary call on the device
 *            first: the per-device set-up must not happen under capture); refused pairs then show only as -1 results. */
#define BSW_STATIC_MAX_ENTRIES   (1 << 20)
#define BSW_STATIC_SCRATCH_BYTES (256 + 8 * (size_t)BSW_STATIC_MAX_ENTRIES)
int bsw_dev_getScores_static(bsw_t *h, bsw_seqpair_t *d_pairArray, const uint8_t *d_seqBufRef,
                             const uint8_t *d_seqBufQer, int32_t numPairs, int32_t maxQueryLen, int32_t w, void *stream);

/* Caller protocol of the extension step on resident pairs, one submission
 * (replaces the host loop around bswLeft/bswRight in
 * mem_chain2aln_across_reads_V2, bwamem.cpp ~L2200-2900, anchor BUG_REPORT.md:28):
 *   left:  attempt 0 with band w on all numLeft pairs; then, for attempt = 1 ..
 *          maxBandTry-1 (upstream MAX_BAND_TRY = 2), the pairs whose score changed
 *          in the previous attempt and whose max_off >= (w' >> 1) + (w' >> 2)
 *          (w' = that attempt's band) are re-extended from scratch with band
 *          w << attempt;
 *   chain: d_right[i].h0 = d_left[d_right[i].regid].score  (regid < 0: h0 kept);
 *   right: the same attempts on the numRight pairs.
 * Both tiers (8/16-bit) share one arithmetic here, so there is no tier routing.
 * Results in place, as bsw_dev_getScores; the host blocks for one histogram per
 * BSW launch and one 4-byte count per retry; returns with the last kernels queued
 * on `stream`. */
void bsw_dev_extendChained(bsw_t *h, bsw_seqpair_t *d_left, const uint8_t *d_leftRef, const uint8_t *d_leftQer,
                           int32_t numLeft, bsw_seqpair_t *d_right, const uint8_t *d_rightRef,
                           const uint8_t *d_rightQer, int32_t numRight, int32_t w, int32_t maxBandTry, void *stream);

void kswv_dev_getScores(kswv_t *h, const bsw_seqpair_t *d_pairArray, const uint8_t *d_seqBufRef,
                        const uint8_t *d_seqBufQer, bsw_kswr_t *d_aln, int32_t numPairs,
                        int is8, int phase, void *stream);

/* ksw_align2 for a resident batch in ONE submission (replaces mem_sam_pe_batch's sequence getScores(phase 0) -> host
 * hand-off -> getScores(phase 1), bwa-mem2 src/bwamem_pair.cpp): forward pass into d_aln, hand-off on
 * the device (hits with KSW_XSTART and not (KSW_XSUBO and score < minsc) get their prefixes target[0..te], query[0..qe]
 * reversed and a KSW_XSTOP | score record), reverse pass refining d_aln[i].tb / .qb in place.  d_pairArray and the
 * sequence buffers are read only.  Blocks the host for three small read-backs; returns with the last kernels queued. */
void kswv_dev_align2(kswv_t *h, const bsw_seqpair_t *d_pairArray, const uint8_t *d_seqBufRef,
                     const uint8_t *d_seqBufQer, bsw_kswr_t *d_aln, int32_t numPairs, int is8, void *stream);

/* The same on HOST buffers, synchronous: one call instead of mem_sam_pe_batch's getScores8/16(phase 0), the host loop
 * that builds the reversed prefixes, and getScores8/16(phase 1) (bwa-mem2 src/bwamem_pair.cpp).
 * aln[i] receives the complete kswr_t of pairArray[i] (score, te, qe, score2, te2, tb, qb; tb = qb = -1 where ksw_align2
 * runs no second pass).  Records and bases cross PCIe once, no reversed copy is built on the host.  is8: the
 * getScores8 (1) or getScores16 (0) arithmetic.  Chunks go round-robin to the handle's devices. */
void kswv_align2(kswv_t *h, const bsw_seqpair_t *pairArray, const uint8_t *seqBufRef, const uint8_t *seqBufQer,
                 bsw_kswr_t *aln, int32_t numPairs, int is8);

/* kswv_align2 with the target windows taken from the resident reference (SURVEY.md 8(f).2 applied to mate rescue; no
 * reference counterpart).  mem_matesw cuts every window out of the packed reference with bns_fetch_seq only to hand it to
 * ksw_align2 (bwa-mem2 src/bwamem_pair.cpp); here the caller passes, instead of target bases, one coordinate per pair in
 * bwa's doubled coordinate system, exactly as for bsw_getScoresRef: rpos[i] >= 0 -> target[k] = base(rpos[i] + k) (what
 * bns_fetch_seq returns for [rb, re), either strand), rpos[i] < 0 -> target[k] = base(~rpos[i] - k).  pairArray[i].idr is
 * ignored; everything else as kswv_align2.  Records, rpos, the mates and the results are all that crosses PCIe. */
void kswv_align2Ref(kswv_t *h, const bsw_seqpair_t *pairArray, const bsw_ref_t *ref, const int64_t *rpos,
                    const uint8_t *seqBufQer, bsw_kswr_t *aln, int32_t numPairs, int is8);

/* ---- CIGAR generation ------------------------------------------------------------
 * Batched bwa ksw.c ksw_global2 (bwa-mem2 src/ksw.cpp), the banded global alignment with backtrack that mem_reg2aln
 * reaches through bwa_gen_cigar2 once per reported alignment: one job per alignment, scoring from the handle
 * (a / -b / ambiguous, gap open / extend; zdrop and end_bonus do not apply).  Outputs per job: the score, n_cigar, and
 * its ops (len << 4 | op, op 0 = M, 1 = I, 2 = D, as bwa stores them) at cigar[cigar_off .. cigar_off + n_cigar):
 * dense, in job order.  Contract: w >= |tlen - qlen| (bwa_gen_cigar2 guarantees + 3; below it ksw_global2's backtrack
 * leaves its matrix and the result is unspecified there as here).  bwa_gen_cigar2's surroundings (the no-gap shortcut, the band rule, NM / MD) stay in the caller. */
typedef struct bsw_global_job {
    int32_t idr, idq;        /* target = seqBufRef + idr, query = seqBufQer + idq */
    int32_t tlen, qlen, w;   /* lengths and band */
    int32_t score;           /* out */
    int32_t n_cigar;         /* out */
    int32_t cigar_off;       /* out */
} bsw_global_job_t;
/* host buffers; synchronous.  Returns the total number of ops; if they exceed cigarCap returns minus the total (scores
 * and n_cigar are valid, the ops of the jobs that end below cigarCap are written). */
int64_t bsw_global2(bsw_t *h, bsw_global_job_t *jobs, int32_t numJobs, const uint8_t *seqBufRef, const uint8_t *seqBufQer,
                    uint32_t *cigar, int64_t cigarCap);
/* resident buffers; maxBand >= every job's w.  Returns the total number of ops (the host blocks for the totals). */
int64_t bsw_dev_global2(bsw_t *h, bsw_global_job_t *d_jobs, int32_t numJobs, const uint8_t *d_seqBufRef,
                        const uint8_t *d_seqBufQer, uint32_t *d_cigar, int64_t cigarCap, int32_t maxBand, void *stream);

/* pinned host arenas for the caller's SeqPair / sequence buffers (optional:
 * pageable memory works, pinned memory lets the copies run at link speed). */
void *bsw_host_alloc(size_t bytes);
void  bsw_host_free(void *p);

/* ---- counters / introspection ------------------------------------------------ */
typedef struct bsw_stats {
    int64_t calls;            /* boundary calls served                       */
    int64_t pairs;            /* pairs processed                             */
    int64_t pairs_packed;     /* pairs served by the packed s16x2 kernel     */
    int64_t pairs_general;    /* pairs served by the general 32-bit kernel   */
    int64_t pairs_wide;       /* pairs served by the lanes-per-pair 16-bit kernel (large / resident calls: as launched;
                                 small staged calls: the pairs only that kernel admits) */
    int64_t cells;            /* DP cells evaluated, oracle definition (8d); BSW counts them
                                 only while bsw_set_cell_counting(h,1) is on    */
    int64_t kernel_launches;  /* CUDA kernels launched by this library       */
    int64_t h2d_bytes, d2h_bytes;
    double  kernel_ms;        /* device time of the DP kernels (CUDA events), only
                                 when BSW_CUDA_TIMING=1 or bsw_set_timing(h,1)   */
} bsw_stats_t;
void bsw_get_stats(bsw_t *h, bsw_stats_t *out);
void bsw_reset_stats(bsw_t *h);
void bsw_set_timing(bsw_t *h, int on);
/* exact oracle-definition DP-cell counting for BSW (tracks the exact trimmed `beg` per row;
 * ~10% slower, results unchanged).  Off by default; BSW_CUDA_COUNT_CELLS=1 turns it on. */
void bsw_set_cell_counting(bsw_t *h, int on);
void kswv_get_stats(kswv_t *h, bsw_stats_t *out);
void kswv_reset_stats(kswv_t *h);
void kswv_set_timing(kswv_t *h, int on);
/* pair count from which a host-pointer call takes the large-call path (default BSW_CUDA_DIRECT_MIN = 131072) */
void bsw_set_direct_min(bsw_t *h, int pairs);
/* calls / merged batches of up to this many pairs run on the lanes-per-pair kernel (latency), larger ones on the
 * thread-owns-two-pairs kernel (throughput); default BSW_CUDA_WIDE_MAX = 32768.  0: never (pairs only the 16-bit
 * kernel admits still use it). */
void bsw_set_wide_max(bsw_t *h, int pairs);
/* the whole kernel-choice policy of a handle: wide_max as above; calls / batches of up to wide_lat pairs (default
 * BSW_CUDA_WIDE_LAT = 32768) use the kernel's shapes with more lanes per pair (lowest latency); staged batches between
 * wide_lat and wide_max keep the bins of up to wide_split x 16 columns on the thread kernel (default
 * BSW_CUDA_WIDE_SPLIT = 0: none).  Results never depend on it. */
void bsw_set_wide_policy(bsw_t *h, int wide_max, int wide_lat, int wide_split);
void kswv_set_direct_min(kswv_t *h, int pairs);

/* counters of the staged-call engine of one device (process-wide) */
typedef struct bsw_engine_stats {
    int64_t batches;          /* device batches submitted                              */
    int64_t pairs;            /* pairs in them                                         */
    int64_t calls;            /* boundary calls (or slot-sized parts of calls) merged  */
    int64_t kernel_launches;
    int64_t h2d_bytes, d2h_bytes;
    int64_t max_batch_pairs;
    int64_t wide_batches;     /* batches run on the lanes-per-pair kernel (<= BSW_CUDA_WIDE_MAX pairs) */
    /* where the time of staged calls goes, nanoseconds.  Caller threads, summed over calls: */
    int64_t scan_ns;          /*   contract checks + class of every pair                              */
    int64_t reserve_ns;       /*   waiting for room in a staging slot                                 */
    int64_t pack_ns;          /*   compact records + named bases into pinned staging                  */
    int64_t wait_ns;          /*   asleep until the batch is back                                     */
    int64_t unpack_ns;        /*   results into the caller's SeqPair / kswr_t array                   */
    /* engine thread, summed over batches: */
    int64_t packers_ns;       /*   waiting for the last packer of a closed batch                      */
    int64_t submit_ns;        /*   enqueueing copies + kernels                                        */
    int64_t flight_ns;        /*   enqueue done -> completion seen (copies + kernels on the device)   */
} bsw_engine_stats_t;
void bsw_engine_get_stats(int device, bsw_engine_stats_t *out);
void bsw_engine_reset_stats(int device);

/* Measures the sustained issue rate of independent packed-16-bit DPX
 * instructions (VIMNMX3.S16x2 / VIADDMNMX.S16x2) on `device`: the roofline
 * denominator of SURVEY.md section 8d.  Returns instructions per second over
 * all SMs (thread-level: one count per lane per instruction), 0 on failure. */
double bsw_measure_dpx_peak(int device, int which /*0 = vimax3, 1 = viaddmax, 2 = mix*/);

/* number of visible CUDA devices (0 when none: compute calls will exit) */
int bsw_device_count(void);
const char *bsw_version(void);

#ifdef __cplusplus
}
#endif
#endif /* BSW_CUDA_H */
