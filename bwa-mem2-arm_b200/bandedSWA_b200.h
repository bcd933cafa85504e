/*
 * bandedSWA_b200.h -- C++ shim classes with the reference's names and method signatures,
 * forwarding to the C ABI of libbswcuda.so (include/bsw_cuda.h).
 *
 * Drop-in for bwa-mem2 v2.2.1 src/bandedSWA.h (class BandedPairWiseSW, struct SeqPair) and
 * src/kswv.h (class kswv) on the BSW / mate-rescue path only: bwamem.cpp
 * (mem_chain2aln_across_reads_V2) and bwamem_pair.cpp (mem_sam_pe_batch) compile against this
 * header unchanged.  Signatures as quoted in R:docs-archive/INTEGRATION_COMPLETE.md:57-71,95-97 and
 * R:docs-archive/WEEK1_WRAPPER_COMPLETE.md:263-268.
 */
#ifndef BANDED_SWA_B200_H
#define BANDED_SWA_B200_H

#include <stdint.h>
#include "../include/bsw_cuda.h"

/* same layout as the reference's `typedef struct dnaSeqPair {...} SeqPair;` */
typedef bsw_seqpair_t SeqPair;
static_assert(sizeof(SeqPair) == 56, "SeqPair must stay 14 x int32");

#ifndef KSW_XBYTE                 /* normally from bwa's ksw.h */
typedef bsw_kswr_t kswr_t;
#define KSW_XBYTE  BSW_KSW_XBYTE
#define KSW_XSTOP  BSW_KSW_XSTOP
#define KSW_XSUBO  BSW_KSW_XSUBO
#define KSW_XSTART BSW_KSW_XSTART
#endif
static_assert(sizeof(bsw_kswr_t) == 28, "kswr_t must stay 7 x int32");

#define MAX_SEQ_LEN8  BSW_MAX_SEQ_LEN8
#define MAX_SEQ_LEN16 BSW_MAX_SEQ_LEN16

class BandedPairWiseSW {
public:
    BandedPairWiseSW(const int o_del, const int e_del, const int o_ins, const int e_ins, const int zdrop,
                     const int end_bonus, const int8_t *mat_, const int8_t w_match, const int8_t w_mismatch,
                     int numThreads)
        : h_(bsw_create(o_del, e_del, o_ins, e_ins, zdrop, end_bonus, mat_, w_match, w_mismatch, numThreads)) {}
    ~BandedPairWiseSW() { bsw_destroy(h_); }
    BandedPairWiseSW(const BandedPairWiseSW &) = delete;
    BandedPairWiseSW &operator=(const BandedPairWiseSW &) = delete;

    void getScores8(SeqPair *pairArray, uint8_t *seqBufRef, uint8_t *seqBufQer, int32_t numPairs,
                    uint16_t numThreads, int32_t w)
    { bsw_getScores8(h_, pairArray, seqBufRef, seqBufQer, numPairs, numThreads, w); }

    void getScores16(SeqPair *pairArray, uint8_t *seqBufRef, uint8_t *seqBufQer, int32_t numPairs,
                     uint16_t numThreads, int32_t w)
    { bsw_getScores16(h_, pairArray, seqBufRef, seqBufQer, numPairs, numThreads, w); }

    void scalarBandedSWAWrapper(SeqPair *seqPairArray, uint8_t *seqBufRef, uint8_t *seqBufQer, int numPairs,
                                int nthreads, int32_t w)
    { bsw_scalarBandedSWAWrapper(h_, seqPairArray, seqBufRef, seqBufQer, numPairs, nthreads, w); }

    /* device-resident pairs: the whole caller protocol of the extension step (left attempts with band retry, right h0
     * chained from the left score by regid, right attempts) in one submission; see bsw_dev_extendChained */
    void extendChainedResident(SeqPair *d_left, const uint8_t *d_leftRef, const uint8_t *d_leftQer, int32_t numLeft,
                               SeqPair *d_right, const uint8_t *d_rightRef, const uint8_t *d_rightQer, int32_t numRight,
                               int32_t w, int32_t maxBandTry, void *stream)
    { bsw_dev_extendChained(h_, d_left, d_leftRef, d_leftQer, numLeft, d_right, d_rightRef, d_rightQer, numRight, w, maxBandTry, stream); }

    /* ---- optional additions next to the unchanged signatures (include/bsw_cuda.h) ---- */
    /* the sequence buffers both tier calls of a batch share go up once: declare them around the calls */
    void arenaBegin(const uint8_t *seqBufRef, size_t refBytes, const uint8_t *seqBufQer, size_t qerBytes)
    { bsw_arena_begin(h_, seqBufRef, refBytes, seqBufQer, qerBytes); }
    void arenaEnd() { bsw_arena_end(h_); }
    void arenaDefer(bool on) { bsw_arena_defer(h_, on ? 1 : 0); }
    /* targets from the reference resident on the GPU (bsw_ref_upload of the .pac): rpos[i] as bns_fetch_seq counts
     * (doubled-strand coordinate; ~x = read downwards from x, the reversed slices of left extensions) */
    void getScoresRef(SeqPair *pairArray, const bsw_ref_t *ref, const int64_t *rpos, uint8_t *seqBufQer, int32_t numPairs, int32_t w)
    { bsw_getScoresRef(h_, pairArray, ref, rpos, seqBufQer, numPairs, w); }
    /* CIGAR generation after the extension: a batch of bwa_gen_cigar2's ksw_global2 calls */
    int64_t global2(bsw_global_job_t *jobs, int32_t numJobs, const uint8_t *seqBufRef, const uint8_t *seqBufQer,
                    uint32_t *cigar, int64_t cigarCap)
    { return bsw_global2(h_, jobs, numJobs, seqBufRef, seqBufQer, cigar, cigarCap); }

    int64_t getTicks() { return 0; }          /* the reference's rdtsc counter; device time is in bsw_get_stats */
    bsw_t *handle() { return h_; }

private:
    bsw_t *h_;
};

class kswv {
public:
    kswv(const int o_del, const int e_del, const int o_ins, const int e_ins, const int8_t w_match,
         const int8_t w_mismatch, int numThreads, int32_t maxRefLen, int32_t maxQerLen)
        : h_(kswv_create(o_del, e_del, o_ins, e_ins, w_match, w_mismatch, numThreads, maxRefLen, maxQerLen)) {}
    ~kswv() { kswv_destroy(h_); }
    kswv(const kswv &) = delete;
    kswv &operator=(const kswv &) = delete;

    void getScores8(SeqPair *pairArray, uint8_t *seqBufRef, uint8_t *seqBufQer, kswr_t *aln, int32_t numPairs,
                    uint16_t numThreads, int phase)
    { kswv_getScores8(h_, pairArray, seqBufRef, seqBufQer, reinterpret_cast<bsw_kswr_t *>(aln), numPairs, numThreads, phase); }

    void getScores16(SeqPair *pairArray, uint8_t *seqBufRef, uint8_t *seqBufQer, kswr_t *aln, int32_t numPairs,
                     uint16_t numThreads, int phase)
    { kswv_getScores16(h_, pairArray, seqBufRef, seqBufQer, reinterpret_cast<bsw_kswr_t *>(aln), numPairs, numThreads, phase); }

    /* host buffers: ksw_align2 in ONE call -- replaces getScores(phase 0) + the host hand-off loop + getScores(phase 1) */
    void align2(const SeqPair *pairArray, const uint8_t *seqBufRef, const uint8_t *seqBufQer, kswr_t *aln, int32_t numPairs, int is8)
    { kswv_align2(h_, pairArray, seqBufRef, seqBufQer, reinterpret_cast<bsw_kswr_t *>(aln), numPairs, is8); }

    /* the same with the windows cut from the resident reference (bsw_ref_upload) at rpos[i] = mem_matesw's rb: no target
     * bases cross PCIe; see kswv_align2Ref */
    void align2Ref(const SeqPair *pairArray, const bsw_ref_t *ref, const int64_t *rpos, const uint8_t *seqBufQer, kswr_t *aln,
                   int32_t numPairs, int is8)
    { kswv_align2Ref(h_, pairArray, ref, rpos, seqBufQer, reinterpret_cast<bsw_kswr_t *>(aln), numPairs, is8); }

    /* device-resident pairs: ksw_align2 (forward pass, hand-off, reverse pass) in one submission; see kswv_dev_align2 */
    void align2Resident(const SeqPair *d_pairArray, const uint8_t *d_seqBufRef, const uint8_t *d_seqBufQer, kswr_t *d_aln,
                        int32_t numPairs, int is8, void *stream)
    { kswv_dev_align2(h_, d_pairArray, d_seqBufRef, d_seqBufQer, reinterpret_cast<bsw_kswr_t *>(d_aln), numPairs, is8, stream); }

    kswv_t *handle() { return h_; }

private:
    kswv_t *h_;
};

#endif /* BANDED_SWA_B200_H */
