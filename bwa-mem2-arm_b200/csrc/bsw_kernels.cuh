/*
 * bsw_kernels.cuh -- __global__ kernels of the BSW / kswv path (sm_100a).
 *
 * Launch sequence of one boundary call (all on one stream, DESIGN.md "Pipeline"):
 *   1. prep_kernel      one thread per pair: admissibility for the packed kernel,
 *                       radix key (query length major, target length minor),
 *                       per-bin histogram (bin = 16 query columns).
 *   2. cub radix sort   descending: longest queries first (heavy tiles first),
 *                       pairs the packed kernel does not admit last (key 0).
 *   3. histogram read-back (one 200-byte D2H): the host derives each bin's sorted range.
 *   4. *_packed_kernel  one launch per non-empty bin, longest queries first.  A block is ONE
 *                       warp = 64 pairs (2 per thread); its dynamic shared memory holds exactly
 *                       that bin's DP rows (NE x 192 B), so the number of resident warps per SM
 *                       is set by the query length, with no block-level barrier anywhere.
 *   5. *_general_kernel 32-bit path for the (rare) pairs step 1 did not admit.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/bsw_cuda.h"
#include "bsw_core.cuh"
#include "bsw_core2.cuh"
#include "kswv_core.cuh"

namespace bswk {

#define MISW_OF(c) (((uint32_t)(-(c).b) & 0xffu) * 0x01010101u)
#define XW_OF(c) (((uint32_t)(c).a ^ (uint32_t)(-(c).b)) & 0xffu)

#define BSW_NBINS 16            /* bins of 16 elements: up to 256 elements per thread */
#define BSW_WARP_ELEM_BYTES 192 /* per element per warp: 32 x (state word + uint16 selector) */

struct Plan {
    int hist[BSW_NBINS + 2];       /* [qb] = packed-admissible pairs whose rows need qb x 16 columns */
    int hist_n[BSW_NBINS + 2];     /* BSW: the subset of them whose query holds an N / invalid base (sorted first; counted here, not in hist) */
    int n_general;                 /* pairs routed to the general kernel */
    int max_len1;                  /* over all pairs                      */
    int max_len2_general;
    int bad_pairs;                 /* contract violations seen by prep    */
    int oob_pairs;                 /* pairs whose bases fall outside the uploaded byte ranges */
};

enum { MODE_BSW = 0, MODE_KSWV = 1 };

/* ------------------------------------------------------------------ 1. prep */
/* does the query hold a base > 3?  Aligned word reads; stays inside the words that hold the query's bytes */
__device__ __forceinline__ bool query_has_n(const uint8_t *__restrict__ qer, long long idq, int len2)
{
    const uintptr_t a = reinterpret_cast<uintptr_t>(qer + idq);
    const uint32_t *wp = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
    const int off = (int)(a & 3), nbytes = off + len2, nw = (nbytes + 3) >> 2;
    uint32_t acc = 0;
    for (int k = 0; k < nw; ++k) {
        uint32_t w = __ldg(wp + k);
        if (k == 0) w &= 0xffffffffu << (8 * off);
        if (k == nw - 1 && (nbytes & 3)) w &= 0xffffffffu >> (8 * (4 - (nbytes & 3)));
        acc |= w;
    }
    return (acc & 0xfcfcfcfcu) != 0;
}

#define BSW_KEY_BITS 26         /* (has-N flag) << 25 | (len2 + 1) << 16 | (len1 / 2) << 8 | h0 */

__global__ void prep_kernel(const bsw_seqpair_t *__restrict__ pairs, const uint8_t *__restrict__ qer, int n, Const c, int mode,
                            int max_elems, int tier8_check, long long rlo, long long rhi, long long qlo, long long qhi,
                            uint32_t *__restrict__ keys, uint32_t *__restrict__ idx, Plan *plan)
{
    __shared__ int s_hist[2 * (BSW_NBINS + 2)];
    __shared__ int s_misc[4];
    if (threadIdx.x < 2 * (BSW_NBINS + 2)) s_hist[threadIdx.x] = 0;
    if (threadIdx.x < 4) s_misc[threadIdx.x] = 0;
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const int len1 = pairs[i].len1, len2 = pairs[i].len2, h0 = pairs[i].h0;
        bool ok = (mode == MODE_BSW) ? packed_eligible(c, len1, len2, h0) : kswv_packed_eligible(c, len1, len2);
        /* columns a thread needs: len2 + 1 (kswv), len2 + 4 (BSW: the sweep's slack, bsw_core2.cuh) */
        const int qb = (len2 + (mode == MODE_BSW ? BSW_SWEEP_SLACK : 1) + 15) >> 4;
        if (ok && qb * 16 > max_elems) ok = false;
        uint32_t key = 0;
        if (ok) {
            /* query length major (sets the bin), then target length and h0, so that the two pairs a
             * thread carries terminate, trim and widen alike */
            const uint32_t l1 = (uint32_t)(len1 >> 1) > 255u ? 255u : (uint32_t)(len1 >> 1);
            const uint32_t hh = mode == MODE_BSW ? ((uint32_t)h0 > 255u ? 255u : (uint32_t)h0) : 0u;
            key = ((uint32_t)(len2 + 1) << 16) | (l1 << 8) | hh;
            /* BSW: queries with an N go to the kernel variant with the general score table (bsw_core2.cuh, QN).  Only
             * pairs inside the uploaded byte ranges are read here (the others make the host redo the chunk). */
            const long long iq0 = pairs[i].idq;
            const bool qn = mode == MODE_BSW && len2 > 0 && iq0 >= qlo && iq0 + len2 <= qhi && query_has_n(qer, iq0, len2);
            if (qn) key |= 1u << 25;
            atomicAdd(&s_hist[qb + (qn ? BSW_NBINS + 2 : 0)], 1);
        } else {
            atomicAdd(&s_misc[0], 1);
            atomicMax(&s_misc[2], len2);
        }
        atomicMax(&s_misc[1], len1);
        bool bad = len1 < 0 || len2 < 0 || (mode == MODE_BSW && h0 <= 0);
        if (tier8_check && (len1 >= BSW_MAX_SEQ_LEN8 || len2 >= BSW_MAX_SEQ_LEN8)) bad = true;
        if (bad) atomicAdd(&s_misc[3], 1);
        /* the host estimates the byte ranges to upload from a sample of the pairs; verify them all here */
        const long long ir = pairs[i].idr, iq = pairs[i].idq;
        if (!bad && ((len1 > 0 && (ir < rlo || ir + len1 > rhi)) || (len2 > 0 && (iq < qlo || iq + len2 > qhi))))
            atomicAdd(&plan->oob_pairs, 1);
        keys[i] = key; idx[i] = (uint32_t)i;
    }
    __syncthreads();
    if (threadIdx.x < 2 * (BSW_NBINS + 2) && s_hist[threadIdx.x])      /* hist and hist_n are adjacent in Plan */
        atomicAdd(&plan->hist[threadIdx.x], s_hist[threadIdx.x]);
    if (threadIdx.x == 0) {
        if (s_misc[0]) atomicAdd(&plan->n_general, s_misc[0]);
        atomicMax(&plan->max_len1, s_misc[1]);
        atomicMax(&plan->max_len2_general, s_misc[2]);
        if (s_misc[3]) atomicAdd(&plan->bad_pairs, s_misc[3]);
    }
}

__device__ __forceinline__ PairIn load_pair(const bsw_seqpair_t *__restrict__ pairs, int pi)
{
    PairIn p;
    if (pi < 0) { p.idr = p.idq = p.len1 = p.len2 = p.h0 = 0; p.valid = 0; return p; }
    p.idr = pairs[pi].idr; p.idq = pairs[pi].idq; p.len1 = pairs[pi].len1; p.len2 = pairs[pi].len2;
    p.h0 = pairs[pi].h0; p.valid = 1;
    return p;
}

/* ------------------------------------------------------------------ 4a. BSW packed */
#ifndef BSW_MINBLOCKS
#define BSW_MINBLOCKS 1
#endif
template <bool COUNT, bool SYM, bool QN>
__global__ void __launch_bounds__(32, BSW_MINBLOCKS)
bsw_packed_kernel(bsw_seqpair_t *__restrict__ pairs, const uint8_t *__restrict__ ref,
                  const uint8_t *__restrict__ qer, const uint32_t *__restrict__ sorted_idx,
                  int first, int bin_end, int NE, Const c, int w, unsigned long long *cells)
{
    extern __shared__ __align__(16) uint8_t smem[];
    const int lane = threadIdx.x;
    uint32_t *he = reinterpret_cast<uint32_t *>(smem) + lane;
    uint16_t *qs = reinterpret_cast<uint16_t *>(smem + (size_t)NE * 128) + lane;
    const int slot = first + blockIdx.x * 64 + lane * 2;
    const int pa = slot < bin_end ? (int)sorted_idx[slot] : -1;
    const int pb = slot + 1 < bin_end ? (int)sorted_idx[slot + 1] : -1;
    /* per-row score tables by target base: byte q of lut_tbl[t] = score(t, q) */
    __shared__ uint32_t lut_tbl[8];
    if (lane < 5) lut_tbl[lane] = bsw_lutrow(MISW_OF(c), XW_OF(c), lut_hi(c), lane * 8u);
    __syncwarp();
    unsigned long long my_cells = 0;
    if (pa >= 0) {
        PairIn A = load_pair(pairs, pa), B = load_pair(pairs, pb);
        PairOut oa, ob;
        if (COUNT) bsw_packed_thread<32, true, SYM>(c, w, A, B, ref, qer, he, qs, oa, ob);
        else bsw_packed_thread2<32, SYM, QN>(c, w, A, B, ref, qer, he, qs, lut_tbl, oa, ob);
        int2 *o = reinterpret_cast<int2 *>(&pairs[pa].score);
        o[0] = make_int2(oa.score, oa.tle); o[1] = make_int2(oa.gtle, oa.qle); o[2] = make_int2(oa.gscore, oa.max_off);
        my_cells += (unsigned long long)oa.cells;
        if (pb >= 0) {
            o = reinterpret_cast<int2 *>(&pairs[pb].score);
            o[0] = make_int2(ob.score, ob.tle); o[1] = make_int2(ob.gtle, ob.qle); o[2] = make_int2(ob.gscore, ob.max_off);
            my_cells += (unsigned long long)ob.cells;
        }
    }
    if (COUNT) {
        for (int o = 16; o; o >>= 1) my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o);
        if (lane == 0 && my_cells) atomicAdd(cells, my_cells);
    }
}

/* ------------------------------------------------------------------ 5a. BSW general */
__global__ void bsw_general_kernel(bsw_seqpair_t *__restrict__ pairs, const uint8_t *__restrict__ ref,
                                   const uint8_t *__restrict__ qer, const uint32_t *__restrict__ sorted_idx,
                                   int first, int count, Const c, int w, int2 *eh, unsigned long long *cells)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long my_cells = 0;
    if (g < count) {
        const int pi = (int)sorted_idx[first + g];
        PairIn P = load_pair(pairs, pi);
        PairOut o;
        bsw_general_thread<long long>(c, w, P, ref, qer, eh + g, (long long)count, o);
        pairs[pi].score = o.score; pairs[pi].tle = o.tle; pairs[pi].gtle = o.gtle; pairs[pi].qle = o.qle;
        pairs[pi].gscore = o.gscore; pairs[pi].max_off = o.max_off;
        my_cells = (unsigned long long)o.cells;
    }
    for (int o = 16; o; o >>= 1) my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o);
    if ((threadIdx.x & 31) == 0 && my_cells) atomicAdd(cells, my_cells);
}

/* ------------------------------------------------------------------ kswv result write-back */
__device__ __forceinline__ void kswv_store(bsw_kswr_t *__restrict__ aln, const bsw_seqpair_t *__restrict__ pairs,
                                           int pi, const KswOut &o, int phase)
{
    if (phase == 0) {
        bsw_kswr_t r; r.score = o.score; r.te = o.te; r.qe = o.qe; r.score2 = o.score2; r.te2 = o.te2; r.tb = -1; r.qb = -1;
        aln[pi] = r;
    } else {
        bsw_kswr_t *a = &aln[pairs[pi].regid];
        if (a->score == o.score) { a->tb = a->te - o.te; a->qb = a->qe - o.qe; }
    }
}

/* ------------------------------------------------------------------ 4b. kswv packed */
template <bool SYM>
__global__ void __launch_bounds__(32)
kswv_packed_kernel(const bsw_seqpair_t *__restrict__ pairs, const uint8_t *__restrict__ ref,
                   const uint8_t *__restrict__ qer, bsw_kswr_t *__restrict__ aln,
                   const uint32_t *__restrict__ sorted_idx, int first, int bin_end, int NE, Const c, int is8, int phase,
                   uint16_t *rowmax, long long rm_stride, unsigned long long *cells)
{
    extern __shared__ __align__(16) uint8_t smem[];
    const int lane = threadIdx.x;
    uint32_t *he = reinterpret_cast<uint32_t *>(smem) + lane;
    uint16_t *qs = reinterpret_cast<uint16_t *>(smem + (size_t)NE * 128) + lane;
    const int slot = first + blockIdx.x * 64 + lane * 2;
    const int pa = slot < bin_end ? (int)sorted_idx[slot] : -1;
    const int pb = slot + 1 < bin_end ? (int)sorted_idx[slot + 1] : -1;
    unsigned long long my_cells = 0;
    if (pa >= 0) {
        PairIn A = load_pair(pairs, pa), B = load_pair(pairs, pb);
        KswOut oa, ob;
        kswv_packed_thread<32, SYM>(c, is8, A, B, ref, qer, he, qs, rowmax + slot, rm_stride, oa, ob);
        kswv_store(aln, pairs, pa, oa, phase);
        my_cells += (unsigned long long)oa.cells;
        if (pb >= 0) { kswv_store(aln, pairs, pb, ob, phase); my_cells += (unsigned long long)ob.cells; }
    }
    for (int o = 16; o; o >>= 1) my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o);
    if (lane == 0 && my_cells) atomicAdd(cells, my_cells);
}

/* ------------------------------------------------------------------ 5b. kswv general */
__global__ void kswv_general_kernel(const bsw_seqpair_t *__restrict__ pairs, const uint8_t *__restrict__ ref,
                                    const uint8_t *__restrict__ qer, bsw_kswr_t *__restrict__ aln,
                                    const uint32_t *__restrict__ sorted_idx, int first, int count,
                                    Const c, int is8, int phase, int2 *he, int *rm, unsigned long long *cells)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long my_cells = 0;
    if (g < count) {
        const int pi = (int)sorted_idx[first + g];
        PairIn P = load_pair(pairs, pi);
        KswOut o;
        kswv_general_thread<long long>(c, is8, P, ref, qer, he + g, (long long)count, rm + g, (long long)count, o);
        kswv_store(aln, pairs, pi, o, phase);
        my_cells = (unsigned long long)o.cells;
    }
    for (int o = 16; o; o >>= 1) my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o);
    if ((threadIdx.x & 31) == 0 && my_cells) atomicAdd(cells, my_cells);
}

/* ------------------------------------------------------------------ DPX issue-rate microbenchmark */
template <int WHICH>
__global__ void dpx_peak_kernel(uint32_t *out, int iters, uint32_t seed)
{
    uint32_t a0 = seed + threadIdx.x, a1 = a0 * 3u, a2 = a0 * 5u, a3 = a0 * 7u,
             a4 = a0 * 11u, a5 = a0 * 13u, a6 = a0 * 17u, a7 = a0 * 19u;
    const uint32_t k1 = seed ^ 0x00030005u, k2 = seed ^ 0x00010002u;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (WHICH == 0) {
                a0 = __vimax3_s16x2(a0, k1, k2); a1 = __vimax3_s16x2(a1, k1, k2); a2 = __vimax3_s16x2(a2, k1, k2); a3 = __vimax3_s16x2(a3, k1, k2);
                a4 = __vimax3_s16x2(a4, k1, k2); a5 = __vimax3_s16x2(a5, k1, k2); a6 = __vimax3_s16x2(a6, k1, k2); a7 = __vimax3_s16x2(a7, k1, k2);
            } else if (WHICH == 1) {
                a0 = __viaddmax_s16x2(a0, k1, k2); a1 = __viaddmax_s16x2(a1, k1, k2); a2 = __viaddmax_s16x2(a2, k1, k2); a3 = __viaddmax_s16x2(a3, k1, k2);
                a4 = __viaddmax_s16x2(a4, k1, k2); a5 = __viaddmax_s16x2(a5, k1, k2); a6 = __viaddmax_s16x2(a6, k1, k2); a7 = __viaddmax_s16x2(a7, k1, k2);
            } else if (WHICH == 2) {
                a0 = __vimax3_s16x2(a0, k1, k2); a1 = __viaddmax_s16x2_relu(a1, k1, k2); a2 = __viaddmin_s16x2(a2, k1, k2); a3 = __vimax3_s16x2(a3, k1, k2);
                a4 = __viaddmax_s16x2_relu(a4, k1, k2); a5 = __viaddmin_s16x2(a5, k1, k2); a6 = __vimax3_s16x2(a6, k1, k2); a7 = __viaddmax_s16x2_relu(a7, k1, k2);
            } else if (WHICH == 3) {   /* plain 32-bit integer add/logic on the same pipe, for comparison */
                a0 = (a0 ^ k1) + k2; a1 = (a1 ^ k1) + k2; a2 = (a2 ^ k1) + k2; a3 = (a3 ^ k1) + k2;
                a4 = (a4 ^ k1) + k2; a5 = (a5 ^ k1) + k2; a6 = (a6 ^ k1) + k2; a7 = (a7 ^ k1) + k2;
            } else {                   /* VIADD.16x2 */
                a0 = __vadd2(a0, k1); a1 = __vadd2(a1, k1); a2 = __vadd2(a2, k1); a3 = __vadd2(a3, k1);
                a4 = __vadd2(a4, k1); a5 = __vadd2(a5, k1); a6 = __vadd2(a6, k1); a7 = __vadd2(a7, k1);
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
}

/* ------------------------------------------------------------------ caller protocol on the device
 * The bookkeeping mem_chain2aln_across_reads_V2 does on the host between its BSW calls (row a9), as kernels over the
 * resident SeqPair array: remember the scores before an attempt, pick the pairs that need the wider band, move them to
 * / from a dense sub-array, and hand every right extension the score of its read's left extension as h0. */
__global__ void proto_save_score_kernel(const bsw_seqpair_t *__restrict__ pairs, int n, int *__restrict__ prev)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) prev[i] = pairs[i].score;
}

/* retry rule: score != score before the attempt && max_off >= (w >> 1) + (w >> 2), w = band of the attempt just done */
__global__ void proto_select_retry_kernel(const bsw_seqpair_t *__restrict__ pairs, const int *__restrict__ prev, int n,
                                          int thr, int *__restrict__ list, int *__restrict__ count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool pick = i < n && pairs[i].score != prev[i] && pairs[i].max_off >= thr;
    const unsigned m = __ballot_sync(0xffffffffu, pick);
    if (!m) return;
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == __ffs(m) - 1) base = atomicAdd(count, __popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    if (pick) list[base + __popc(m & ((1u << lane) - 1u))] = i;
}

__global__ void proto_gather_kernel(const bsw_seqpair_t *__restrict__ pairs, const int *__restrict__ list, int m,
                                    bsw_seqpair_t *__restrict__ sub)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < m) sub[j] = pairs[list[j]];
}

__global__ void proto_scatter_kernel(const bsw_seqpair_t *__restrict__ sub, const int *__restrict__ list, int m,
                                     bsw_seqpair_t *__restrict__ pairs)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < m) pairs[list[j]] = sub[j];
}

/* right h0 = score of the left extension the pair's regid names (regid < 0: the pair keeps its h0) */
__global__ void proto_chain_h0_kernel(const bsw_seqpair_t *__restrict__ left, int n_left, bsw_seqpair_t *__restrict__ right,
                                      int n_right)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_right) return;
    const int r = right[i].regid;
    if (r >= 0 && r < n_left) right[i].h0 = left[r].score;
}


/* host-pointer path with gathered sequences: exchanges (idr, idq) of the device copy of every pair with the offsets
 * into the gathered buffers; a second run puts the caller's offsets back before the results are downloaded */
__global__ void swap_offsets_kernel(bsw_seqpair_t *__restrict__ pairs, int2 *__restrict__ off, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int2 o = off[i];
    off[i] = make_int2(pairs[i].idr, pairs[i].idq);
    pairs[i].idr = o.x; pairs[i].idq = o.y;
}

} /* namespace bswk */
