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
 *   4. bsw_packed_kernel one launch per non-empty bin, longest queries first.  A block is ONE
 *                       warp = 64 pairs (2 per thread); its dynamic shared memory holds exactly
 *                       that bin's DP rows (NE x 192 B), so the number of resident warps per SM
 *                       is set by the query length, with no block-level barrier anywhere.
 *      kswv_sys_kernel  one launch per non-empty bin: a warp = 4 groups of 8 lanes, two pairs per group,
 *                       DP state in registers (kswv_sys.cuh).
 *   5. *_general_kernel 32-bit path for the (rare) pairs step 1 did not admit.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <stdlib.h>
#include "../../include/bsw_cuda.h"
#include "bsw_core.cuh"
#include "bsw_core2.cuh"
#include "kswv_core.cuh"
#include "kswv_sys.cuh"
#include "bsw_wide.cuh"
#include "global_core.cuh"

namespace bswk {

#define MISW_OF(c) (((uint32_t)(-(c).b) & 0xffu) * 0x01010101u)
#define XW_OF(c) (((uint32_t)(c).a ^ (uint32_t)(-(c).b)) & 0xffu)

#define BSW_NBINS 16            /* bins of 16 elements: up to 256 elements per thread */
#define BSW_WARP_ELEM_BYTES 192 /* per element per warp: 32 x (state word + uint16 selector) */

struct Plan {
    int hist[BSW_NBINS + 2];       /* [qb] = packed-admissible pairs whose rows need qb x 16 columns */
    int hist_n[BSW_NBINS + 2];     /* BSW: the subset of them whose query holds an N / invalid base (sorted first; counted here, not in hist) */
    int hist_w[BSW_WIDE_NCFG + 1]; /* BSW: pairs only the lanes-per-pair kernel admits (16-bit state), by configuration */
    int n_general;                 /* pairs routed to the general kernel */
    int max_len1;                  /* over all pairs                      */
    int max_len1_packed;           /* over the pairs the packed kernels take (sizes kswv's `b` lists in shared memory) */
    int max_len2_general;
    int bad_pairs;                 /* contract violations seen by prep    */
    int oob_pairs;                 /* pairs whose bases fall outside the uploaded byte ranges */
    int max_len2_bin[BSW_NBINS + 2]; /* kswv: longest query of each bin (N and plain classes together): columns per lane of its launch */
};

enum { MODE_BSW = 0, MODE_KSWV = 1 };
static_assert(KSWV_SYSW_NCFG <= BSW_WIDE_NCFG, "the 16-bit kswv shapes share Plan::hist_w and the CLS_WIDE0 classes with the BSW lanes-per-pair configurations");

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

#define BSW_KEY_BITS 26         /* packed pairs: (has-N flag) << 25 | (len2 + 1) << 16 | (len1 / 2) << 8 | h0  (>= 0x10000);
                                 * wide-only pairs: (configuration + 1) << 12 | len1 / 2  (< 0x10000); general pairs: 0 */

__global__ void prep_kernel(const bsw_seqpair_t *__restrict__ pairs, const uint8_t *__restrict__ qer, int n, Const c, int mode,
                            int max_elems, int tier8_check, int allow_wide, long long rlo, long long rhi, long long qlo, long long qhi,
                            uint32_t *__restrict__ keys, uint32_t *__restrict__ idx, Plan *plan)
{
    __shared__ int s_hist[2 * (BSW_NBINS + 2)];
    __shared__ int s_misc[5];
    __shared__ int s_wide[BSW_WIDE_NCFG];
    __shared__ int s_ml2[BSW_NBINS + 2];
    if (threadIdx.x < 2 * (BSW_NBINS + 2)) s_hist[threadIdx.x] = 0;
    if (threadIdx.x < 5) s_misc[threadIdx.x] = 0;
    if (threadIdx.x < BSW_WIDE_NCFG) s_wide[threadIdx.x] = 0;
    if (threadIdx.x < BSW_NBINS + 2) s_ml2[threadIdx.x] = 0;
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const int len1 = pairs[i].len1, len2 = pairs[i].len2, h0 = pairs[i].h0;
        bool ok = (mode == MODE_BSW) ? packed_eligible(c, len1, len2, h0) : kswv_sys_eligible(c, len1, len2);
        /* BSW: columns a thread needs = len2 + 4 (the sweep's slack, bsw_core2.cuh); kswv: the query's columns (kswv_sys.cuh) */
        int qb = (len2 + (mode == MODE_BSW ? BSW_SWEEP_SLACK : 0) + 15) >> 4;
        qb = qb < 1 ? 1 : qb;
        if (ok && qb * 16 > max_elems) ok = false;
        uint32_t key = 0;
        if (ok) {
            /* query length major (sets the bin), then target length and h0, so that the two pairs a
             * thread carries terminate, trim and widen alike */
            const uint32_t l1 = (uint32_t)(len1 >> 1) > 255u ? 255u : (uint32_t)(len1 >> 1);
            const uint32_t hh = mode == MODE_BSW ? ((uint32_t)h0 > 255u ? 255u : (uint32_t)h0) : 0u;
            key = mode == MODE_BSW ? ((uint32_t)(len2 + 1) << 16) | (l1 << 8) | hh
                                   : ((uint32_t)(len2 + 1) << 16) | ((uint32_t)len1 << 5);     /* len1 <= KSWV_SYS_MAX_ROWS < 2048 */
            /* queries with an N go to the kernel variant with the general score table (bsw_core2.cuh / kswv_sys.cuh, QN).
             * Only pairs inside the uploaded byte ranges are read here (the others make the host redo the chunk). */
            const long long iq0 = pairs[i].idq;
            const bool qn = len2 > 0 && iq0 >= qlo && iq0 + len2 <= qhi && query_has_n(qer, iq0, len2);
            if (qn) key |= 1u << 25;
            atomicAdd(&s_hist[qb + (qn ? BSW_NBINS + 2 : 0)], 1);
            atomicMax(&s_misc[4], len1);
            if (mode == MODE_KSWV) atomicMax(&s_ml2[qb], len2);
        } else if (mode == MODE_KSWV && kswv_sysw_eligible(c, tier8_check, len1, len2)) {
            /* scores or columns beyond the byte key: the 16-bit shapes of the registers kernel (kswv_sys.cuh, W) */
            const int cfg = kswv_sysw_cfg_for(len2);
            key = ((uint32_t)(cfg + 1) << 12) | ((uint32_t)len1 < 4095u ? (uint32_t)len1 : 4095u);
            atomicAdd(&s_wide[cfg], 1);
            atomicMax(&s_misc[4], len1);
        } else if (mode == MODE_BSW && allow_wide && wide_eligible(c, len1, len2, h0)) {
            /* too long or too high-scoring for the byte-packed rows: the 16-bit lanes-per-pair kernel takes it */
            const int cfg = wide_cfg_for(len2);
            const uint32_t l1 = (uint32_t)(len1 >> 1) > 4095u ? 4095u : (uint32_t)(len1 >> 1);
            key = ((uint32_t)(cfg + 1) << 12) | l1;
            atomicAdd(&s_wide[cfg], 1);
        } else {
            atomicAdd(&s_misc[0], 1);
            atomicMax(&s_misc[2], len2);
        }
        atomicMax(&s_misc[1], len1);
        bool bad = len1 < 0 || len2 < 0 || (mode == MODE_BSW && h0 <= 0);
        if (mode == MODE_BSW && tier8_check && (len1 >= BSW_MAX_SEQ_LEN8 || len2 >= BSW_MAX_SEQ_LEN8)) bad = true;   /* kswv: tier8_check = is8 */
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
    if (threadIdx.x < BSW_WIDE_NCFG && s_wide[threadIdx.x]) atomicAdd(&plan->hist_w[threadIdx.x], s_wide[threadIdx.x]);
    if (threadIdx.x < BSW_NBINS + 2 && s_ml2[threadIdx.x]) atomicMax(&plan->max_len2_bin[threadIdx.x], s_ml2[threadIdx.x]);
    if (threadIdx.x == 0) {
        if (s_misc[0]) atomicAdd(&plan->n_general, s_misc[0]);
        atomicMax(&plan->max_len1, s_misc[1]);
        atomicMax(&plan->max_len1_packed, s_misc[4]);
        atomicMax(&plan->max_len2_general, s_misc[2]);
        if (s_misc[3]) atomicAdd(&plan->bad_pairs, s_misc[3]);
    }
}

/* Where a kernel finds pair i's inputs and puts its results.  Inputs are the first 24 bytes of a SeqPair
 * ({idr, idq, id, len1, len2, h0}) at in + i * in_stride; the six BSW result words go to out + i * out_stride.
 *   resident / direct calls: the caller's 56-byte records, in place (in = pairs, out = &pairs[0].score, strides 56);
 *   staged calls (engine.cuh): compact 24-byte input records and a separate dense 24-byte result array. */
struct PairIO {
    const char *in; char *out;
    int in_stride, out_stride;
};
static inline PairIO pairio_records(bsw_seqpair_t *pairs)
{
    PairIO io; io.in = (const char *)pairs; io.out = (char *)pairs + offsetof(bsw_seqpair_t, score);
    io.in_stride = io.out_stride = (int)sizeof(bsw_seqpair_t);
    return io;
}
static inline PairIO pairio_compact(const void *in24, void *out24)
{
    PairIO io; io.in = (const char *)in24; io.out = (char *)out24; io.in_stride = io.out_stride = 24;
    return io;
}

__device__ __forceinline__ PairIn load_pair(const PairIO &io, int pi)
{
    PairIn p;
    if (pi < 0) { p.idr = p.idq = p.len1 = p.len2 = p.h0 = 0; p.valid = 0; return p; }
    const int2 *r = reinterpret_cast<const int2 *>(io.in + (size_t)pi * io.in_stride);
    const int2 a = r[0], b = r[1], c = r[2];        /* {idr, idq} {id, len1} {len2, h0} */
    p.idr = a.x; p.idq = a.y; p.len1 = b.y; p.len2 = c.x; p.h0 = c.y; p.valid = 1;
    return p;
}

__device__ __forceinline__ void store_bsw(const PairIO &io, int pi, const PairOut &o)
{
    int2 *w = reinterpret_cast<int2 *>(io.out + (size_t)pi * io.out_stride);
    w[0] = make_int2(o.score, o.tle); w[1] = make_int2(o.gtle, o.qle); w[2] = make_int2(o.gscore, o.max_off);
}

/* Which pairs a block of a packed kernel serves.  One launch covers one or more query-length bins: bin k owns
 * blocks [first_block[k], first_block[k+1]) and the sorted pair slots [first_pair[k], first_pair[k+1]), 64 per
 * block, and lays its rows out for ne[k] columns.  The per-bin launches of the large-call paths pass one bin; the
 * staged small-call path (engine.cuh) passes all bins of a batch in ONE launch sized for the longest bin. */
#define BSW_MAX_TBL 17
struct BinTable {
    int nbins;
    int first_block[BSW_MAX_TBL + 1];
    int first_pair[BSW_MAX_TBL + 1];
    int ne[BSW_MAX_TBL];
};
static inline BinTable bintable_single(int first, int count, int NE)
{
    BinTable t; memset(&t, 0, sizeof t);
    t.nbins = 1; t.first_block[0] = 0; t.first_block[1] = (count + 63) / 64;
    t.first_pair[0] = first; t.first_pair[1] = first + count; t.ne[0] = NE;
    return t;
}
/* block -> (first sorted slot of the block, end of its bin, columns) */
__device__ __forceinline__ void bintable_lookup(const BinTable &t, int block, int &slot0, int &bin_end, int &NE)
{
    int k = 0;
    while (k + 1 < t.nbins && block >= t.first_block[k + 1]) ++k;
    slot0 = t.first_pair[k] + (block - t.first_block[k]) * 64;
    bin_end = t.first_pair[k + 1];
    NE = t.ne[k];
}

/* staged batches: pair i's compact record carries (class << 24 | position within the class) in its `id` word
 * (written by the host thread that packed it); class_start[] is known once the batch is closed.  Deterministic
 * inverse: sorted[class_start[class] + position] = i. */
#define BSW_NCLASS 48
struct ClassStart { int start[BSW_NCLASS]; };
__global__ void class_scatter_kernel(const char *__restrict__ rec24, int n, ClassStart cs, uint32_t *__restrict__ sorted)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t id = reinterpret_cast<const uint32_t *>(rec24 + (size_t)i * 24)[2];
    sorted[cs.start[id >> 24] + (int)(id & 0xffffffu)] = (uint32_t)i;
}

/* ------------------------------------------------------------------ 4a. BSW packed */
#ifndef BSW_MINBLOCKS
#define BSW_MINBLOCKS 1
#endif
template <bool COUNT, bool SYM, bool QN>
__global__ void __launch_bounds__(32, BSW_MINBLOCKS)
bsw_packed_kernel(PairIO io, const uint8_t *__restrict__ ref,
                  const uint8_t *__restrict__ qer, const uint32_t *__restrict__ sorted_idx,
                  const __grid_constant__ BinTable tbl, Const c, int w, unsigned long long *cells)
{
    extern __shared__ __align__(16) uint8_t smem[];
    const int lane = threadIdx.x;
    int slot0, bin_end, NE;
    bintable_lookup(tbl, blockIdx.x, slot0, bin_end, NE);
    uint32_t *he = reinterpret_cast<uint32_t *>(smem) + lane;
    uint16_t *qs = reinterpret_cast<uint16_t *>(smem + (size_t)NE * 128) + lane;
    const int slot = slot0 + lane * 2;
    const int pa = slot < bin_end ? (int)sorted_idx[slot] : -1;
    const int pb = slot + 1 < bin_end ? (int)sorted_idx[slot + 1] : -1;
    /* per-row score tables by target base: byte q of lut_tbl[t] = score(t, q) */
    __shared__ uint32_t lut_tbl[8];
    if (lane < 5) lut_tbl[lane] = bsw_lutrow(MISW_OF(c), XW_OF(c), lut_hi(c), lane * 8u);
    __syncwarp();
    unsigned long long my_cells = 0;
    if (pa >= 0) {
        PairIn A = load_pair(io, pa), B = load_pair(io, pb);
        PairOut oa, ob;
        bsw_packed_thread2<32, SYM, QN, COUNT>(c, w, A, B, ref, qer, he, qs, lut_tbl, oa, ob);
        store_bsw(io, pa, oa);
        my_cells += (unsigned long long)oa.cells;
        if (pb >= 0) { store_bsw(io, pb, ob); my_cells += (unsigned long long)ob.cells; }
    }
    if (COUNT) {
        for (int o = 16; o; o >>= 1) my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o);
        if (lane == 0 && my_cells) atomicAdd(cells, my_cells);
    }
}

/* ------------------------------------------------------------------ 4c. BSW wide (lanes per pair, bsw_wide.cuh)
 * G lanes per pair, 128 threads per block; pair of group g = sorted slot first + g. */
/* Kernel shapes (G lanes x R registers x 2 halves).  0..7 are the configurations of bsw_wide.cuh (the throughput choice);
 * 8..10 are the latency alternates: the same columns on more lanes, i.e. a shorter chain per row, for batches too
 * small to fill the device anyway. */
#define BSW_WIDE_NSHAPE 11
static inline int wide_shape_lanes(int sh) { return sh <= 7 ? wide_cfg_lanes(sh) : sh == 8 ? 8 : sh == 9 ? 16 : 32; }
static inline int wide_shape_regs(int sh) { return sh <= 7 ? wide_cfg_regs(sh) : 5; }
/* (4,10) -> (8,5); (4,16), (8,9) -> (16,5) = 160 columns; (8,16), (16,9) -> (32,5) = 320 columns; larger: unchanged */
static inline int wide_latency_shape(int cfg) { return cfg == 0 ? 8 : cfg <= 2 ? 9 : cfg <= 4 ? 10 : cfg; }

#define BSW_WIDE_MAXSEG 8
struct WideTable {
    int nseg;
    int first_block[BSW_WIDE_MAXSEG + 1];
    int first_pair[BSW_WIDE_MAXSEG], count[BSW_WIDE_MAXSEG], shape[BSW_WIDE_MAXSEG];
};

template <int G, int R, bool SYM>
__device__ __forceinline__ void wide_block(const PairIO &io, const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                                           const uint32_t *__restrict__ sorted_idx, int first, int count, const Const &c, int w, int block)
{
    const int gid = (block * 128 + (int)threadIdx.x) / G, gl = (int)threadIdx.x % G;
    const int pi = gid < count ? (int)sorted_idx[first + gid] : -1;
    const PairIn P = load_pair(io, pi);
    PairOut o;
    bsw_wide_lane<G, R, SYM>(c, w, P, ref, qer, gl, o);
    if (pi >= 0 && gl == 0) store_bsw(io, pi, o);
}

/* several shapes in ONE launch: a small batch is two or three short segments, and launching them one after another
 * on a stream adds their latencies (each 50-150 us).  Two kernels by register need: R <= 10 (96 registers, 20 warps/SM)
 * and R = 16 (<= 168 registers, 12 warps/SM). */
template <bool SYM>
__global__ void __launch_bounds__(128)
bsw_wide_multi_kernel(PairIO io, const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                      const uint32_t *__restrict__ sorted_idx, const __grid_constant__ WideTable t, Const c, int w)
{
    int s = 0;
    while (s + 1 < t.nseg && (int)blockIdx.x >= t.first_block[s + 1]) ++s;
    const int block = (int)blockIdx.x - t.first_block[s], first = t.first_pair[s], count = t.count[s];
    switch (t.shape[s]) {
    case 0: wide_block<4, 10, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    case 2: wide_block<8, 9, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    case 4: wide_block<16, 9, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    case 6: wide_block<32, 9, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    case 8: wide_block<8, 5, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    case 9: wide_block<16, 5, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    default: wide_block<32, 5, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    }
}

template <bool SYM>
__global__ void __launch_bounds__(128, 3)
bsw_wide_multi16_kernel(PairIO io, const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                        const uint32_t *__restrict__ sorted_idx, const __grid_constant__ WideTable t, Const c, int w)
{
    int s = 0;
    while (s + 1 < t.nseg && (int)blockIdx.x >= t.first_block[s + 1]) ++s;
    const int block = (int)blockIdx.x - t.first_block[s], first = t.first_pair[s], count = t.count[s];
    switch (t.shape[s]) {
    case 1: wide_block<4, 16, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    case 3: wide_block<8, 16, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    case 5: wide_block<16, 16, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    default: wide_block<32, 16, SYM>(io, ref, qer, sorted_idx, first, count, c, w, block); break;
    }
}

/* one sorted range per configuration */
struct WideSeg { int cfg, first, count; };

/* launches the segments (configuration cfg of bsw_wide.cuh over sorted slots [first, first + count)); returns the number
 * of kernel launches.  latency = true picks the shapes with more lanes per pair. */
static inline int launch_wide_segs(const WideSeg *segs, int nseg, bool latency, bool sym, PairIO io, const uint8_t *ref,
                                   const uint8_t *qer, const uint32_t *sorted, const Const &c, int w, cudaStream_t s)
{
    int launches = 0;
    for (int pass = 0; pass < 2; ++pass) {                  /* pass 0: the R <= 10 shapes, pass 1: the R = 16 shapes */
        WideTable t; memset(&t, 0, sizeof t);
        int blocks = 0;
        auto flush = [&] {
            t.first_block[t.nseg] = blocks;
            if (blocks > 0) {
                if (pass == 0) {
                    if (sym) bsw_wide_multi_kernel<true><<<blocks, 128, 0, s>>>(io, ref, qer, sorted, t, c, w);
                    else bsw_wide_multi_kernel<false><<<blocks, 128, 0, s>>>(io, ref, qer, sorted, t, c, w);
                } else {
                    if (sym) bsw_wide_multi16_kernel<true><<<blocks, 128, 0, s>>>(io, ref, qer, sorted, t, c, w);
                    else bsw_wide_multi16_kernel<false><<<blocks, 128, 0, s>>>(io, ref, qer, sorted, t, c, w);
                }
                ++launches;
            }
            memset(&t, 0, sizeof t); blocks = 0;
        };
        for (int k = 0; k < nseg; ++k) {
            if (segs[k].count <= 0) continue;
            const int sh = latency ? wide_latency_shape(segs[k].cfg) : segs[k].cfg;
            if ((wide_shape_regs(sh) == 16) != (pass == 1)) continue;
            if (t.nseg == BSW_WIDE_MAXSEG) flush();
            const int per_block = 128 / wide_shape_lanes(sh);
            t.first_block[t.nseg] = blocks; t.first_pair[t.nseg] = segs[k].first; t.count[t.nseg] = segs[k].count; t.shape[t.nseg] = sh;
            blocks += (segs[k].count + per_block - 1) / per_block;
            t.nseg++;
        }
        flush();
    }
    return launches;
}

/* configuration that serves the packed bin qb (queries of up to qb * 16 - 4 bases) */
static inline int wide_cfg_of_bin(int qb) { return wide_cfg_for(qb * 16 - BSW_SWEEP_SLACK); }

/* ------------------------------------------------------------------ 5a. BSW general */
__global__ void bsw_general_kernel(PairIO io, const uint8_t *__restrict__ ref,
                                   const uint8_t *__restrict__ qer, const uint32_t *__restrict__ sorted_idx,
                                   int first, int count, Const c, int w, int2 *eh, unsigned long long *cells)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long my_cells = 0;
    if (g < count) {
        const int pi = (int)sorted_idx[first + g];
        PairIn P = load_pair(io, pi);
        PairOut o;
        bsw_general_thread<long long>(c, w, P, ref, qer, eh + g, (long long)count, o);
        store_bsw(io, pi, o);
        my_cells = (unsigned long long)o.cells;
    }
    for (int o = 16; o; o >>= 1) my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o);
    if ((threadIdx.x & 31) == 0 && my_cells) atomicAdd(cells, my_cells);
}

/* ------------------------------------------------------------------ kswv result write-back */
/* phase 0 (and the staged calls of either phase, whose host side applies the phase-1 rule): aln[pi] <- the pass result;
 * phase 1 on resident records: refine aln[regid] (regid is word 7 of the 56-byte record) */
__device__ __forceinline__ void kswv_store(bsw_kswr_t *__restrict__ aln, const PairIO &io,
                                           int pi, const KswOut &o, int phase)
{
    if (phase == 0) {
        bsw_kswr_t r; r.score = o.score; r.te = o.te; r.qe = o.qe; r.score2 = o.score2; r.te2 = o.te2; r.tb = -1; r.qb = -1;
        aln[pi] = r;
    } else {
        const int regid = reinterpret_cast<const int *>(io.in + (size_t)pi * io.in_stride)[7];
        bsw_kswr_t *a = &aln[regid];
        if (a->score == o.score) { a->tb = a->te - o.te; a->qb = a->qe - o.qe; }
    }
}

/* ------------------------------------------------------------------ 4b. kswv in registers (kswv_sys.cuh)
 * One warp per block = 32 / G groups of two pairs; group g of block b serves the sorted slots first + (b * 32 / G + g) * 2
 * and + 1.  Dynamic shared memory: the `b` lists, cap words per pair. */
template <int G, int C, bool SYM, bool QN, bool W = false>
__global__ void __launch_bounds__(32)
kswv_sys_kernel(PairIO io, const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer, bsw_kswr_t *__restrict__ aln,
                const uint32_t *__restrict__ sorted_idx, int first, int count, Const c, int is8, int phase, int cap,
                unsigned long long *cells)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint32_t *bl = reinterpret_cast<uint32_t *>(smem);
    constexpr int NG = 32 / G;
    const int lane = threadIdx.x, g = lane / G, gl = lane % G;
    const int rel = ((int)blockIdx.x * NG + g) * 2;
    const int pa = rel < count ? (int)sorted_idx[first + rel] : -1;
    const int pb = rel + 1 < count ? (int)sorted_idx[first + rel + 1] : -1;
    const PairIn A = load_pair(io, pa), B = load_pair(io, pb);
    KswOut oa, ob;
    kswv_sys_lane<G, C, SYM, QN, W>(c, is8, A, B, ref, qer, gl, bl + (size_t)(g * 2) * cap, bl + (size_t)(g * 2 + 1) * cap, oa, ob);
    unsigned long long my_cells = 0;
    if (gl == G - 1) {
        if (pa >= 0) { kswv_store(aln, io, pa, oa, phase); my_cells += (unsigned long long)oa.cells; }
        if (pb >= 0) { kswv_store(aln, io, pb, ob, phase); my_cells += (unsigned long long)ob.cells; }
    }
    for (int o = 16; o; o >>= 1) my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o);
    if (lane == 0 && my_cells) atomicAdd(cells, my_cells);
}

/* Columns per lane of a launch with 8 lanes per group that holds queries of up to max_len2 bases.  The common variant
 * (symmetric gap sums, no N in the query) has one instantiation per column count 1..32; the three others share four shapes. */
static inline int kswv_sys_cols(int max_len2, bool sym, bool qn)
{
    int c = (max_len2 + 7) >> 3;
    c = c < 1 ? 1 : c > 32 ? 32 : c;
    return (sym && !qn) ? c : (c + 7) & ~7;
}

template <int G, int C, bool SYM, bool QN, bool W = false>
static inline void launch_kswv_sys_shape(int first, int count, int max_rows, PairIO io, const uint8_t *ref, const uint8_t *qer,
                                         bsw_kswr_t *aln, const uint32_t *sorted, const Const &c, int is8, int phase,
                                         unsigned long long *cells, cudaStream_t s)
{
    constexpr int NG = 32 / G;
    const int cap = kswv_sys_list_cap(max_rows);
    const int grid = (count + 2 * NG - 1) / (2 * NG);
    const size_t smem = (size_t)NG * 2 * cap * sizeof(uint32_t);
    kswv_sys_kernel<G, C, SYM, QN, W><<<grid, 32, smem, s>>>(io, ref, qer, aln, sorted, first, count, c, is8, phase, cap, cells);
}

/* one launch over the sorted slots [first, first + count), all of whose queries have at most max_len2 bases.
 * lanes = lanes per group: 8 is the throughput shape; 16 and 32 put the same columns on more lanes (a row step is half /
 * a quarter as long, 2 / 4 times the warps per pair) for batches too small to fill the device anyway -- the staged
 * small-call route picks them by batch size (kswv_sys_lanes_for) and runs the whole batch -- all query lengths, with and
 * without N -- in ONE such launch (qn: the general score table for everybody).  Symmetric gap sums only. */
static inline int kswv_sys_lanes_for(int pairs)
{
    /* the device holds ~2400 warps of this kernel: more lanes per group while the batch stays within a wave or two.
     * Measured (tools/kswv_lanes_sweep.sh, device flight of one call of 150 x 300-500 bp pairs, 8 / 16 / 32 lanes): 1000 pairs
     * 514 / 207 / 147 us, 2452: 610 / 364 / 344, 5000: 697 / 355 / 483, 10000: 851 / 635 / 707
     * (BSW_CUDA_KSWV_L32 / BSW_CUDA_KSWV_L16: the largest batch on 32 / 16 lanes) */
    static const int l32 = getenv("BSW_CUDA_KSWV_L32") ? atoi(getenv("BSW_CUDA_KSWV_L32")) : 4096;
    static const int l16 = getenv("BSW_CUDA_KSWV_L16") ? atoi(getenv("BSW_CUDA_KSWV_L16")) : 20000;
    return pairs <= l32 ? 32 : pairs <= l16 ? 16 : 8;
}

static inline void launch_kswv_sys(int max_len2, bool sym, bool qn, int lanes, int first, int count, int max_rows, PairIO io, const uint8_t *ref,
                                   const uint8_t *qer, bsw_kswr_t *aln, const uint32_t *sorted, const Const &c, int is8, int phase,
                                   unsigned long long *cells, cudaStream_t s)
{
#define KSWV_SYS_LAUNCH(G, CC, SYM, QN) \
    launch_kswv_sys_shape<G, CC, SYM, QN>(first, count, max_rows, io, ref, qer, aln, sorted, c, is8, phase, cells, s)
#define KSWV_SYS_COARSE(SYM, QN) \
    switch (kswv_sys_cols(max_len2, SYM, QN)) { \
    case 8: KSWV_SYS_LAUNCH(8, 8, SYM, QN); break; case 16: KSWV_SYS_LAUNCH(8, 16, SYM, QN); break; \
    case 24: KSWV_SYS_LAUNCH(8, 24, SYM, QN); break; default: KSWV_SYS_LAUNCH(8, 32, SYM, QN); break; }
#define KSWV_SYS_CASE(G, CC) case CC: KSWV_SYS_LAUNCH(G, CC, true, false); break;
#define KSWV_SYS_CASEQ(G, CC) case CC: KSWV_SYS_LAUNCH(G, CC, true, true); break;
    if (sym && lanes >= 32) {
        if (qn) switch ((std::max(max_len2, 1) + 31) >> 5) {
            KSWV_SYS_CASEQ(32, 1) KSWV_SYS_CASEQ(32, 2) KSWV_SYS_CASEQ(32, 3) KSWV_SYS_CASEQ(32, 4) KSWV_SYS_CASEQ(32, 5) KSWV_SYS_CASEQ(32, 6) KSWV_SYS_CASEQ(32, 7)
            default: KSWV_SYS_LAUNCH(32, 8, true, true); break;
        } else switch ((std::max(max_len2, 1) + 31) >> 5) {
            KSWV_SYS_CASE(32, 1) KSWV_SYS_CASE(32, 2) KSWV_SYS_CASE(32, 3) KSWV_SYS_CASE(32, 4) KSWV_SYS_CASE(32, 5) KSWV_SYS_CASE(32, 6) KSWV_SYS_CASE(32, 7)
            default: KSWV_SYS_LAUNCH(32, 8, true, false); break;
        }
        return;
    }
    if (sym && lanes >= 16) {
        if (qn) switch (((std::max(max_len2, 1) + 31) >> 5) * 2) {       /* even column counts: 32-column steps */
            KSWV_SYS_CASEQ(16, 2) KSWV_SYS_CASEQ(16, 4) KSWV_SYS_CASEQ(16, 6) KSWV_SYS_CASEQ(16, 8) KSWV_SYS_CASEQ(16, 10) KSWV_SYS_CASEQ(16, 12) KSWV_SYS_CASEQ(16, 14)
            default: KSWV_SYS_LAUNCH(16, 16, true, true); break;
        } else switch (((std::max(max_len2, 1) + 31) >> 5) * 2) {
            KSWV_SYS_CASE(16, 2) KSWV_SYS_CASE(16, 4) KSWV_SYS_CASE(16, 6) KSWV_SYS_CASE(16, 8) KSWV_SYS_CASE(16, 10) KSWV_SYS_CASE(16, 12) KSWV_SYS_CASE(16, 14)
            default: KSWV_SYS_LAUNCH(16, 16, true, false); break;
        }
        return;
    }
    if (sym && !qn) {
        switch (kswv_sys_cols(max_len2, true, false)) {
        KSWV_SYS_CASE(8, 1) KSWV_SYS_CASE(8, 2) KSWV_SYS_CASE(8, 3) KSWV_SYS_CASE(8, 4) KSWV_SYS_CASE(8, 5) KSWV_SYS_CASE(8, 6) KSWV_SYS_CASE(8, 7) KSWV_SYS_CASE(8, 8)
        KSWV_SYS_CASE(8, 9) KSWV_SYS_CASE(8, 10) KSWV_SYS_CASE(8, 11) KSWV_SYS_CASE(8, 12) KSWV_SYS_CASE(8, 13) KSWV_SYS_CASE(8, 14) KSWV_SYS_CASE(8, 15) KSWV_SYS_CASE(8, 16)
        KSWV_SYS_CASE(8, 17) KSWV_SYS_CASE(8, 18) KSWV_SYS_CASE(8, 19) KSWV_SYS_CASE(8, 20) KSWV_SYS_CASE(8, 21) KSWV_SYS_CASE(8, 22) KSWV_SYS_CASE(8, 23) KSWV_SYS_CASE(8, 24)
        KSWV_SYS_CASE(8, 25) KSWV_SYS_CASE(8, 26) KSWV_SYS_CASE(8, 27) KSWV_SYS_CASE(8, 28) KSWV_SYS_CASE(8, 29) KSWV_SYS_CASE(8, 30) KSWV_SYS_CASE(8, 31)
        default: KSWV_SYS_LAUNCH(8, 32, true, false); break;
        }
    } else if (sym) { KSWV_SYS_COARSE(true, true) }
    else if (qn) { KSWV_SYS_COARSE(false, true) }
    else { KSWV_SYS_COARSE(false, false) }
#undef KSWV_SYS_CASE
#undef KSWV_SYS_CASEQ
#undef KSWV_SYS_COARSE
#undef KSWV_SYS_LAUNCH
}

/* the 16-bit tier: configuration cfg of kswv_sys.cuh (kswv_sysw_cols) over the sorted slots [first, first + count); always
 * the general score table (queries with and without N together) */
static inline void launch_kswv_sysw(int cfg, bool sym, int first, int count, int max_rows, PairIO io, const uint8_t *ref,
                                    const uint8_t *qer, bsw_kswr_t *aln, const uint32_t *sorted, const Const &c, int is8, int phase,
                                    unsigned long long *cells, cudaStream_t s)
{
#define KSWV_SYSW_LAUNCH(G, CC) \
    if (sym) launch_kswv_sys_shape<G, CC, true, true, true>(first, count, max_rows, io, ref, qer, aln, sorted, c, is8, phase, cells, s); \
    else launch_kswv_sys_shape<G, CC, false, true, true>(first, count, max_rows, io, ref, qer, aln, sorted, c, is8, phase, cells, s)
    switch (cfg) {
    case 0: KSWV_SYSW_LAUNCH(8, 8); break;   case 1: KSWV_SYSW_LAUNCH(8, 16); break;
    case 2: KSWV_SYSW_LAUNCH(8, 20); break;  case 3: KSWV_SYSW_LAUNCH(8, 24); break;
    case 4: KSWV_SYSW_LAUNCH(8, 32); break;  case 5: KSWV_SYSW_LAUNCH(16, 20); break;
    case 6: KSWV_SYSW_LAUNCH(16, 26); break; default: KSWV_SYSW_LAUNCH(16, 32); break;
    }
#undef KSWV_SYSW_LAUNCH
}

/* ------------------------------------------------------------------ 5b. kswv general */
__global__ void kswv_general_kernel(PairIO io, const uint8_t *__restrict__ ref,
                                    const uint8_t *__restrict__ qer, bsw_kswr_t *__restrict__ aln,
                                    const uint32_t *__restrict__ sorted_idx, int first, int count,
                                    Const c, int is8, int phase, int2 *he, int *rm, unsigned long long *cells)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long my_cells = 0;
    if (g < count) {
        const int pi = (int)sorted_idx[first + g];
        PairIn P = load_pair(io, pi);
        KswOut o;
        kswv_general_thread<long long>(c, is8, P, ref, qer, he + g, (long long)count, rm + g, (long long)count, o);
        kswv_store(aln, io, pi, o, phase);
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


/* ------------------------------------------------------------------ kswv hand-off on the device
 * ksw_align2 between its two passes (mem_sam_pe_batch does it on the host between the getScores calls): a hit with
 * KSW_XSTART and not (KSW_XSUBO and score < minsc) gets a second pass over the REVERSED prefixes target[0..te],
 * query[0..qe] with xtra = KSW_XSTOP | score; its result refines aln[i] (kswv_store, phase 1).
 * handoff_len: per pair the lengths of the two reversed prefixes (0: no second pass), rounded up to 4 for the scan;
 * handoff_fill: 8 lanes per pair write the reversed bases and the phase-1 record (regid = seqid = i). */
struct HandoffLens {
    const bsw_seqpair_t *p; const bsw_kswr_t *a; int which;       /* 0: 1 per qualifying pair, 1: target bytes, 2: query bytes */
    __host__ __device__ int operator()(int i) const
    {
        const int xtra = p[i].h0, sc = a[i].score;
        const bool ok = (xtra & KSW_XSTART_) != 0 && !((xtra & KSW_XSUBO_) != 0 && sc < (xtra & 0xffff));
        if (!ok) return 0;
        if (which == 0) return 1;
        const int l = (which == 1 ? a[i].te : a[i].qe) + 1;
        return l > 0 ? (l + 3) & ~3 : 0;
    }
};

__global__ void kswv_handoff_totals_kernel(const bsw_seqpair_t *__restrict__ pairs, const bsw_kswr_t *__restrict__ aln, int n,
                                           const int *__restrict__ off_k, const int *__restrict__ off_r, const int *__restrict__ off_q,
                                           int *__restrict__ totals)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int g = n - 1;
    const HandoffLens one{ pairs, aln, 0 }, lr{ pairs, aln, 1 }, lq{ pairs, aln, 2 };
    totals[0] = off_k[g] + one(g); totals[1] = off_r[g] + lr(g); totals[2] = off_q[g] + lq(g);
}

__global__ void kswv_handoff_fill_kernel(const bsw_seqpair_t *__restrict__ pairs, const bsw_kswr_t *__restrict__ aln, int n,
                                         const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                                         const int *__restrict__ off_k, const int *__restrict__ off_r, const int *__restrict__ off_q,
                                         bsw_seqpair_t *__restrict__ out, uint8_t *__restrict__ rref, uint8_t *__restrict__ rqer,
                                         int *__restrict__ totals)
{
    const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 3, lane = threadIdx.x & 7;
    if (g >= n) return;
    const HandoffLens one{ pairs, aln, 0 };
    const bool ok = one(g) != 0;
    (void)totals;
    if (!ok) return;
    const int l1 = aln[g].te + 1, l2 = aln[g].qe + 1, k = off_k[g];
    const long long r0 = pairs[g].idr, q0 = pairs[g].idq;
    uint8_t *dr = rref + off_r[g], *dq = rqer + off_q[g];
    for (int j = lane; j < l1; j += 8) dr[j] = ref[r0 + (l1 - 1 - j)];
    for (int j = lane; j < l2; j += 8) dq[j] = qer[q0 + (l2 - 1 - j)];
    if (lane == 0) {
        bsw_seqpair_t o;
        o.idr = off_r[g]; o.idq = off_q[g]; o.id = k; o.len1 = l1 > 0 ? l1 : 0; o.len2 = l2 > 0 ? l2 : 0;
        o.h0 = KSW_XSTOP_ | aln[g].score; o.seqid = g; o.regid = g;
        o.score = o.tle = o.gtle = o.qle = o.gscore = o.max_off = 0;
        out[k] = o;
    }
}

/* ------------------------------------------------------------------ CIGAR generation (global_core.cuh)
 * global2_dp_kernel: DP + op count of jobs [first, first + count); thread t keeps its rolling row in eh[slot * count + t]
 * and its direction words at z + zoff[t].  global2_emit_kernel: second backtrack, ops to cigar[base + coff[t] ...]. */
struct GlobalZWords {
    const GJob *j;
    __host__ __device__ long long operator()(int i) const { return global2_zwords(j[i].tlen, j[i].qlen, j[i].w); }
};
struct GlobalNCigar {
    const GJob *j;
    __host__ __device__ int operator()(int i) const { return j[i].n_cigar > 0 ? j[i].n_cigar : 0; }
};

__global__ void global2_dp_kernel(GJob *__restrict__ jobs, int count, const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                                  Const c, int2 *__restrict__ eh, int M, uint32_t *__restrict__ z, const long long *__restrict__ zoff,
                                  unsigned long long *cells)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long my_cells = 0;
    if (t < count) {
        GJob g = jobs[t];
        if (g.tlen < 0 || g.qlen < 0 || g.w < 0 || 2 * g.w + 2 > M) { jobs[t].score = 0; jobs[t].n_cigar = -1; }     /* contract: reported by the host */
        else {
            long long cl = 0;
            uint32_t *zt = z + zoff[t];
            g.score = global2_dp(c, g.tlen, g.qlen, g.w, ref + g.idr, qer + g.idq, eh + t, (long long)count, M, zt, cl);
            g.n_cigar = global2_backtrack<false>(g.tlen, g.qlen, g.w, zt, nullptr, 0);
            jobs[t].score = g.score; jobs[t].n_cigar = g.n_cigar;
            my_cells = (unsigned long long)cl;
        }
    }
    for (int o = 16; o; o >>= 1) my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o);
    if ((threadIdx.x & 31) == 0 && my_cells) atomicAdd(cells, my_cells);
}

__global__ void global2_emit_kernel(GJob *__restrict__ jobs, int count, const uint32_t *__restrict__ z, const long long *__restrict__ zoff,
                                    const int *__restrict__ coff, long long base, uint32_t *__restrict__ cigar, long long cap)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const GJob g = jobs[t];
    if (g.n_cigar < 0) return;
    const long long off = base + coff[t];
    jobs[t].cigar_off = (int)off;
    if (off + g.n_cigar <= cap) global2_backtrack<true>(g.tlen, g.qlen, g.w, z + zoff[t], cigar + off, g.n_cigar);
}

/* ------------------------------------------------------------------ resident reference
 * The reference lives in HBM in bwa's .pac layout: base x of the forward strand in byte x >> 2, bits ((~x & 3) << 1).
 * Coordinates are bwa's doubled ones (bns_get_seq / bns_fetch_seq): 0 <= x < l_pac is the forward strand, l_pac <= x <
 * 2 l_pac the reverse complement, base(x) = 3 - fwd(2 l_pac - 1 - x).  A pair's target is len1 bases starting at
 * rpos: read upwards (x = rpos + k) when rpos >= 0, downwards (x = ~rpos - k) when rpos < 0 -- the reversed slices
 * mem_chain2aln_across_reads_V2 builds for left extensions.  ref_expand_kernel writes the bases, one byte each, to
 * dst + toff[i] (toff = exclusive prefix sum of the lengths rounded up to 4) and points the device copy of the record at
 * them; the caller's idr is parked in toff[i] and put back by restore_idr_kernel before the records go home. */
struct Len1Rounded {
    const bsw_seqpair_t *p;
    __host__ __device__ int operator()(int i) const { const int l = p[i].len1; return l > 0 ? (l + 3) & ~3 : 0; }
};

__device__ __forceinline__ uint32_t pac_base(const uint8_t *__restrict__ pac, long long l_pac, long long x)
{
    const bool rc = x >= l_pac;
    const long long f = rc ? 2 * l_pac - 1 - x : x;
    const uint32_t b = (__ldg(pac + (f >> 2)) >> ((~(uint32_t)f & 3u) << 1)) & 3u;
    return rc ? 3u - b : b;
}

__global__ void ref_expand_kernel(bsw_seqpair_t *__restrict__ pairs, int n, const long long *__restrict__ rpos,
                                  int *__restrict__ toff, const uint8_t *__restrict__ pac, long long l_pac,
                                  uint8_t *__restrict__ dst, long long cap, Plan *plan)
{
    const int i = blockIdx.x * (blockDim.x >> 3) + (threadIdx.x >> 3), sub = threadIdx.x & 7;
    if (i >= n) return;
    const int len = pairs[i].len1, off = toff[i];
    const long long rp = rpos[i];
    const bool down = rp < 0;
    const long long x0 = down ? ~rp : rp, xl = down ? x0 - (len - 1) : x0 + (len - 1);
    const bool bad = len > 0 && (x0 < 0 || x0 >= 2 * l_pac || xl < 0 || xl >= 2 * l_pac);
    const bool over = len > 0 && (long long)off + ((len + 3) & ~3) > cap;
    __syncwarp();                                   /* every lane of the pair has read len / off before lane 0 rewrites the record */
    if (sub == 0) {
        if (bad) atomicAdd(&plan->bad_pairs, 1);
        else if (over) atomicAdd(&plan->oob_pairs, 1);
        toff[i] = pairs[i].idr;
        pairs[i].idr = off;
    }
    if (bad || over || len <= 0) return;
    uint32_t *out = reinterpret_cast<uint32_t *>(dst + off);
    for (int k4 = sub; k4 * 4 < len; k4 += 8) {
        uint32_t w = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int k = k4 * 4 + u;
            if (k < len) w |= pac_base(pac, l_pac, down ? x0 - k : x0 + k) << (8 * u);
        }
        out[k4] = w;
    }
}

__global__ void restore_idr_kernel(bsw_seqpair_t *__restrict__ pairs, int n, const int *__restrict__ saved)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pairs[i].idr = saved[i];
}

} /* namespace bswk */
