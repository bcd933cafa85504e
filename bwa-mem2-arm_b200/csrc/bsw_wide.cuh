/*
 * bsw_wide.cuh -- lanes-per-pair banded extension: G lanes of a warp sweep ONE pair row by row
 * (BandedPairWiseSW::getScores8/16; semantics = bwa ksw_extend2).
 *
 * Why a second mapping next to the thread-owns-two-pairs kernel (bsw_core2.cuh):
 *   latency   a thread walks its two pairs serially (0.3-0.9 ms for a 130 x 260 pair), which is the floor of every
 *             small call; here a 130-column row is 16 columns per lane, and a pair takes ~30 us;
 *   16 bits   the DP state is 16 bits wide and lives in REGISTERS, so scores up to 16383 and queries up to
 *             1023 columns stay on chip (the byte-packed shared-memory rows of the thread kernel stop at 255 / 252).
 *
 * Layout.  Lane l of the group owns the 2R columns [l*2R, (l+1)*2R): register r holds column l*2R + r in its
 * low and column l*2R + R + r in its high 16 bits, i.e. every lane runs TWO independent half-block chains in
 * one s16x2 register file: Hs[r] (H(i-1, j-1), the oracle's eh[j].h), E[r] (E(i, j)), Q[r] (PRMT selector
 * of the query base).  The cell arithmetic is the thread kernel's (PRMT score, IMAD + VIADDMNMX zero gate,
 * VIMNMX3, VIADD, 2 x VIADDMNMX.RELU).
 *
 * The row-direction gap.  In ksw_extend2 e and f open from M = Hdiag + s, not from h (A.1 note 3), so
 * f(i, j+1) = max(f(i, j) - e_ins, t_j) with t_j independent of f: a prefix maximum.  Every half-block first runs
 * its chain with f = 0 on entry (local pass); the true entry values follow from one exclusive max-scan over the
 * blocks' exit values with the ramp (block index) * R * e_ins added (log2 G shuffle steps), and h -- the only
 * thing f feeds -- is patched with one VIADDMNMX per register: h = max(h, f_in - r * e_ins) (post pass).
 *
 * Exactness (all relative to A.1):
 *   - cells are evaluated for the columns [lb, end) with lb = max(i - w, 0): left of the oracle's trimmed beg
 *     the state is zero, so those cells evaluate to zero (value-neutral), but lb itself is exact (band);
 *   - the registers of columns outside [lb, end) are never modified (the H input is masked, stores are
 *     bit-selects), which reproduces the stale entries of the oracle's rolling array, initial row included;
 *     E is stored unmasked: it is provably zero right of `end` and never read again left of lb;
 *   - eh[end].h = h of the last cell arrives through the one-column shift, eh[lb].h = 0 (or the first-column
 *     boundary in lane 0) through the destination mask;
 *   - row maximum with its LAST column, the right trim (last non-zero entry), z-drop, gscore / gtle and
 *     max_off are evaluated per row from group-wide reductions, in the oracle's order.
 * Admission (wide_eligible): len2 + 1 columns must fit the group, and (a + 1) * (h0 + min(len1, len2) * a) must
 * fit 15 bits (the zero gate multiplies).  Everything else goes to the general 32-bit kernel.
 *
 * __host__ __device__ like the other cores: tests/emu/warp_emu.h runs the same source with the G lanes of a
 * group as coroutines that meet at every shuffle / vote.
 */
#pragma once
#include "bsw_core.cuh"
#include "bsw_core2.cuh"

namespace bswk {

#if defined(__CUDA_ARCH__)
__device__ __forceinline__ uint32_t w_shfl_up(uint32_t v, int d, int G) { return __shfl_up_sync(0xffffffffu, v, d, G); }
__device__ __forceinline__ uint32_t w_shfl_xor(uint32_t v, int m, int G) { return __shfl_xor_sync(0xffffffffu, v, m, G); }
__device__ __forceinline__ uint32_t w_shfl(uint32_t v, int src, int G) { return __shfl_sync(0xffffffffu, v, src, G); }
__device__ __forceinline__ bool w_any(bool p) { return __any_sync(0xffffffffu, p) != 0; }
/* maximum over the G-lane group (REDUX.SYNC: one instruction instead of log2 G shuffle + max steps) */
__device__ __forceinline__ uint32_t w_redmax(uint32_t v, int G)
{
    unsigned lane; asm("mov.u32 %0, %%laneid;" : "=r"(lane));
    const unsigned mask = G >= 32 ? 0xffffffffu : ((1u << G) - 1u) << (lane & ~(unsigned)(G - 1));
    return __reduce_max_sync(mask, v);
}
__device__ __forceinline__ uint32_t p_addmax(uint32_t a, uint32_t b, uint32_t c) { return __viaddmax_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t p_minu(uint32_t a, uint32_t b) { return __vminu2(a, b); }
__device__ __forceinline__ int ctz32(uint32_t v) { return __ffs((int)v) - 1; }
#else
uint32_t w_shfl_up(uint32_t v, int d, int G);       /* tests/emu/warp_emu.h */
uint32_t w_shfl_xor(uint32_t v, int m, int G);
uint32_t w_shfl(uint32_t v, int src, int G);
bool w_any(bool p);
uint32_t w_redmax(uint32_t v, int G);
static inline uint32_t p_addmax(uint32_t a, uint32_t b, uint32_t c)
{ return pk(imax((int16_t)(lo16(a) + lo16(b)), lo16(c)), imax((int16_t)(hi16(a) + hi16(b)), hi16(c))); }
static inline uint32_t p_minu(uint32_t a, uint32_t b)
{ uint32_t l = (a & 0xffff) < (b & 0xffff) ? (a & 0xffff) : (b & 0xffff), h = (a >> 16) < (b >> 16) ? (a >> 16) : (b >> 16); return l | (h << 16); }
static inline int ctz32(uint32_t v) { return __builtin_ctz(v); }
#endif

#define BSW_WIDE_MAX_COLS 1024

/* the (G, R) configurations a pair is classified into, smallest first; columns = G * 2R.  These are the THROUGHPUT
 * shapes: as few lanes per pair as the registers allow (the per-row bookkeeping -- masks, scan, shift, reductions -- costs
 * ~320 instructions per warp whatever G is, so it is shared by 8 pairs at G = 4 and by one at G = 32).  Small batches
 * run the same columns on more lanes (bsw_kernels.cuh wide_latency_shape). */
#define BSW_WIDE_NCFG 8
BSW_HD int wide_cfg_lanes(int cfg) { return cfg <= 1 ? 4 : cfg <= 3 ? 8 : cfg <= 5 ? 16 : 32; }
BSW_HD int wide_cfg_regs(int cfg) { return cfg == 0 ? 10 : (cfg & 1) ? 16 : 9; }
BSW_HD int wide_cfg_cols(int cfg) { return wide_cfg_lanes(cfg) * 2 * wide_cfg_regs(cfg); }   /* 80 128 144 256 288 512 576 1024 */

/* smallest configuration that holds columns 0 .. len2, or -1 */
BSW_HD int wide_cfg_for(int len2)
{
    for (int k = 0; k < BSW_WIDE_NCFG; ++k) if (len2 + 1 <= wide_cfg_cols(k)) return k;
    return -1;
}

BSW_HD bool wide_eligible(const Const &c, int len1, int len2, int h0)
{
    if (len1 < 0 || len2 < 0 || h0 <= 0 || len2 + 1 > BSW_WIDE_MAX_COLS) return false;
    if (c.a > 120 || c.b > 120 || c.oe_del > 8000 || c.oe_ins > 8000 || c.e_ins > 1000) return false;
    const long long bound = (long long)h0 + (long long)mn(len1, len2) * c.a;
    return bound * (c.a + 1) <= 32767;
}

/* 0xffff in every half whose bit (bit r: low half, bit 16 + r: high half) is set in B */
BSW_HD uint32_t wide_mask(uint32_t B, int r)
{
#if defined(__CUDA_ARCH__)
    return p_perm(B << (15 - r), 0u, 0xBB99u);              /* sign replication of bytes 1 and 3 */
#else
    return (((B >> r) & 1u) ? 0x0000ffffu : 0u) | (((B >> (16 + r)) & 1u) ? 0xffff0000u : 0u);
#endif
}

/* bits [lo, hi) clamped to [0, R) */
BSW_HD uint32_t wide_bits(int lo, int hi, int R)
{
    lo = lo < 0 ? 0 : lo; hi = hi > R ? R : hi;
    return hi > lo ? ((1u << hi) - 1u) & ~((1u << lo) - 1u) : 0u;
}

/*
 * One lane of a G-lane group.  P is the group's pair (every lane passes the same record), gl the lane's index
 * in the group.  All lanes of a warp must call this together (shuffles and votes are warp-wide); a group whose
 * pair is invalid or finished keeps stepping without effect until every group of the warp is done.
 */
template <int G, int R, bool SYM>
BSW_HD void bsw_wide_lane(const Const &c, int w_in, const PairIn &P, const uint8_t *__restrict__ ref,
                          const uint8_t *__restrict__ qer, int gl, PairOut &out)
{
    constexpr int C = 2 * R;
    const int qlen = P.valid ? P.len2 : 0, tlen = P.valid ? P.len1 : 0, h0 = P.valid ? P.h0 : 0;
    const int wb = eff_band(c, qlen, w_in);
    const int base_lo = gl * C, base_hi = gl * C + R;

    uint32_t Hs[R], E[R], Q[R], h[R], K[R];
    /* initial row (A.1): H(-1, 0) = h0, H(-1, 1) = relu(h0 - oe_ins), then - e_ins per column; 0 right of the query */
#pragma unroll
    for (int r = 0; r < R; ++r) {
        int hv[2]; uint32_t qv[2];
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
            const int col = (hf ? base_hi : base_lo) + r;
            int v = col == 0 ? h0 : h0 - c.oe_ins - (col - 1) * c.e_ins;
            v = (col > qlen || v < 0) ? 0 : v;
            hv[hf] = v;
            uint32_t q = 0;
            if (P.valid && col < qlen) { q = qer[P.idq + col]; q = q > 4u ? 4u : q; }
            qv[hf] = q;
        }
        Hs[r] = pack2(hv[0], hv[1]); E[r] = 0u;
        Q[r] = bsw_selector<true>(qv[0], qv[1]);
    }

    const uint32_t MISW = ((uint32_t)(-c.b) & 0xffu) * 0x01010101u, AMBW = lut_hi(c), XW = ((uint32_t)c.a ^ (uint32_t)(-c.b)) & 0xffu;
    const uint32_t NOD = pack2(-c.oe_del, -c.oe_del), NOI = pack2(-c.oe_ins, -c.oe_ins);
    const uint32_t NED = pack2(-c.e_del, -c.e_del), NEI = pack2(-c.e_ins, -c.e_ins);
    const uint32_t GATE = (uint32_t)(c.a + 1);
    const uint32_t L1 = launder(AMBW);
    const int D = R * c.e_ins;
    const int NEGBIG = -0x3fffffff;

    int end = qlen, smax = h0, max_i = -1, max_j = -1, max_ie = -1, gscore = -1, max_off = 0;
    bool act = P.valid && tlen > 0;
    TWin tw;
    twin_init(tw, ref, P.idr, tlen, act);
    uint32_t win = 0;
    int i = 0;
    while (w_any(act)) {
        if ((i & 3) == 0) win = twin_refill(tw, i >> 2);
        const uint32_t L0 = bsw_lutrow(MISW, XW, AMBW, (win & 0xffu) * 8u);
        win >>= 8;
        const int lb = mx(i - wb, 0);
        const int endc = mn(mn(end, i + wb + 1), qlen);
        int bnd = h0 - (c.o_del + c.e_del * (i + 1)); bnd = bnd > 0 ? bnd : 0;       /* H(i, -1) */
        if (act && lb >= endc) {
            /* no cell in this row: m == 0, the pair stops here (A.1: gscore check with j = beg, then break) */
            const int h1 = lb == 0 ? bnd : 0;
            if (lb == qlen) { max_ie = gscore > h1 ? max_ie : i; gscore = gscore > h1 ? gscore : h1; }
            act = false;
        }
        const bool live = act;
        /* in-range columns [lb, endc) of this lane's two half-blocks */
        const uint32_t B = live ? (wide_bits(lb - base_lo, endc - base_lo, R) | (wide_bits(lb - base_hi, endc - base_hi, R) << 16)) : 0u;

        /* ---- local pass: both half-block chains with f = 0 on entry ---- */
        uint32_t f = 0u;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            K[r] = wide_mask(B, r);
            const uint32_t hs = Hs[r] & K[r];
            const uint32_t s = p_perm(L0, L1, Q[r]);
            const uint32_t M = p_addmin(hs, s, hs * GATE);
            h[r] = p_max3(M, E[r], f);
            if (SYM) {
                const uint32_t u = p_add(M, NOD);
                E[r] = p_addmax_relu(E[r], NED, u);
                f = p_addmax_relu(f, NEI, u);
            } else {
                E[r] = p_addmax_relu(M, NOD, p_add(E[r], NED));
                f = p_addmax_relu(M, NOI, p_add(f, NEI));
            }
        }

        /* ---- entry values of f: exclusive max-scan over the blocks (block b = 2 * lane + half) ---- */
        const int Alo = (int)(f & 0xffffu) + (2 * gl) * D, Ahi = (int)(f >> 16) + (2 * gl + 1) * D;
        int X = Alo > Ahi ? Alo : Ahi;
#pragma unroll
        for (int d = 1; d < G; d <<= 1) {
            const int y = (int)w_shfl_up((uint32_t)X, d, G);
            if (gl >= d) X = X > y ? X : y;
        }
        int Xex = (int)w_shfl_up((uint32_t)X, 1, G);
        if (gl == 0) Xex = NEGBIG;
        int fin_lo = Xex - (2 * gl - 1) * D; fin_lo = fin_lo > 0 ? fin_lo : 0;
        int fin_hi = (Xex > Alo ? Xex : Alo) - (2 * gl) * D; fin_hi = fin_hi > 0 ? fin_hi : 0;
        const uint32_t FIN = pack2(fin_lo, fin_hi);

        /* ---- post pass: patch h with the entry value, mask, row maximum, last in-range h ---- */
        uint32_t hmaxp = 0u, hl = 0u;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const uint32_t ramp = pack2(-r * c.e_ins, -r * c.e_ins);
            const uint32_t hh = p_addmax(FIN, ramp, h[r]);
            const uint32_t hm = hh & K[r];
            h[r] = hm;
            hmaxp = p_maxu(hmaxp, hm);
            hl = hm | (hl & ~K[r]);
        }

        /* ---- one-column shift into the state: eh[j + 1].h = h(i, j) for j in [lb - 1, endc) ---- */
        {
            const uint32_t up = w_shfl_up(h[R - 1], 1, G);
            /* destination column base_lo / base_hi is written iff it lies in [lb, endc] */
            const bool dlo = live && base_lo >= lb && base_lo <= endc, dhi = live && base_hi >= lb && base_hi <= endc;
            const uint32_t KD0 = (dlo ? 0x0000ffffu : 0u) | (dhi ? 0xffff0000u : 0u);
            uint32_t carry = (gl == 0 ? (uint32_t)(lb == 0 ? bnd : 0) : (up >> 16)) | (h[R - 1] << 16);
#pragma unroll
            for (int r = R - 1; r >= 1; --r) {
                const uint32_t KD = K[r] | K[r - 1];
                Hs[r] = h[r - 1] | (Hs[r] & ~KD);
            }
            Hs[0] = (carry & KD0) | (Hs[0] & ~KD0);
        }

        /* ---- row bookkeeping (A.1), identical in every lane of the group ---- */
        /* row maximum and its LAST column in one group reduction: key = m << 10 | column.  Per lane: the maximum of its
         * own columns, then per half a bit string of "differs from it" flags, register 0 first. */
        int m, mj;
        {
            const uint32_t ml = (uint32_t)mx((int)(hmaxp & 0xffffu), (int)(hmaxp >> 16));
            const uint32_t mm = ml * 0x00010001u;
            uint32_t acc = 0u;
#pragma unroll
            for (int r = 0; r < R; ++r) acc = acc * 2u + p_minu(h[r] ^ mm, 0x00010001u);
            const uint32_t rm = (1u << R) - 1u, eq_lo = ~acc & rm, eq_hi = (~acc >> 16) & rm;
            const int cand = eq_hi ? base_hi + R - 1 - ctz32(eq_hi) : base_lo + R - 1 - ctz32(eq_lo | (1u << (R - 1)));
            const uint32_t key = w_redmax(ml ? (ml << 10) | (uint32_t)cand : 0u, G);
            m = (int)(key >> 10); mj = (int)(key & 1023u);
        }
        int h1 = 0;
        {
            const int ce = live ? endc - 1 : 0;
            const uint32_t v = w_shfl(hl, ce / C, G);
            h1 = (ce % C) >= R ? (int)(v >> 16) : (int)(v & 0xffffu);
        }
        if (live && endc == qlen) { max_ie = gscore > h1 ? max_ie : i; gscore = gscore > h1 ? gscore : h1; }
        bool stop = false, scan = false;
        if (live) {
            stop = m == 0;
            if (!stop) {
                if (m > smax) {
                    int off = mj - i; off = off < 0 ? -off : off;
                    smax = m; max_i = i; max_j = mj; max_off = max_off > off ? max_off : off;
                } else if (c.zdrop > 0 && smax - m > c.zdrop) {
                    const int di = i - max_i, dj = mj - max_j;
                    if (di > dj) stop = smax - m - (di - dj) * c.e_del > c.zdrop;
                    else         stop = smax - m - (dj - di) * c.e_ins > c.zdrop;
                }
            }
            if (!stop && i + 1 >= tlen) stop = true;
            if (!stop) {
                if (h1 != 0) end = mn(endc + 2, qlen);
                else scan = true;
            }
        }
        if (w_any(scan)) {
            /* right trim with eh[end].h == 0: last non-zero entry among the columns [lb, endc) (m > 0 guarantees one) */
            uint32_t acc = 0u;
#pragma unroll
            for (int r = 0; r < R; ++r) acc = acc * 2u + p_minu((Hs[r] | E[r]) & K[r], 0x00010001u);
            const uint32_t nz_lo = acc & 0xffffu, nz_hi = acc >> 16;
            uint32_t cand = 0u;                      /* column + 1; 0: none in this lane */
            if (nz_hi) cand = (uint32_t)(base_hi + R - ctz32(nz_hi));
            else if (nz_lo) cand = (uint32_t)(base_lo + R - ctz32(nz_lo));
            cand = w_redmax(cand, G);
            if (scan) end = mn((int)cand + 1, qlen);
        }
        if (stop) act = false;
        ++i;
    }
    out.score = smax; out.qle = max_j + 1; out.tle = max_i + 1;
    out.gtle = max_ie + 1; out.gscore = gscore; out.max_off = max_off;
    out.cells = 0;
}

} /* namespace bswk */
