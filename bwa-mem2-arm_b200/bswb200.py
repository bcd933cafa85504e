"""Python host mirror of the reference's BSW / kswv operator interface.

Thin ctypes binding over libbswcuda.so (the C ABI in include/bsw_cuda.h).  The
class and method names, argument order and error behaviour follow bwa-mem2's
`BandedPairWiseSW` (src/bandedSWA.h) and `kswv` (src/kswv.h) as the reference
docs quote them (R:docs-archive/INTEGRATION_COMPLETE.md:57-71,95-97,
R:docs-archive/WEEK1_WRAPPER_COMPLETE.md:263-268).  There is no CPU path: if
the CUDA library is missing, import of the symbols fails; if no GPU is present,
constructing an object terminates the process with the library's error, as the
reference does on fatal errors.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BSW_CUDA_LIB") or os.path.join(_HERE, "libbswcuda.so")   # override: A/B builds (tools/ab.sh)

SEQPAIR = np.dtype([(n, "<i4") for n in (
    "idr", "idq", "id", "len1", "len2", "h0", "seqid", "regid",
    "score", "tle", "gtle", "qle", "gscore", "max_off")])
KSWR = np.dtype([(n, "<i4") for n in ("score", "te", "qe", "score2", "te2", "tb", "qb")])
GLOBALJOB = np.dtype([(n, "<i4") for n in ("idr", "idq", "tlen", "qlen", "w", "score", "n_cigar", "cigar_off")])
KSW_XBYTE, KSW_XSTOP, KSW_XSUBO, KSW_XSTART = 0x10000, 0x20000, 0x40000, 0x80000

# every symbol include/bsw_cuda.h declares (tests check the .so exports all of them)
EXPORTS = (
    "bsw_create", "bsw_destroy", "bsw_getScores8", "bsw_getScores16", "bsw_scalarBandedSWAWrapper",
    "kswv_create", "kswv_destroy", "kswv_getScores8", "kswv_getScores16",
    "bsw_dev_getScores", "bsw_dev_getScores_static", "bsw_dev_extendChained", "kswv_dev_getScores", "kswv_dev_align2", "kswv_align2", "kswv_align2Ref", "bsw_global2", "bsw_dev_global2", "bsw_host_alloc", "bsw_host_free",
    "bsw_get_stats", "bsw_reset_stats", "bsw_set_timing", "bsw_set_cell_counting",
    "kswv_get_stats", "kswv_reset_stats", "kswv_set_timing",
    "bsw_measure_dpx_peak", "bsw_device_count", "bsw_version",
    "bsw_arena_begin", "bsw_arena_end", "bsw_arena_defer", "bsw_ref_upload", "bsw_ref_free", "bsw_getScoresRef",
    "bsw_set_direct_min", "kswv_set_direct_min", "bsw_set_wide_max", "bsw_set_wide_policy", "bsw_engine_get_stats", "bsw_engine_reset_stats",
)


class Stats(C.Structure):
    _fields_ = [("calls", C.c_int64), ("pairs", C.c_int64), ("pairs_packed", C.c_int64),
                ("pairs_general", C.c_int64), ("pairs_wide", C.c_int64), ("cells", C.c_int64), ("kernel_launches", C.c_int64),
                ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64), ("kernel_ms", C.c_double)]

    def asdict(self):
        return {f: getattr(self, f) for f, _ in self._fields_}


class EngineStats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("batches", "pairs", "calls", "kernel_launches", "h2d_bytes", "d2h_bytes",
                                          "max_batch_pairs", "wide_batches", "scan_ns", "reserve_ns", "pack_ns",
                                          "wait_ns", "unpack_ns", "packers_ns", "submit_ns", "flight_ns")]

    def asdict(self):
        return {f: getattr(self, f) for f, _ in self._fields_}


_lib = None


def lib():
    """Loads libbswcuda.so; raises (never falls back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for this path)")
    L = C.CDLL(LIB_PATH)
    vp, i32, i8 = C.c_void_p, C.c_int32, C.c_int8
    L.bsw_create.restype = vp
    L.bsw_create.argtypes = [C.c_int] * 6 + [vp, i8, i8, C.c_int]
    L.bsw_destroy.argtypes = [vp]
    for f in (L.bsw_getScores8, L.bsw_getScores16):
        f.restype = None
        f.argtypes = [vp, vp, vp, vp, i32, C.c_uint16, i32]
    L.bsw_scalarBandedSWAWrapper.restype = None
    L.bsw_scalarBandedSWAWrapper.argtypes = [vp, vp, vp, vp, i32, C.c_int, i32]
    L.kswv_create.restype = vp
    L.kswv_create.argtypes = [C.c_int] * 4 + [i8, i8, C.c_int, i32, i32]
    L.kswv_destroy.argtypes = [vp]
    for f in (L.kswv_getScores8, L.kswv_getScores16):
        f.restype = None
        f.argtypes = [vp, vp, vp, vp, vp, i32, C.c_uint16, C.c_int]
    L.bsw_dev_getScores.restype = None
    L.bsw_dev_getScores.argtypes = [vp, vp, vp, vp, i32, i32, vp]
    L.bsw_dev_getScores_static.restype = C.c_int
    L.bsw_dev_getScores_static.argtypes = [vp, vp, vp, vp, i32, i32, i32, vp]
    L.bsw_dev_extendChained.restype = None
    L.bsw_dev_extendChained.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp, i32, i32, i32, vp]
    L.kswv_dev_getScores.restype = None
    L.kswv_dev_getScores.argtypes = [vp, vp, vp, vp, vp, i32, C.c_int, C.c_int, vp]
    L.bsw_global2.restype = C.c_int64
    L.bsw_global2.argtypes = [vp, vp, i32, vp, vp, vp, C.c_int64]
    L.bsw_dev_global2.restype = C.c_int64
    L.bsw_dev_global2.argtypes = [vp, vp, i32, vp, vp, vp, C.c_int64, i32, vp]
    L.kswv_align2.restype = None
    L.kswv_align2.argtypes = [vp, vp, vp, vp, vp, i32, C.c_int]
    L.kswv_align2Ref.restype = None
    L.kswv_align2Ref.argtypes = [vp, vp, vp, vp, vp, vp, i32, C.c_int]
    L.kswv_dev_align2.restype = None
    L.kswv_dev_align2.argtypes = [vp, vp, vp, vp, vp, i32, C.c_int, vp]
    L.bsw_host_alloc.restype = vp
    L.bsw_host_alloc.argtypes = [C.c_size_t]
    L.bsw_host_free.argtypes = [vp]
    for f in (L.bsw_get_stats, L.kswv_get_stats):
        f.argtypes = [vp, C.POINTER(Stats)]
    for f in (L.bsw_reset_stats, L.kswv_reset_stats):
        f.argtypes = [vp]
    for f in (L.bsw_set_timing, L.kswv_set_timing, L.bsw_set_cell_counting):
        f.argtypes = [vp, C.c_int]
    L.bsw_arena_begin.restype = None
    L.bsw_arena_begin.argtypes = [vp, vp, C.c_size_t, vp, C.c_size_t]
    L.bsw_arena_end.argtypes = [vp]
    L.bsw_arena_defer.restype = None
    L.bsw_arena_defer.argtypes = [vp, C.c_int]
    L.bsw_ref_upload.restype = vp
    L.bsw_ref_upload.argtypes = [vp, C.c_int64]
    L.bsw_ref_free.argtypes = [vp]
    L.bsw_getScoresRef.restype = None
    L.bsw_getScoresRef.argtypes = [vp, vp, vp, vp, vp, i32, i32]
    for f in (L.bsw_set_direct_min, L.kswv_set_direct_min, L.bsw_set_wide_max):
        f.argtypes = [vp, C.c_int]
    L.bsw_set_wide_policy.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    L.bsw_engine_get_stats.argtypes = [C.c_int, C.POINTER(EngineStats)]
    L.bsw_engine_reset_stats.argtypes = [C.c_int]
    L.bsw_measure_dpx_peak.restype = C.c_double
    L.bsw_measure_dpx_peak.argtypes = [C.c_int, C.c_int]
    L.bsw_device_count.restype = C.c_int
    L.bsw_version.restype = C.c_char_p
    _lib = L
    return L


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags.c_contiguous
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):          # torch tensor (host pinned or device)
        return C.c_void_p(a.data_ptr())
    return C.c_void_p(int(a))


def fill_scmat(a, b, ambig=-1):
    """bwa `bwa_fill_scmat` (the 5x5 `mat` the reference constructor receives)."""
    m = np.full((5, 5), ambig, dtype=np.int8)
    for i in range(4):
        for j in range(4):
            m[i, j] = a if i == j else -b
    return np.ascontiguousarray(m.reshape(-1))


class BandedPairWiseSW:
    """Mirror of bwa-mem2 `class BandedPairWiseSW` (src/bandedSWA.h).

    Constructor arguments in the reference's order:
    (o_del, e_del, o_ins, e_ins, zdrop, end_bonus, mat, w_match, w_mismatch, numThreads).
    """

    def __init__(self, o_del=6, e_del=1, o_ins=6, e_ins=1, zdrop=100, end_bonus=5, mat=None,
                 w_match=1, w_mismatch=4, numThreads=1):
        self._L = lib()
        self._mat = fill_scmat(w_match, abs(w_mismatch)) if mat is None else np.ascontiguousarray(mat, dtype=np.int8)
        self._h = self._L.bsw_create(o_del, e_del, o_ins, e_ins, zdrop, end_bonus, _ptr(self._mat),
                                     w_match, w_mismatch, numThreads)
        if not self._h:
            raise RuntimeError("bsw_create failed")

    def close(self):
        if getattr(self, "_h", None):
            self._L.bsw_destroy(self._h)
            self._h = None

    __del__ = close

    def getScores8(self, pairArray, seqBufRef, seqBufQer, numPairs=None, numThreads=1, w=100):
        n = len(pairArray) if numPairs is None else numPairs
        self._L.bsw_getScores8(self._h, _ptr(pairArray), _ptr(seqBufRef), _ptr(seqBufQer), n, numThreads, w)

    def getScores16(self, pairArray, seqBufRef, seqBufQer, numPairs=None, numThreads=1, w=100):
        n = len(pairArray) if numPairs is None else numPairs
        self._L.bsw_getScores16(self._h, _ptr(pairArray), _ptr(seqBufRef), _ptr(seqBufQer), n, numThreads, w)

    def scalarBandedSWAWrapper(self, pairArray, seqBufRef, seqBufQer, numPairs=None, nthreads=1, w=100):
        n = len(pairArray) if numPairs is None else numPairs
        self._L.bsw_scalarBandedSWAWrapper(self._h, _ptr(pairArray), _ptr(seqBufRef), _ptr(seqBufQer), n, nthreads, w)

    def dev_getScores(self, d_pairs, d_ref, d_qer, numPairs, w=100, stream=0):
        """Device-resident variant: arguments are device pointers / CUDA tensors."""
        self._L.bsw_dev_getScores(self._h, _ptr(d_pairs), _ptr(d_ref), _ptr(d_qer), numPairs, w, C.c_void_p(stream))

    def dev_getScores_static(self, d_pairs, d_ref, d_qer, numPairs, maxQueryLen, w=100, stream=0):
        """bsw_dev_getScores_static: the same on the library's static device scratch, no allocation; returns the
        C entry point's code (0 = done, -2 = beyond BSW_STATIC_MAX_ENTRIES, ...)."""
        return int(self._L.bsw_dev_getScores_static(self._h, _ptr(d_pairs), _ptr(d_ref), _ptr(d_qer), numPairs, maxQueryLen, w,
                                                    C.c_void_p(stream)))

    def dev_extendChained(self, d_left, d_lref, d_lqer, numLeft, d_right, d_rref, d_rqer, numRight, w=100, maxBandTry=2,
                          stream=0):
        """Caller protocol on resident pairs in one submission: left attempts with band retry,
        right h0 = left score by regid, right attempts."""
        self._L.bsw_dev_extendChained(self._h, _ptr(d_left), _ptr(d_lref), _ptr(d_lqer), numLeft, _ptr(d_right), _ptr(d_rref),
                                      _ptr(d_rqer), numRight, w, maxBandTry, C.c_void_p(stream))

    def arena_begin(self, seqBufRef, seqBufQer):
        """Declares the two sequence arenas of a batch (uploaded once; see bsw_arena_begin)."""
        self._L.bsw_arena_begin(self._h, _ptr(seqBufRef), seqBufRef.nbytes, _ptr(seqBufQer), seqBufQer.nbytes)

    def arena_end(self):
        self._L.bsw_arena_end(self._h)

    def arena_defer(self, on=True):
        """Large calls inside an arena bracket complete at arena_end (see bsw_arena_defer)."""
        self._L.bsw_arena_defer(self._h, 1 if on else 0)

    def getScoresRef(self, pairArray, ref, rpos, seqBufQer, numPairs=None, w=100):
        """Targets from the resident reference `ref` (a Reference object) at doubled-strand positions `rpos` (int64)."""
        n = len(pairArray) if numPairs is None else numPairs
        assert rpos.dtype == np.int64
        self._L.bsw_getScoresRef(self._h, _ptr(pairArray), ref._r, _ptr(rpos), _ptr(seqBufQer), n, w)

    def set_direct_min(self, pairs):
        self._L.bsw_set_direct_min(self._h, int(pairs))

    def set_wide_max(self, pairs):
        self._L.bsw_set_wide_max(self._h, int(pairs))

    def global2(self, jobs, seqBufRef, seqBufQer, cigar=None):
        """Batched ksw_global2 (CIGAR generation).  jobs: GLOBALJOB array, filled in place.  Returns the dense op array."""
        assert jobs.dtype == GLOBALJOB and jobs.flags.c_contiguous
        if cigar is None:
            cigar = np.zeros(int((jobs["tlen"].astype(np.int64) + jobs["qlen"]).sum()) + 16, dtype=np.uint32)
        tot = self._L.bsw_global2(self._h, _ptr(jobs), len(jobs), _ptr(seqBufRef), _ptr(seqBufQer), _ptr(cigar), len(cigar))
        if tot < 0:
            raise RuntimeError(f"cigar buffer too small: {-tot} ops needed")
        return cigar[:tot]

    def dev_global2(self, d_jobs, numJobs, d_ref, d_qer, d_cigar, cigarCap, maxBand, stream=0):
        return self._L.bsw_dev_global2(self._h, _ptr(d_jobs), numJobs, _ptr(d_ref), _ptr(d_qer), _ptr(d_cigar), int(cigarCap),
                                       int(maxBand), C.c_void_p(stream))

    def set_wide_policy(self, wide_max, wide_lat, wide_split):
        self._L.bsw_set_wide_policy(self._h, int(wide_max), int(wide_lat), int(wide_split))

    def stats(self):
        s = Stats()
        self._L.bsw_get_stats(self._h, C.byref(s))
        return s.asdict()

    def reset_stats(self):
        self._L.bsw_reset_stats(self._h)

    def set_timing(self, on=True):
        self._L.bsw_set_timing(self._h, int(on))

    def set_cell_counting(self, on=True):
        self._L.bsw_set_cell_counting(self._h, int(on))


class kswv:
    """Mirror of bwa-mem2 `class kswv` (src/kswv.h).

    Constructor: (o_del, e_del, o_ins, e_ins, w_match, w_mismatch, numThreads, maxRefLen, maxQerLen);
    getScores8/16(pairArray, seqBufRef, seqBufQer, aln, numPairs, numThreads, phase).
    """

    def __init__(self, o_del=6, e_del=1, o_ins=6, e_ins=1, w_match=1, w_mismatch=4, numThreads=1,
                 maxRefLen=2048, maxQerLen=256):
        self._L = lib()
        self._h = self._L.kswv_create(o_del, e_del, o_ins, e_ins, w_match, w_mismatch, numThreads, maxRefLen, maxQerLen)
        if not self._h:
            raise RuntimeError("kswv_create failed")

    def close(self):
        if getattr(self, "_h", None):
            self._L.kswv_destroy(self._h)
            self._h = None

    __del__ = close

    def getScores8(self, pairArray, seqBufRef, seqBufQer, aln, numPairs=None, numThreads=1, phase=0):
        n = len(pairArray) if numPairs is None else numPairs
        self._L.kswv_getScores8(self._h, _ptr(pairArray), _ptr(seqBufRef), _ptr(seqBufQer), _ptr(aln), n, numThreads, phase)

    def getScores16(self, pairArray, seqBufRef, seqBufQer, aln, numPairs=None, numThreads=1, phase=0):
        n = len(pairArray) if numPairs is None else numPairs
        self._L.kswv_getScores16(self._h, _ptr(pairArray), _ptr(seqBufRef), _ptr(seqBufQer), _ptr(aln), n, numThreads, phase)

    def dev_getScores(self, d_pairs, d_ref, d_qer, d_aln, numPairs, is8=1, phase=0, stream=0):
        self._L.kswv_dev_getScores(self._h, _ptr(d_pairs), _ptr(d_ref), _ptr(d_qer), _ptr(d_aln), numPairs,
                                   int(is8), int(phase), C.c_void_p(stream))

    def align2(self, pairArray, seqBufRef, seqBufQer, aln, numPairs=None, is8=1):
        """ksw_align2 on host buffers in one call: both passes and the hand-off on the device (kswv_align2)."""
        n = len(pairArray) if numPairs is None else numPairs
        self._L.kswv_align2(self._h, _ptr(pairArray), _ptr(seqBufRef), _ptr(seqBufQer), _ptr(aln), n, int(is8))

    def align2Ref(self, pairArray, ref, rpos, seqBufQer, aln, numPairs=None, is8=1):
        """kswv_align2 with the windows fetched on the device from the resident reference `ref` (a Reference object) at
        doubled-strand positions `rpos` (int64, as for getScoresRef)."""
        n = len(pairArray) if numPairs is None else numPairs
        assert rpos.dtype == np.int64
        self._L.kswv_align2Ref(self._h, _ptr(pairArray), ref._r, _ptr(rpos), _ptr(seqBufQer), _ptr(aln), n, int(is8))

    def dev_align2(self, d_pairs, d_ref, d_qer, d_aln, numPairs, is8=1, stream=0):
        """Forward pass, device-side hand-off and reverse pass of ksw_align2 in one submission (resident buffers)."""
        self._L.kswv_dev_align2(self._h, _ptr(d_pairs), _ptr(d_ref), _ptr(d_qer), _ptr(d_aln), numPairs, int(is8), C.c_void_p(stream))

    def stats(self):
        s = Stats()
        self._L.kswv_get_stats(self._h, C.byref(s))
        return s.asdict()

    def reset_stats(self):
        self._L.kswv_reset_stats(self._h)

    def set_timing(self, on=True):
        self._L.kswv_set_timing(self._h, int(on))

    def set_direct_min(self, pairs):
        self._L.kswv_set_direct_min(self._h, int(pairs))


class Reference:
    """2-bit packed reference resident on the GPU (bsw_ref_upload): `pac` in bwa's .pac layout, `l_pac` bases."""

    def __init__(self, pac, l_pac):
        self._L = lib()
        pac = np.ascontiguousarray(pac, dtype=np.uint8)
        self.l_pac = int(l_pac)
        self._r = self._L.bsw_ref_upload(_ptr(pac), self.l_pac)

    def close(self):
        if getattr(self, "_r", None):
            self._L.bsw_ref_free(self._r)
            self._r = None

    __del__ = close


def pack_pac(bases):
    """bases (uint8, values 0..3) -> bwa .pac bytes: base x in byte x >> 2, bits ((~x & 3) << 1)."""
    b = np.asarray(bases, dtype=np.uint8)
    n = len(b)
    pad = (-n) % 4
    if pad:
        b = np.concatenate([b, np.zeros(pad, np.uint8)])
    q = b.reshape(-1, 4)
    return ((q[:, 0] << 6) | (q[:, 1] << 4) | (q[:, 2] << 2) | q[:, 3]).astype(np.uint8)


def engine_stats(device=0):
    s = EngineStats()
    lib().bsw_engine_get_stats(device, C.byref(s))
    return s.asdict()


def engine_reset_stats(device=0):
    lib().bsw_engine_reset_stats(device)


def measure_dpx_peak(device=0, which=2):
    return lib().bsw_measure_dpx_peak(device, which)


def device_count():
    return lib().bsw_device_count()
