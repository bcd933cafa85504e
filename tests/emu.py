"""Builds and binds tests/emu/libbswemu.so: the kernels' per-thread DP compiled for the
host with emulated DPX intrinsics (TEST-ONLY; not part of the product library)."""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
_lib = None


def lib():
    global _lib
    if _lib is None:
        so = os.path.join(HERE, "emu", "libbswemu.so")
        src = os.path.join(HERE, "emu", "emu.cpp")
        deps = [src, os.path.join(HERE, "emu", "warp_emu.h")] + [os.path.join(ROOT, "bwa-mem2-arm_b200", "csrc", f)
                                                                 for f in ("bsw_core.cuh", "bsw_core2.cuh", "kswv_core.cuh", "kswv_sys.cuh", "bsw_wide.cuh", "global_core.cuh")]
        if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, src])
        _lib = C.CDLL(so)
        _lib.emu_bsw.restype = C.c_int
        _lib.emu_kswv.restype = C.c_int
        _lib.emu_bsw_wide.restype = C.c_int
        _lib.emu_global2.restype = C.c_longlong
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def bsw(pairs, ref, qer, w=100, o_del=6, e_del=1, o_ins=6, e_ins=1, zdrop=100, end_bonus=5, a=1, b=4,
        force_general=0, count=True):
    out = pairs.copy()
    cells = np.zeros(len(out), np.int64)
    npk = lib().emu_bsw(_p(out), _p(ref), _p(qer), len(out), o_del, e_del, o_ins, e_ins, zdrop, end_bonus,
                        a, b, -1, w, force_general, _p(cells) if count else None)
    return out, cells, npk


def kswv(pairs, ref, qer, aln, is8, phase, o_del=6, e_del=1, o_ins=6, e_ins=1, a=1, b=4, force_general=0):
    cells = np.zeros(len(pairs), np.int64)
    npk = lib().emu_kswv(_p(pairs), _p(ref), _p(qer), _p(aln), len(pairs), o_del, e_del, o_ins, e_ins, a, b, -1,
                         is8, phase, force_general, _p(cells))
    return cells, npk


def bsw_wide(pairs, ref, qer, w=100, o_del=6, e_del=1, o_ins=6, e_ins=1, zdrop=100, end_bonus=5, a=1, b=4, min_cfg=0, latency=False):
    """Lanes-per-pair kernel (bsw_wide.cuh) on an emulated warp.  Returns (results, mask of the pairs it admits).
    latency: the shapes with more lanes per pair that small batches run on."""
    out = pairs.copy()
    taken = np.zeros(len(out), np.uint8)
    lib().emu_bsw_wide(_p(out), _p(ref), _p(qer), len(out), o_del, e_del, o_ins, e_ins, zdrop, end_bonus, a, b, -1, w,
                       min_cfg | (0x100 if latency else 0), _p(taken))
    return out, taken.astype(bool)


def global2(jobs, ref, qer, o_del=6, e_del=1, o_ins=6, e_ins=1, a=1, b=4, nthreads=32):
    """CIGAR generation (global_core.cuh) with the library's scratch layout.  Returns (jobs, dense op array)."""
    out = jobs.copy()
    cap = int((out["tlen"].astype(np.int64) + out["qlen"]).sum()) + 16
    cig = np.zeros(cap, np.uint32)
    tot = lib().emu_global2(_p(out), len(out), _p(ref), _p(qer), o_del, e_del, o_ins, e_ins, a, b, -1, _p(cig), C.c_longlong(cap), nthreads)
    return out, cig[:tot]
