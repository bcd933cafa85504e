"""The C-ABI library loads and exports every symbol include/bsw_cuda.h declares
(no compute call: this runs in the GPU-less container)."""
import ctypes as C
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "bsw_cuda.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    names = set(re.findall(r"\b((?:bsw|kswv)_[A-Za-z0-9_]+)\s*\(", txt))
    return sorted(names)


def test_header_symbols_exported(bsw):
    L = C.CDLL(bsw.LIB_PATH)
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/bsw_cuda.h but not exported"
    assert set(declared) == set(bsw.EXPORTS)


def test_struct_layouts(bsw):
    # bandedSWA.h SeqPair: 14 x int32; ksw.h kswr_t: 7 x int32
    assert bsw.SEQPAIR.itemsize == 56 and bsw.KSWR.itemsize == 28
    assert bsw.SEQPAIR.names[8:] == ("score", "tle", "gtle", "qle", "gscore", "max_off")
    assert bsw.SEQPAIR.fields["score"][1] == 32
    assert bsw.KSWR.names == ("score", "te", "qe", "score2", "te2", "tb", "qb")


def test_version_and_device_count(bsw):
    assert b"sm_100a" in bsw.lib().bsw_version()
    assert bsw.device_count() >= 0


def test_product_does_not_touch_oracle():
    """The product path must not import, link or execute anything under oracle/ (comments may name the oracle's
    definitions, e.g. its DP-cell count; code may not reach it)."""
    import re
    pkg = os.path.join(ROOT, "bwa-mem2-arm_b200")
    banned = re.compile(r'#\s*include\s*"[^"]*oracle|liboracle|oracle/|\bimport\s+ora\b|\bfrom\s+ora\b|\bora\.')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".cu", ".cuh", ".h", ".py", ".cpp")) or f == "Makefile":
                txt = open(os.path.join(dirpath, f)).read()
                m = banned.search(txt)
                assert m is None, (f, m.group(0))
    mk = open(os.path.join(pkg, "Makefile")).read()
    assert "-lcudart" in mk and "oracle" not in mk


def test_shim_header_compiles_and_links(tmp_path):
    """bandedSWA_b200.h (the drop-in classes a bwa-mem2 tree would include) compiles as C++ and links against the
    library: the C5 harness is its translation unit."""
    import subprocess
    exe = tmp_path / "stream_c5"
    pkg = os.path.join(ROOT, "bwa-mem2-arm_b200")
    cmd = ["g++", "-O1", "-std=c++17", "-Wall", "-Werror", "-pthread", "-o", str(exe), os.path.join(ROOT, "tools", "stream_c5.cpp"),
           "-L" + pkg, "-lbswcuda", "-Wl,-rpath," + pkg, "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run([str(exe)], capture_output=True, text=True)     # no --in: usage, exit 2, no CUDA call
    assert r.returncode == 2 and "--in" in r.stderr
