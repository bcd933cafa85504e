"""One process, one handle over two devices (BSW_CUDA_DEVICES=0,1): chunks dealt round-robin; parity vs oracle."""
import os, sys
os.environ["BSW_CUDA_DEVICES"] = os.environ.get("BSW_CUDA_DEVICES", "0,1")
os.environ.setdefault("BSW_CUDA_CHUNK_PAIRS", "20000")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, bswb200, ora, common
d = ora.gen_bsw(60000, seed=0x2D)
P, r, q = d["left"]
o = P.copy(); ora.bsw_batch(o, r, q, ora.make_params(), 100, 16)
sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
g = P.copy(); sw.getScores16(g, r, q, len(g), 1, 100)
assert not common.diff_fields(o, g, common.BSW_FIELDS).any()
print("bsw 2-device ok", sw.stats()["pairs"], len(P))
K, kr, kq = ora.gen_kswv(30000)
ao = np.zeros(len(K), ora.KSWR); ora.kswv_batch(K, kr, kq, ao, ora.make_params(), 1, 0, 16)
ks = bswb200.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
ag = np.zeros(len(K), bswb200.KSWR); ks.getScores8(K, kr, kq, ag, len(K), 1, 0)
assert (ao == ag).all()
p1, r1, q1 = ora.kswv_phase1_inputs(K, kr, kq, ao)
ora.kswv_batch(p1, r1, q1, ao, ora.make_params(), 1, 1, 16); ks.getScores8(p1, r1, q1, ag, len(p1), 1, 1)
assert (ao == ag).all()
print("kswv 2-device ok (phase 0 and 1)")
