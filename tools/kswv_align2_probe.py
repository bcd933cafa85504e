"""Where does the kswv_align2 / kswv_align2Ref step of config 4 spend its time?  The 8-bit call and the 16-bit slice timed
apart, one after the other, and side by side from two host threads (two handles)."""
import json, os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, torch, bswb200, ora

keep = []
def pin(a):
    t = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True)
    o = t.numpy().view(a.dtype).reshape(a.shape); o[...] = a
    keep.append(t)
    return o

n = int(os.environ.get("PROBE_PAIRS", "1000000")); n16 = n // 20
K, kr, kq = ora.gen_kswv(n)
K16, kr16, kq16 = ora.gen_kswv(n16, qlen=250, tmin=400, tmax=700, seed=0xC5)
fw = np.concatenate([kr, kr16]); nn = fw > 3; fw[nn] = (np.nonzero(nn)[0] & 3).astype(np.uint8)
Rf = bswb200.Reference(bswb200.pack_pac(fw), len(fw))
hK, hkr, hkq, hK16, hkr16, hkq16 = pin(K), pin(fw[:len(kr)].copy()), pin(kq), pin(K16), pin(fw[len(kr):].copy()), pin(kq16)
rp8, rp16 = pin(K["idr"].astype(np.int64)), pin(K16["idr"].astype(np.int64) + len(kr))
ha, hb = pin(np.zeros(n, bswb200.KSWR)), pin(np.zeros(n16, bswb200.KSWR))
ks, ks2 = bswb200.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256), bswb200.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)

def timed(fn, reps=5):
    fn(); fn()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t0) / reps * 1e3

def both_threads(f8, f16):
    t = threading.Thread(target=f16); t.start(); f8(); t.join()

a8 = lambda: ks.align2(hK, hkr, hkq, ha, n, 1)
a16 = lambda: ks.align2(hK16, hkr16, hkq16, hb, n16, 0)
b16 = lambda: ks2.align2(hK16, hkr16, hkq16, hb, n16, 0)
r8 = lambda: ks.align2Ref(hK, Rf, rp8, hkq, ha, n, 1)
r16 = lambda: ks.align2Ref(hK16, Rf, rp16, hkq16, hb, n16, 0)
s16 = lambda: ks2.align2Ref(hK16, Rf, rp16, hkq16, hb, n16, 0)
out = {"pairs": n, "pairs16": n16,
       "align2_8bit_ms": timed(a8), "align2_16bit_ms": timed(a16), "align2_serial_ms": timed(lambda: (a8(), a16())),
       "align2_two_threads_ms": timed(lambda: both_threads(a8, b16)),
       "align2Ref_8bit_ms": timed(r8), "align2Ref_16bit_ms": timed(r16), "align2Ref_serial_ms": timed(lambda: (r8(), r16())),
       "align2Ref_two_threads_ms": timed(lambda: both_threads(r8, s16))}
ref8 = ha.copy(); r8(); out["ref_equals_bytes_8bit"] = bool((ha == ref8).all())
if os.environ.get("PROBE_SKIP_RESIDENT"):
    print(json.dumps(out), flush=True); sys.exit(0)
# the same work on resident buffers: kswv_dev_align2 (hand-off on the device) and the two getScores calls with the
# phase-1 inputs prepared beforehand (what bench.py's resident step times)
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()
prm = ora.make_params()
a0 = np.zeros(n, ora.KSWR); ora.kswv_batch(K, hkr, kq, a0, prm, 1, 0, 16)
p1, r1, q1 = ora.kswv_phase1_inputs(K, hkr, kq, a0)
pad = np.zeros(64, np.uint8)
dK, dkr, dkq = dev(K), dev(np.concatenate([hkr, pad])), dev(np.concatenate([kq, pad]))
dp1, dr1, dq1 = dev(p1), dev(np.concatenate([r1, pad])), dev(np.concatenate([q1, pad]))
da = torch.zeros(n * 28, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def d_align2(): ks.dev_align2(dK, dkr, dkq, da, n, 1, st); torch.cuda.synchronize()
def d_two(): ks.dev_getScores(dK, dkr, dkq, da, n, 1, 0, st); ks.dev_getScores(dp1, dr1, dq1, da, len(p1), 1, 1, st); torch.cuda.synchronize()
def d_p0(): ks.dev_getScores(dK, dkr, dkq, da, n, 1, 0, st); torch.cuda.synchronize()
out["resident_dev_align2_8bit_ms"] = timed(d_align2); out["resident_two_calls_8bit_ms"] = timed(d_two); out["resident_phase0_8bit_ms"] = timed(d_p0)
out["pairs_phase1"] = int(len(p1))
print(json.dumps(out), flush=True)
