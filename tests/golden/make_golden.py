"""Generates tests/golden/{bsw,kswv}_golden.json from the CPU oracle.

The reference snapshot holds no source, test or golden vector for this path
("parity unpinned"), so these fixtures cannot come from the
reference; they freeze the oracle's answers on the fork's documented case list
(R:GRAVITON4_SVE2_RESULTS.md:398-413: empty / max-length / all-N / repetitive /
unequal lengths) plus seeded random pairs, and are what the GPU box compares
the CUDA path with.   Run:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import ora      # noqa: E402
import common   # noqa: E402


def main():
    prm_kw = dict(o_del=6, e_del=1, o_ins=6, e_ins=1, zdrop=100, end_bonus=5, a=1, b=4)
    prm = ora.make_params(**prm_kw)
    cases = []
    pairs, ref, qer = common.edge_case_pairs()
    names = ["all_match", "all_match_long_target", "all_mismatch", "empty_query", "empty_target", "both_empty",
             "all_N", "repetitive", "homopolymer_127", "very_unequal", "deletion3", "insertion4", "single_cell",
             "N_in_query", "unrelated_h0_120", "unrelated_h0_3"]
    for i, p in enumerate(pairs):
        q = qer[p["idq"]:p["idq"] + p["len2"]]
        t = ref[p["idr"]:p["idr"] + p["len1"]]
        for w in (100, 7):
            r = ora.bsw_extend(q, t, prm, w, int(p["h0"]))
            cases.append(dict(name=f"{names[i]}_w{w}", query=q.tolist(), target=t.tolist(), w=w, h0=int(p["h0"]),
                              expect=[r[k] for k in ("score", "tle", "gtle", "qle", "gscore", "max_off")]))
    P, r_, q_ = ora.gen_random(40, 5, 90, 5, 140, h0min=1, h0max=90, sim=0.85, p_n=0.01, seed=2024)
    for i, p in enumerate(P):
        q = q_[p["idq"]:p["idq"] + p["len2"]]
        t = r_[p["idr"]:p["idr"] + p["len1"]]
        r = ora.bsw_extend(q, t, prm, 100, int(p["h0"]))
        cases.append(dict(name=f"random_{i}", query=q.tolist(), target=t.tolist(), w=100, h0=int(p["h0"]),
                          expect=[r[k] for k in ("score", "tle", "gtle", "qle", "gscore", "max_off")]))
    with open(os.path.join(HERE, "bsw_golden.json"), "w") as f:
        json.dump(dict(params=prm_kw, cases=cases), f, separators=(",", ":"))

    kcases = []
    P, r_, q_ = ora.gen_kswv(30, qlen=60, tmin=120, tmax=200, seed=77, p_contain=0.8, div=0.04)
    for i, p in enumerate(P):
        q = q_[p["idq"]:p["idq"] + p["len2"]]
        t = r_[p["idr"]:p["idr"] + p["len1"]].copy()
        if i % 3 == 0:                                       # tandem copy -> second-best hits
            h = len(t) // 2
            t[h:2 * h] = t[:h]
        for xtra in (int(p["h0"]), ora.KSW_XSUBO | ora.KSW_XSTART | 19):
            r, _ = ora.ksw_align2(q, t, prm, xtra)
            kcases.append(dict(name=f"kswv_{i}_{xtra:x}", query=q.tolist(), target=t.tolist(), xtra=xtra,
                               expect=[int(r[k]) for k in common.KSW_FIELDS]))
    with open(os.path.join(HERE, "kswv_golden.json"), "w") as f:
        json.dump(dict(params=prm_kw, cases=kcases), f, separators=(",", ":"))
    print(len(cases), "bsw cases,", len(kcases), "kswv cases")


if __name__ == "__main__":
    main()
