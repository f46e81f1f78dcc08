#!/usr/bin/env python
"""Generates tests/golden/kfdb_golden.npz: outputs of the REFERENCE's own KeyFrameDatabase.cc (compiled unmodified,
oracle/_ref/libref_kfdb.so -- only buildable where /root/reference is mounted) on the seeded cases of kfdb_cases.py.
Per case and query: DetectLoopCandidates (both only_global_map settings) -> candidate slots, mnLoopWords, mLoopScore bits;
DetectRelocalizationCandidates run as ONE sequence over all queries (mRelocScore persists) -> candidates + the final
mRelocScore bits.  usage: python tests/golden/make_kfdb_golden.py"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, HERE)


def run(impl, db):
    """impl: module with detect_loop_candidates / detect_reloc_candidates (ref_py or oracle_py)."""
    out = {}
    for qi, q in enumerate(db["queries"]):
        for og in (0, 1):
            cand, common, score = impl.detect_loop_candidates(db, q["word"], q["val"], q["connected"], q["min_score"], bool(og))
            out[f"loop{qi}_{og}_cand"] = cand.astype(np.int32)
            if og == 0:
                out[f"loop{qi}_common"] = common.astype(np.int32)
                out[f"loop{qi}_score"] = score.view(np.uint32).copy()
    rs = np.zeros(db["n_kf"], np.float32)
    for qi, q in enumerate(db["queries"]):
        cand, common = impl.detect_reloc_candidates(db, q["word"], q["val"], rs, bool(qi & 1))
        out[f"reloc{qi}_cand"] = cand.astype(np.int32)
        out[f"reloc{qi}_common"] = common.astype(np.int32)
    out["reloc_score_final"] = rs.view(np.uint32).copy()
    return out


def main():
    import kfdb_cases
    import ref_py
    pkg = importlib.import_module("asd-slam_b200")
    assert ref_py.kfdb_available(), "needs /root/reference (oracle/Makefile ref)"
    gold = {}
    for name in kfdb_cases.CASES:
        db = kfdb_cases.build(pkg, name)
        for k, v in run(ref_py, db).items():
            gold[f"{name}/{k}"] = v
        print(name, "queries", len(db["queries"]), "loop candidates",
              [len(gold[f"{name}/loop{qi}_0_cand"]) for qi in range(len(db["queries"]))])
    np.savez_compressed(os.path.join(HERE, "kfdb_golden.npz"), **gold)


if __name__ == "__main__":
    main()
