#!/usr/bin/env python
"""Times the keyframe database (asdm_bowdb_*, kernel k_bowdb_query = K6 of DESIGN.md) on a synthetic map: n_kf keyframes
x words_per_kf BowVector entries from a vocabulary of n_words words (k=10, L=6 -> 1e6 words), queries = perturbed
copies of stored keyframes.  Prints one JSON line: ABI call latency of the device pass alone (asdm_bowdb_query) and of
DetectLoopCandidates, the algorithmic bytes (12 B per stored entry) and -- with --cpu -- the CPU oracle's
time for the same detection (checked equal).  Kernel time: run under `ncu --metrics gpu__time_duration.sum`.
usage: python tools/kfdb_probe.py [--kf 20000] [--words 1200] [--vocab 1000000] [--reps 20] [--cpu]"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kf", type=int, default=20000)
    ap.add_argument("--words", type=int, default=1200)
    ap.add_argument("--vocab", type=int, default=1000000)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--filter-bits", type=int, default=0, help="Bloom filter bits per query word (0: library default)")
    ap.add_argument("--bsearch", action="store_true", help="the binary-search kernel instead of the shared-memory hash table")
    a = ap.parse_args()
    pkg = importlib.import_module("asd-slam_b200")
    rng = np.random.default_rng(0xA5D0B0)
    ctx = pkg.Context(0)
    kdb = pkg.KeyFrameDatabase(ctx)
    ctx.set_option("bowdb_search", int(a.bsearch))
    if a.filter_bits:
        ctx.set_option("bowdb_filter_bits", a.filter_bits)
    # keyframes of one "place" share most of their words: 100 keyframes per place
    n_places = max(a.kf // 100, 1)
    place_words = [np.unique(rng.integers(0, a.vocab, int(a.words * 1.3))) for _ in range(n_places)]
    off, words, vals = [0], [], []
    t0 = time.perf_counter()
    for s in range(a.kf):
        pw = place_words[s % n_places]
        w = np.unique(np.concatenate([rng.choice(pw, int(a.words * 0.8), replace=False),
                                      rng.integers(0, a.vocab, int(a.words * 0.2))])).astype(np.uint32)
        v = rng.random(len(w))
        v /= v.sum()
        assert kdb.add(w, v) == s
        words.append(w), vals.append(v), off.append(off[-1] + len(w))
    t_add = time.perf_counter() - t0
    entries = off[-1]
    q_of = [int(x) for x in rng.integers(0, a.kf, a.reps)]
    connected = np.zeros(a.kf, np.uint8)
    neigh_off = np.arange(a.kf + 1, dtype=np.int32) * 2
    neigh_slot = np.stack([(np.arange(a.kf) + n_places) % a.kf, (np.arange(a.kf) + 2 * n_places) % a.kf], 1).astype(np.int32).ravel()
    kdb.query(words[0], vals[0])  # warm-up (staging, result buffers)
    tq, tl, cands = [], [], []
    for s in q_of:
        t0 = time.perf_counter()
        common, first, score = kdb.query(words[s], vals[s])
        tq.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        c = kdb.DetectLoopCandidates(words[s], vals[s], connected, 0.05, False, None, neigh_off, neigh_slot)
        tl.append(time.perf_counter() - t0)
        cands.append(c)
    out = {"tool": "kfdb_probe", "kernel": "k_bowdb_query" if a.bsearch else "k_bowdb_query_hash", "filter_bits_per_word": a.filter_bits or "default", "keyframes": a.kf, "entries": int(entries), "vocabulary_words": a.vocab,
           "algorithmic_bytes_per_query": int(entries) * 12, "add_us_per_keyframe": 1e6 * t_add / a.kf,
           "query_call_ms_median": 1e3 * float(np.median(tq)), "query_call_ms_min": 1e3 * float(np.min(tq)),
           "detect_loop_call_ms_median": 1e3 * float(np.median(tl)),
           "algorithmic_GBps_at_call_latency": entries * 12 / float(np.min(tq)) / 1e9,
           "candidates_per_query_mean": float(np.mean([len(c) for c in cands]))}
    if a.cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_py
        oracle_py.build()
        db = {"n_kf": a.kf, "bow_off": np.asarray(off, np.int32), "bow_word": np.concatenate(words),
              "bow_val": np.concatenate(vals), "alive": np.ones(a.kf, np.uint8), "global_flag": np.zeros(a.kf, np.uint8),
              "neigh_off": neigh_off, "neigh_slot": neigh_slot, "n_words": a.vocab}
        tc, same = [], True
        for s, c in list(zip(q_of, cands))[:5]:
            t0 = time.perf_counter()
            oc, _, _ = oracle_py.detect_loop_candidates(db, words[s], vals[s], connected, 0.05, False)
            tc.append(time.perf_counter() - t0)
            same = same and np.array_equal(oc, c)
        out["cpu_oracle_detect_loop_ms_median"] = 1e3 * float(np.median(tc))
        out["cpu_oracle_note"] = "flat-array oracle incl. building its inverted file per call"
        out["equal_to_oracle"] = bool(same)
    print(json.dumps(out))
    kdb.release()
    ctx.close()


if __name__ == "__main__":
    main()
