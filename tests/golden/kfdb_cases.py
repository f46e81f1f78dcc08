"""Seeded KeyFrameDatabase cases shared by make_kfdb_golden.py (reference outputs) and tests/test_kfdb.py.
Each case = synth.config_kfdb arguments; queries come with the database. Edge cases: ties in the first common word and
common-word counts (small vocabulary), an empty database, all keyframes erased, a query sharing no word, keyframes with
empty BowVectors, neighbours that point at erased / connected keyframes."""
import numpy as np

CASES = {
    "trajectory_300": dict(n_kf=300, n_words=100000, words_per_kf=900, n_places=60, seed_extra=1),
    "small_vocab_ties": dict(n_kf=120, n_words=400, words_per_kf=60, n_places=12, seed_extra=2, erase_frac=0.1),
    "dense_150": dict(n_kf=150, n_words=20000, words_per_kf=1500, n_places=10, seed_extra=3, global_frac=0.4),
    "many_erased": dict(n_kf=80, n_words=5000, words_per_kf=200, n_places=8, seed_extra=4, erase_frac=0.6),
}


def build(pkg, name):
    db = pkg.synth.config_kfdb(**CASES[name])
    rng = np.random.default_rng(0xDB00 + len(name))
    if name == "small_vocab_ties":  # identical BowVectors (exact score ties) and two empty ones
        off, w, v = db["bow_off"], db["bow_word"], db["bow_val"]
        bows = [(w[off[s]:off[s + 1]].copy(), v[off[s]:off[s + 1]].copy()) for s in range(db["n_kf"])]
        bows[7] = bows[3]
        bows[8] = bows[3]
        bows[20] = (np.zeros(0, np.uint32), np.zeros(0, np.float64))
        bows[21] = (np.zeros(0, np.uint32), np.zeros(0, np.float64))
        db["bow_off"] = np.concatenate([[0], np.cumsum([len(a) for a, _ in bows])]).astype(np.int32)
        db["bow_word"] = np.concatenate([a for a, _ in bows]).astype(np.uint32)
        db["bow_val"] = np.concatenate([b for _, b in bows]).astype(np.float64)
    # extra queries: no shared word (words no stored keyframe has), an empty BowVector, a copy of a stored keyframe
    n = db["n_kf"]
    conn0 = np.zeros(n, np.uint8)
    unused = np.setdiff1d(np.arange(db["n_words"], dtype=np.uint32), db["bow_word"])[:45]  # (ids beyond the vocabulary
    if len(unused):                                                                          # index past mvInvertedFile)
        db["queries"].append(dict(word=unused.astype(np.uint32), val=np.full(len(unused), 1.0 / len(unused)), connected=conn0,
                                  min_score=0.01))
    db["queries"].append(dict(word=np.zeros(0, np.uint32), val=np.zeros(0, np.float64), connected=conn0, min_score=0.01))
    s = int(rng.integers(0, n))
    o = db["bow_off"]
    db["queries"].append(dict(word=db["bow_word"][o[s]:o[s + 1]].copy(), val=db["bow_val"][o[s]:o[s + 1]].copy(),
                              connected=conn0, min_score=0.0))
    return db
