"""Seeded vocabulary-transform cases shared by make_vocab_golden.py (reference outputs) and tests/test_vocab.py.

Each case = (name, tree kwargs for synth.vocab_tree | "trained", n query descriptors, levelsup list). The "trained"
tree is the reference's own TemplatedVocabulary::create output and is stored in the golden file itself."""
import hashlib
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
pkg = importlib.import_module("asd-slam_b200")
synth = pkg.synth

TREE_KEYS = ("child_off", "child_idx", "desc", "weight", "word_id")

CASES = [
    ("k10L3", dict(k=10, L=3), 900, (0, 1, 2, 4)),
    ("k4L4_ragged_ties", dict(k=4, L=4, ragged=0.3, dup_children=True, seed_extra=1), 700, (0, 2, 4)),
    ("k40L2_idf_l2", dict(k=40, L=2, weighting=2, scoring=1, seed_extra=2), 600, (1, 4)),
    ("k3L5_tf_dot", dict(k=3, L=5, weighting=1, scoring=5, ragged=0.2, seed_extra=3), 500, (1, 4)),
    ("k10L5_binary_chi2", dict(k=10, L=5, weighting=3, scoring=2, ragged=0.5, seed_extra=4), 800, (4,)),
    ("k2L6_kl", dict(k=2, L=6, scoring=3, stop_frac=0.3, seed_extra=5), 300, (2, 4)),
]
TRAINED = ("trained_k5L3", dict(k=5, L=3, n_train=3000, n_images=10, seed=7), 500, (1, 2, 4))


def queries(tree, n, extra):
    q = synth.vocab_queries(tree, n, seed_extra=extra)
    q[5] = q[6]          # duplicate rows: same word twice (BowVector::addWeight accumulation)
    q[7] = tree["desc"][tree["child_idx"][0]]  # exactly on a level-1 centre
    return q


def training_set(spec):
    rng = synth.rng_for(11, spec["seed"])
    return synth.make_descriptors(rng, spec["n_train"])


def digest(tree, q):
    h = hashlib.sha256()
    for k in TREE_KEYS:
        h.update(np.ascontiguousarray(tree[k]).tobytes())
    h.update(np.ascontiguousarray(q).tobytes())
    return h.hexdigest()


def leaf_depth_reaches(tree, word_id, level):
    """True where the leaf of word_id lies at or below `level` (the reference's FeatureVector node is defined there)."""
    off, idx = tree["child_off"], tree["child_idx"]
    nn = len(off) - 1
    depth = np.zeros(nn, np.int32)
    for p in range(nn):  # node ids are created parents-first in both generators
        depth[idx[off[p]:off[p + 1]]] = depth[p] + 1
    leaves = np.nonzero(off[1:] == off[:-1])[0]
    leaf_of_word = np.zeros(len(leaves), np.int64)
    leaf_of_word[tree["word_id"][leaves]] = leaves
    return depth[leaf_of_word[word_id]] >= level
