#!/usr/bin/env python
"""Generates tests/golden/vocab_golden.npz by running the REFERENCE's own vocabulary code (oracle/_ref/libref_vocab.so:
TemplatedVocabulary.h + FSift.cpp + BowVector.cpp + FeatureVector.cpp compiled from /root/reference by oracle/Makefile
`ref`) on the cases of vocab_cases.py.  Run in the container that mounts /root/reference:

    python tests/golden/make_vocab_golden.py

Synthetic trees go through the reference's load(); the "trained" tree is built by the reference's create() (k-means++,
rand() seeded) and stored in the file, since it cannot be regenerated without the reference."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle"))
import vocab_cases as V  # noqa: E402
import ref_py as ref  # noqa: E402


def run(out, name, voc, tree, n, levels, extra):
    q = V.queries(tree, n, extra)
    out[name + "/sha"] = np.frombuffer(bytes.fromhex(V.digest(tree, q)), np.uint8)
    for lu in levels:
        r = voc.transform(q, lu)
        for k in ("word_id", "fv_node", "fv_order", "bow_id", "bow_val"):
            out[f"{name}/{lu}/{k}"] = r[k]
        print(name, "levelsup", lu, "words", len(r["bow_id"]), "in fv", len(r["fv_order"]), "nodes", len(np.unique(r["fv_node"])))


def main():
    out = {}
    for i, (name, kw, n, levels) in enumerate(V.CASES):
        tree = V.synth.vocab_tree(**kw)
        voc = ref.RefVocabulary.load(tree)
        back = voc.export()
        assert all(np.array_equal(back[k], tree[k]) for k in V.TREE_KEYS), "load() did not reproduce the tree"
        run(out, name, voc, tree, n, levels, i)
        voc.free()
    name, spec, n, levels = V.TRAINED
    voc = ref.RefVocabulary.train(V.training_set(spec), spec["n_images"], spec["k"], spec["L"], seed=spec["seed"])
    tree = voc.export()
    for k in V.TREE_KEYS + ("k", "L", "weighting", "scoring"):
        out[f"{name}/tree/{k}"] = np.asarray(tree[k])
    run(out, name, voc, tree, n, levels, 100)
    voc.free()
    path = os.path.join(HERE, "vocab_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
