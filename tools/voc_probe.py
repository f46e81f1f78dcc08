#!/usr/bin/env python
"""Vocabulary descent, warm-L2 kernel times for both descent kernels:
   ncu --cache-control none --clock-control none --metrics gpu__time_duration.sum -k regex:k_voc python tools/voc_probe.py"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("asd-slam_b200")
syn = pkg.synth
ctx = pkg.Context(0)
tree = syn.vocab_tree(k=10, L=5, seed_extra=30)
voc = pkg.ORBVocabulary(ctx, tree)
sizes = [int(x) for x in sys.argv[1:]] or [1000, 2000, 4000]
for n in sizes:
    S = pkg.DescriptorSet(ctx, syn.vocab_queries(tree, n, seed_extra=31, sigma=0.08))
    for mode in ("fast", "exact", "direct"):
        ctx.set_option("voc_direct", mode == "direct")
        ctx.set_option("voc_exact", mode == "exact")
        for _ in range(3):
            voc.transform(S, 4)
        print(n, mode, "exact_levels", ctx.stat("voc_exact_levels"), flush=True)
    S.release()
ctx.close()
