#!/usr/bin/env python
"""One call of each tracking-side entry point (C2, init, ComputeBoW) after a warm-up call, for an ncu launch list:
   ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/kernel_times.py"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("asd-slam_b200")
syn = pkg.synth
ctx = pkg.Context(0)
c = syn.config2()
F = pkg.DescriptorSet(ctx, c["desc_t"], c["kps_t"], c["grid"])
q = c["queries"]
mt = pkg.ORBmatcher(ctx, c["nnratio"], True)
for _ in range(2):
    mt.SearchByProjection_LocalMap(F, q["valid"], q["u"], q["v"], c["level"], c["view_cos"], q["blocks"], c["desc_q"],
                                   syn.scale_factors(), th=1.0, blocked0=c["blocked0"])
c = syn.config_init()
F1 = pkg.DescriptorSet(ctx, c["desc1"], c["kps1"], c["grid2"])
F2 = pkg.DescriptorSet(ctx, c["desc2"], c["kps2"], c["grid2"])
mt = pkg.ORBmatcher(ctx, 0.9, True)
for _ in range(2):
    mt.SearchForInitialization(F1, F2, c["prev_matched"].copy(), c["window"])
tree = syn.vocab_tree(k=10, L=5, seed_extra=30)
voc = pkg.ORBVocabulary(ctx, tree)
S = pkg.DescriptorSet(ctx, syn.vocab_queries(tree, 4000, seed_extra=31, sigma=0.08))
for direct in (0, 1):  # both descent kernels (option "voc_direct")
    ctx.set_option("voc_direct", direct)
    for _ in range(2):
        voc.transform(S, 4)
ctx.set_option("voc_direct", 0)
c = syn.config1()
A = pkg.DescriptorSet(ctx, c["desc_a"], c["kps_a"], c["grid"], c["fv_a"])
B = pkg.DescriptorSet(ctx, c["desc_b"], c["kps_b"], c["grid"], c["fv_b"])
mt = pkg.ORBmatcher(ctx, 0.7, True)
for _ in range(2):
    mt.SearchByBoW_KF_F(A, c["valid_a"], B)
ctx.close()
