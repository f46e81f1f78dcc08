#!/usr/bin/env python
"""Per-call latency of the tracking-side entry points (BASELINE configs C1-C4 + initialisation) through the C ABI,
sets resident on the GPU, host buffers in / host results out; the CPU oracle (single thread, the reference matcher is
serial) timed on the same inputs beside it.  Prints one JSON object; `bench.py` embeds the same numbers.

usage: python tools/latency.py [--reps 200] [--no-cpu]
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timeit(fn, reps, warm=5):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        t = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t)
    ts = np.asarray(ts) * 1e3
    return {"median_ms": float(np.median(ts)), "p10_ms": float(np.percentile(ts, 10)), "p90_ms": float(np.percentile(ts, 90))}


def measure(pkg, ctx, reps=200, cpu=True, oracle=None):
    syn = pkg.synth
    out = {}

    def cpu_time(fn, n=3):
        if not cpu:
            return None
        fn()
        ts = []
        for _ in range(n):
            t = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t)
        return float(np.min(ts) * 1e3)

    # ---- C1: one frame pair, 2000 x 2000, SearchByBoW(KF, F) ratio 0.7 + rotation check (dense: one vocabulary node)
    c = syn.config1()
    A = pkg.DescriptorSet(ctx, c["desc_a"], c["kps_a"], c["grid"], c["fv_a"])
    B = pkg.DescriptorSet(ctx, c["desc_b"], c["kps_b"], c["grid"], c["fv_b"])
    mt = pkg.ORBmatcher(ctx, 0.7, True)
    r = timeit(lambda: mt.SearchByBoW_KF_F(A, c["valid_a"], B), reps)
    r["launches"] = ctx.stat("gpu_launches")
    r["cpu_oracle_ms"] = cpu_time(lambda: oracle.search_by_bow_kf_f(c["desc_a"], c["kps_a"], c["fv_a"], c["valid_a"], c["desc_b"],
                                                                    c["kps_b"], c["fv_b"], 0.7, 1))
    # including the per-frame upload of the current frame (asdm_set_create + featvec), as Tracking would pay it
    def c1_with_upload():
        Bn = pkg.DescriptorSet(ctx, c["desc_b"], c["kps_b"], c["grid"], c["fv_b"])
        mt.SearchByBoW_KF_F(A, c["valid_a"], Bn)
        Bn.release()
    r["with_frame_upload_ms"] = timeit(c1_with_upload, reps)["median_ms"]
    out["C1_frame_pair_2000x2000_SearchByBoW"] = r
    A.release(); B.release()

    # ---- C2: SearchByProjection(Frame, local map): 2000 keypoints vs 8000 map points (2000 in view)
    c = syn.config2()
    F = pkg.DescriptorSet(ctx, c["desc_t"], c["kps_t"], c["grid"])
    q = c["queries"]
    mt = pkg.ORBmatcher(ctx, c["nnratio"], True)
    sf = syn.scale_factors()
    r = timeit(lambda: mt.SearchByProjection_LocalMap(F, q["valid"], q["u"], q["v"], c["level"], c["view_cos"], q["blocks"],
                                                      c["desc_q"], sf, th=1.0, blocked0=c["blocked0"]), reps)
    r["launches"] = ctx.stat("gpu_launches")
    # the C-ABI call alone, on the query records a C++ adaptor would have filled (asdm_proj_query array built beforehand)
    r["abi_call_ms"] = timeit(lambda: mt.SearchByProjection(F, q, desc_q=c["desc_q"], blocked0=c["blocked0"], mode=pkg.ASDM_PROJ_RATIO_LEVEL,
                                                            th_accept=1.5, best_init=256.0, count_twice=True, check_ori=False), reps)["median_ms"]
    # map-point descriptors resident in a device pool (they only change in ComputeDistinctiveDescriptors): no descriptor
    # crosses PCIe per frame, the call ships the 8000 query records only
    pool = pkg.DescriptorSet(ctx, c["desc_q"])
    res_pool = mt.SearchByProjection(F, q, q_set=pool, blocked0=c["blocked0"], mode=pkg.ASDM_PROJ_RATIO_LEVEL, th_accept=1.5,
                                     best_init=256.0, count_twice=True, check_ori=False)
    res_host = mt.SearchByProjection(F, q, desc_q=c["desc_q"], blocked0=c["blocked0"], mode=pkg.ASDM_PROJ_RATIO_LEVEL,
                                     th_accept=1.5, best_init=256.0, count_twice=True, check_ori=False)
    assert res_pool[0] == res_host[0] and np.array_equal(res_pool[1], res_host[1])
    r["abi_call_resident_pool_ms"] = timeit(lambda: mt.SearchByProjection(F, q, q_set=pool, blocked0=c["blocked0"],
                                                                          mode=pkg.ASDM_PROJ_RATIO_LEVEL, th_accept=1.5, best_init=256.0,
                                                                          count_twice=True, check_ori=False), reps)["median_ms"]
    pool.release()
    r["cpu_oracle_ms"] = cpu_time(lambda: oracle.search_by_projection(c["desc_t"], c["kps_t"], c["grid"], q, c["desc_q"], c["blocked0"],
                                                                      c["mode"], c["th_accept"], c["best_init"], c["nnratio"],
                                                                      c["check_ori"], c["count_twice"]))
    out["C2_SearchByProjection_2000kp_vs_8000mp"] = r
    F.release()

    # ---- C3: EuRoC-shaped stereo row match + frame-to-frame projection search, 1500 keypoints per image
    c = syn.config3()
    s, f = c["stereo"], c["f2f"]
    L = pkg.DescriptorSet(ctx, s["desc_l"], s["kps_l"], f["grid"])
    R = pkg.DescriptorSet(ctx, s["desc_r"], s["kps_r"], f["grid"])
    Cur = pkg.DescriptorSet(ctx, f["desc_t"], f["kps_t"], f["grid"])
    mt = pkg.ORBmatcher(ctx, 0.8, True)
    r = timeit(lambda: mt.StereoRowMatch(L, R, s["scale_factors"], s["n_rows"], s["min_d"], s["max_d"]), reps)
    r["launches"] = ctx.stat("gpu_launches")
    r["cpu_oracle_ms"] = cpu_time(lambda: oracle.stereo_row_match(s["desc_l"], s["kps_l"], s["desc_r"], s["kps_r"], s["scale_factors"],
                                                                  s["n_rows"], s["min_d"], s["max_d"]))
    out["C3_stereo_row_match_1500"] = r
    r = timeit(lambda: mt.SearchByProjection(Cur, f["queries"], q_set=L, blocked0=f["blocked0"], mode=pkg.ASDM_PROJ_BEST_ONLY,
                                             th_accept=1.5, best_init=100.0), reps)
    r["launches"] = ctx.stat("gpu_launches")
    r["cpu_oracle_ms"] = cpu_time(lambda: oracle.search_by_projection(f["desc_t"], f["kps_t"], f["grid"], f["queries"], f["desc_q"],
                                                                      f["blocked0"], f["mode"], f["th_accept"], f["best_init"],
                                                                      f["nnratio"], f["check_ori"], f["count_twice"]))
    out["C3_frame_to_frame_projection_1500_th15"] = r
    for x in (L, R, Cur):
        x.release()

    # ---- C4: local-mapping triangulation, 20 covisible pairs x 2000 x 2000 as one batched job (prepare + run + fetch)
    c = syn.config4()
    kfs = c["kfs"]
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    masks = [k["has_mp"] for k in kfs]

    def c4():
        job = pkg.PairJob(ctx, pkg.ASDM_JOB_TRIANGULATION, sets, masks, c["ia"], c["ib"], 0.6, 0, c["F12"], c["epipoles"],
                          c["scale_factors"], c["level_sigma2"])
        job.run()
        job.fetch()
        job.destroy()
    r = timeit(c4, max(20, reps // 4))
    r["launches"] = ctx.stat("gpu_launches")
    r["cpu_oracle_ms"] = cpu_time(lambda: oracle.batch_triangulation(kfs, c["ia"], c["ib"], c["F12"], c["epipoles"], c["scale_factors"],
                                                                     c["level_sigma2"], 0, threads=1), n=1)
    r["keyframe_pairs_per_sec"] = 20.0 / (r["median_ms"] * 1e-3)
    out["C4_triangulation_20_pairs_2000x2000"] = r
    for x in sets:
        x.release()

    # ---- monocular initialisation: SearchForInitialization, 4000 keypoints, window 100
    c = syn.config_init()
    F1 = pkg.DescriptorSet(ctx, c["desc1"], c["kps1"], c["grid2"])
    F2 = pkg.DescriptorSet(ctx, c["desc2"], c["kps2"], c["grid2"])
    mt = pkg.ORBmatcher(ctx, 0.9, True)
    r = timeit(lambda: mt.SearchForInitialization(F1, F2, c["prev_matched"], c["window"]), max(20, reps // 4))
    r["launches"] = ctx.stat("gpu_launches")
    r["cpu_oracle_ms"] = cpu_time(lambda: oracle.search_for_initialization(c["desc1"], c["kps1"], c["desc2"], c["kps2"], c["grid2"],
                                                                           c["prev_matched"], c["window"], 0.9, 1), n=1)
    out["init_SearchForInitialization_4000_w100"] = r
    F1.release(); F2.release()

    # ---- Frame::ComputeBoW: vocabulary transform of 2000 / 4000 features, k=10 L=5 (DBoW2 default shape), levelsup 4
    tree = syn.vocab_tree(k=10, L=5, seed_extra=30)
    voc = pkg.ORBVocabulary(ctx, tree)
    for n in (2000, 4000):
        qd = syn.vocab_queries(tree, n, seed_extra=31, sigma=0.08)
        S = pkg.DescriptorSet(ctx, qd)
        r = timeit(lambda: voc.transform(S, 4), reps)
        r["launches"] = ctx.stat("gpu_launches")
        r["device_only_ms"] = timeit(lambda: (voc.transform(S, 4, host=False), ctx.synchronize()), reps)["median_ms"]
        r["cpu_oracle_ms"] = cpu_time(lambda: oracle.voc_transform(tree, qd, 4))
        out[f"ComputeBoW_transform_{n}_k10_L5"] = r
        S.release()
    voc.release()
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=200)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    pkg = importlib.import_module("asd-slam_b200")
    oracle = None
    if not a.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_py as oracle  # CPU baseline leg only
        oracle.build()
    ctx = pkg.Context(0)
    res = measure(pkg, ctx, a.reps, not a.no_cpu, oracle)
    ctx.close()
    print(json.dumps(res, indent=1))
