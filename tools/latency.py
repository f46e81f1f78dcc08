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


def abi_closure(fn, *args):
    """The C-ABI call alone: arguments converted once (what a C++ caller passes), only `fn(*args)` is timed."""
    def call():
        rc = fn(*args)
        if rc != 0:
            raise RuntimeError(f"ABI call failed with {rc}")
    return call


def abi_proj(pkg, ctx, target, q, desc_q=None, q_set=None, blocked0=None, mode=1, th_accept=1.5, best_init=256.0, nnratio=0.8,
             check_ori=False, count_twice=False):
    import ctypes as C
    L = pkg._lib
    q = np.ascontiguousarray(q, dtype=L.PROJ_QUERY_DTYPE)
    p = L.ProjParams(int(mode), float(th_accept), float(best_init), float(nnratio), int(check_ori), int(count_twice))
    assign, dist, nm = np.empty(target.n, np.int32), np.empty(target.n, np.float32), C.c_int32()
    keep = (q, p, assign, dist, nm, desc_q, blocked0)
    call = abi_closure(ctx.lib.asdm_search_by_projection, ctx.h, target.h, L.ptr(q), len(q), L.ptr(desc_q),
                       q_set.h if q_set is not None else None, None, L.ptr(blocked0), C.byref(p), L.ptr(assign), L.ptr(dist),
                       C.byref(nm))
    call.keep = keep
    call.result = lambda: (nm.value, assign, dist)
    return call


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
    import ctypes as C
    L_ = pkg._lib
    va = np.ascontiguousarray(c["valid_a"], np.uint8)
    m1, d1, n1 = np.empty(B.n, np.int32), np.empty(B.n, np.float32), C.c_int32()
    r["abi_call_ms"] = timeit(abi_closure(ctx.lib.asdm_search_by_bow_kf_f, ctx.h, A.h, L_.ptr(va), B.h, 0.7, 1, L_.ptr(m1), L_.ptr(d1),
                                          C.byref(n1)), reps)["median_ms"]
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
    kw2 = dict(blocked0=c["blocked0"], mode=pkg.ASDM_PROJ_RATIO_LEVEL, th_accept=1.5, best_init=256.0, nnratio=c["nnratio"],
               count_twice=True, check_ori=False)
    r["abi_call_ms"] = timeit(abi_proj(pkg, ctx, F, q, desc_q=c["desc_q"], **kw2), reps)["median_ms"]
    # the same with the map-point descriptors gathered by the adaptor into page-locked memory (asdm_host_alloc): the
    # kernel reads the rows in place over PCIe, nothing is staged
    pinned_q = ctx.pinned(c["desc_q"].shape, np.float32)
    pinned_q[...] = c["desc_q"]
    call = abi_proj(pkg, ctx, F, q, desc_q=pinned_q, **kw2)
    r["abi_call_pinned_desc_ms"] = timeit(call, reps)["median_ms"]
    res_pin = call.result()
    # map-point descriptors resident in a device pool (they only change in ComputeDistinctiveDescriptors): no descriptor
    # crosses PCIe per frame, the call ships the 8000 query records only
    pool = pkg.DescriptorSet(ctx, c["desc_q"])
    res_pool = mt.SearchByProjection(F, q, q_set=pool, blocked0=c["blocked0"], mode=pkg.ASDM_PROJ_RATIO_LEVEL, th_accept=1.5,
                                     best_init=256.0, count_twice=True, check_ori=False)
    res_host = mt.SearchByProjection(F, q, desc_q=c["desc_q"], blocked0=c["blocked0"], mode=pkg.ASDM_PROJ_RATIO_LEVEL,
                                     th_accept=1.5, best_init=256.0, count_twice=True, check_ori=False)
    assert res_pool[0] == res_host[0] and np.array_equal(res_pool[1], res_host[1])
    assert res_pin[0] == res_host[0] and np.array_equal(res_pin[1], res_host[1])
    r["abi_call_resident_pool_ms"] = timeit(abi_proj(pkg, ctx, F, q, q_set=pool, **kw2), reps)["median_ms"]
    # k_proj_fused of the last call: [phase 1 of the resolving CTA, ticket, phase-2 set-up, sweeps, selection] ns + sweeps
    r["kernel_phases_ns"] = [ctx.stat(f"proj_ns_{k}") for k in range(6)]
    r["host_phases_ns"] = [ctx.stat(f"proj_host_ns_{k}") for k in range(3)]  # staging | launch | wait
    r["algorithmic_bytes"] = int(8000 * 512 + 2000 * 512 + 2000 * 16 + (64 * 48 * 4 + 2000 * 4) + 8000 * 24 + 2000 * 8)
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
    sfa = np.ascontiguousarray(s["scale_factors"], np.float32)
    bs, ds = np.empty(L.n, np.int32), np.empty(L.n, np.float32)
    r["abi_call_ms"] = timeit(abi_closure(ctx.lib.asdm_stereo_row_match, ctx.h, L.h, R.h, L_.ptr(sfa), len(sfa), int(s["n_rows"]),
                                          float(s["min_d"]), float(s["max_d"]), L_.ptr(bs), L_.ptr(ds)), reps)["median_ms"]
    out["C3_stereo_row_match_1500"] = r
    r = timeit(lambda: mt.SearchByProjection(Cur, f["queries"], q_set=L, blocked0=f["blocked0"], mode=pkg.ASDM_PROJ_BEST_ONLY,
                                             th_accept=1.5, best_init=100.0), reps)
    r["launches"] = ctx.stat("gpu_launches")
    r["cpu_oracle_ms"] = cpu_time(lambda: oracle.search_by_projection(f["desc_t"], f["kps_t"], f["grid"], f["queries"], f["desc_q"],
                                                                      f["blocked0"], f["mode"], f["th_accept"], f["best_init"],
                                                                      f["nnratio"], f["check_ori"], f["count_twice"]))
    r["abi_call_ms"] = timeit(abi_proj(pkg, ctx, Cur, f["queries"], q_set=L, blocked0=f["blocked0"], mode=pkg.ASDM_PROJ_BEST_ONLY,
                                       th_accept=1.5, best_init=100.0, nnratio=0.8, check_ori=True), reps)["median_ms"]
    r["kernel_phases_ns"] = [ctx.stat(f"proj_ns_{k}") for k in range(6)]
    r["host_phases_ns"] = [ctx.stat(f"proj_host_ns_{k}") for k in range(3)]
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
