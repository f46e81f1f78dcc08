#!/usr/bin/env python
"""Per-call latency of the tracking-side entry points (BASELINE configs C1-C4 + initialisation + ComputeBoW) through the
C ABI: sets resident on the GPU, host buffers in / host results out, arguments converted once (what a C++ caller passes)
so that only the library call is timed.  The CPU oracle (single thread: the reference matcher is serial) is timed on the
same inputs beside it and its results are compared with the call's (indices, counts and distance bits) -> `parity`.
`bench.py` embeds C1-C4 in its JSON line.

usage: python tools/latency.py [--reps 200] [--no-cpu] [--only C1,C2,C3]
"""
import argparse
import ctypes as C
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


def bits(a):
    return np.asarray(a, np.float32).view(np.uint32)


def abi_closure(fn, *args):
    """The C-ABI call alone: arguments converted once, only `fn(*args)` is timed."""
    def call():
        rc = fn(*args)
        if rc != 0:
            raise RuntimeError(f"ABI call failed with {rc}")
    return call


def abi_proj(pkg, ctx, target, q, desc_q=None, q_set=None, blocked0=None, mode=1, th_accept=1.5, best_init=256.0, nnratio=0.8,
             check_ori=False, count_twice=False):
    L = pkg._lib
    q = np.ascontiguousarray(q, dtype=L.PROJ_QUERY_DTYPE)
    p = L.ProjParams(int(mode), float(th_accept), float(best_init), float(nnratio), int(check_ori), int(count_twice))
    assign, dist, nm = np.empty(target.n, np.int32), np.empty(target.n, np.float32), C.c_int32()
    call = abi_closure(ctx.lib.asdm_search_by_projection, ctx.h, target.h, L.ptr(q), len(q), L.ptr(desc_q),
                       q_set.h if q_set is not None else None, None, L.ptr(blocked0), C.byref(p), L.ptr(assign), L.ptr(dist),
                       C.byref(nm))
    call.keep = (q, p, assign, dist, nm, desc_q, blocked0)
    call.result = lambda: (nm.value, assign, dist)
    return call


def cpu_time(fn, n=3):
    res = fn()
    ts = []
    for _ in range(n):
        t = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t)
    return float(np.min(ts) * 1e3), res


def c1(pkg, ctx, reps, oracle):
    """C1: one frame pair, 2000 x 2000, SearchByBoW(KF, F) ratio 0.7 + rotation check (dense: one vocabulary node)."""
    syn, L_ = pkg.synth, pkg._lib
    c = syn.config1()
    A = pkg.DescriptorSet(ctx, c["desc_a"], c["kps_a"], c["grid"], c["fv_a"])
    B = pkg.DescriptorSet(ctx, c["desc_b"], c["kps_b"], c["grid"], c["fv_b"])
    va = np.ascontiguousarray(c["valid_a"], np.uint8)
    m1, d1, n1 = np.empty(B.n, np.int32), np.empty(B.n, np.float32), C.c_int32()
    call = abi_closure(ctx.lib.asdm_search_by_bow_kf_f, ctx.h, A.h, L_.ptr(va), B.h, 0.7, 1, L_.ptr(m1), L_.ptr(d1), C.byref(n1))
    r = {"ms_per_frame_pair": timeit(call, reps)["median_ms"], "launches": ctx.stat("gpu_launches")}
    mt = pkg.ORBmatcher(ctx, 0.7, True)

    def with_upload():  # including the per-frame upload of the current frame (asdm_set_create + featvec), as Tracking pays it
        Bn = pkg.DescriptorSet(ctx, c["desc_b"], c["kps_b"], c["grid"], c["fv_b"])
        mt.SearchByBoW_KF_F(A, c["valid_a"], Bn)
        Bn.release()
    r["ms_per_frame_pair_with_upload"] = timeit(with_upload, reps)["median_ms"]
    r["descriptor_pairs_per_sec"] = 2000.0 * 2000.0 / (r["ms_per_frame_pair"] * 1e-3)
    if oracle is not None:
        r["cpu_oracle_ms"], (onm, om, od) = cpu_time(lambda: oracle.search_by_bow_kf_f(c["desc_a"], c["kps_a"], c["fv_a"], c["valid_a"], c["desc_b"],
                                                                                       c["kps_b"], c["fv_b"], 0.7, 1), n=1)
        r["parity"] = bool(n1.value == onm and np.array_equal(m1, om) and np.array_equal(bits(d1), bits(od)))
    A.release(); B.release()
    return r


def c2(pkg, ctx, reps, oracle):
    """C2: SearchByProjection(Frame, local map): 2000 keypoints vs 8000 map points (2000 in view), th = 1."""
    syn = pkg.synth
    c = syn.config2()
    F = pkg.DescriptorSet(ctx, c["desc_t"], c["kps_t"], c["grid"])
    q = c["queries"]
    kw = dict(blocked0=c["blocked0"], mode=pkg.ASDM_PROJ_RATIO_LEVEL, th_accept=1.5, best_init=256.0, nnratio=c["nnratio"],
              count_twice=True, check_ori=False)
    r = {}
    # (a) the call as host/ORBmatcher_cuda.cc makes it: QueryPack gathers the MapPoint::GetDescriptor() rows into
    # page-locked memory (asdm_host_alloc) and the kernel reads them in place over PCIe
    pinned_q = ctx.pinned(c["desc_q"].shape, np.float32)
    pinned_q[...] = c["desc_q"]
    call = abi_proj(pkg, ctx, F, q, desc_q=pinned_q, **kw)
    r["abi_call_ms"] = timeit(call, reps)["median_ms"]
    res_pin = tuple(np.copy(x) for x in call.result())
    # (b) a caller that hands over rows in ordinary (pageable) memory: the library stages the valid rows through its
    # pinned arena first (one more host copy of the same megabyte)
    call = abi_proj(pkg, ctx, F, q, desc_q=c["desc_q"], **kw)
    r["abi_call_pageable_desc_ms"] = timeit(call, reps)["median_ms"]
    res_host = tuple(np.copy(x) for x in call.result())
    # (c) resident in a device pool (map-point descriptors only change in ComputeDistinctiveDescriptors): only the query
    # records cross PCIe
    pool = pkg.DescriptorSet(ctx, c["desc_q"])
    call = abi_proj(pkg, ctx, F, q, q_set=pool, **kw)
    r["abi_call_resident_pool_ms"] = timeit(call, reps)["median_ms"]
    res_pool = tuple(np.copy(x) for x in call.result())
    r["launches"] = ctx.stat("gpu_launches")
    # k_proj_fused of the last call: [phase 1 of the resolving CTA, ticket, phase-2 set-up, sweeps, selection] ns + sweeps
    r["kernel_phases_ns"] = [ctx.stat(f"proj_ns_{k}") for k in range(6)]
    r["host_phases_ns"] = [ctx.stat(f"proj_host_ns_{k}") for k in range(3)]  # staging | launch | wait
    r["algorithmic_bytes"] = int(8000 * 512 + 2000 * 512 + 2000 * 16 + (64 * 48 * 4 + 2000 * 4) + 8000 * 24 + 2000 * 8)
    pool.release()
    same = all(a[0] == res_host[0] and np.array_equal(a[1], res_host[1]) and np.array_equal(bits(a[2]), bits(res_host[2]))
               for a in (res_pin, res_pool))
    if oracle is not None:
        r["cpu_oracle_ms"], (onm, oa, od) = cpu_time(lambda: oracle.search_by_projection(
            c["desc_t"], c["kps_t"], c["grid"], q, c["desc_q"], c["blocked0"], c["mode"], c["th_accept"], c["best_init"], c["nnratio"],
            c["check_ori"], c["count_twice"]))
        r["parity"] = bool(same and res_host[0] == onm and np.array_equal(res_host[1], oa) and np.array_equal(bits(res_host[2]), bits(od)))
    else:
        r["variants_agree"] = bool(same)
    F.release()
    return r


def c3(pkg, ctx, reps, oracle):
    """C3: EuRoC-shaped stereo row match + frame-to-frame projection search (th = 15), 1500 keypoints per image."""
    syn, L_ = pkg.synth, pkg._lib
    c = syn.config3()
    s, f = c["stereo"], c["f2f"]
    L = pkg.DescriptorSet(ctx, s["desc_l"], s["kps_l"], f["grid"])
    R = pkg.DescriptorSet(ctx, s["desc_r"], s["kps_r"], f["grid"])
    Cur = pkg.DescriptorSet(ctx, f["desc_t"], f["kps_t"], f["grid"])
    sfa = np.ascontiguousarray(s["scale_factors"], np.float32)
    bs, ds = np.empty(L.n, np.int32), np.empty(L.n, np.float32)
    call = abi_closure(ctx.lib.asdm_stereo_row_match, ctx.h, L.h, R.h, L_.ptr(sfa), len(sfa), int(s["n_rows"]), float(s["min_d"]),
                       float(s["max_d"]), L_.ptr(bs), L_.ptr(ds))
    st = {"abi_call_ms": timeit(call, reps)["median_ms"], "launches": ctx.stat("gpu_launches"),
          "algorithmic_bytes": int(2 * 1500 * 512 + 2 * 1500 * 16 + 1500 * 8)}
    call2 = abi_proj(pkg, ctx, Cur, f["queries"], q_set=L, blocked0=f["blocked0"], mode=pkg.ASDM_PROJ_BEST_ONLY, th_accept=1.5,
                     best_init=100.0, nnratio=0.8, check_ori=True)
    ff = {"abi_call_ms": timeit(call2, reps)["median_ms"], "launches": ctx.stat("gpu_launches"),
          "kernel_phases_ns": [ctx.stat(f"proj_ns_{k}") for k in range(6)],
          "host_phases_ns": [ctx.stat(f"proj_host_ns_{k}") for k in range(3)],
          "algorithmic_bytes": int(2 * 1500 * 512 + 1500 * 16 + (64 * 48 * 4 + 1500 * 4) + 1500 * 32 + 1500 * 8)}
    if oracle is not None:
        st["cpu_oracle_ms"], (ob, od) = cpu_time(lambda: oracle.stereo_row_match(s["desc_l"], s["kps_l"], s["desc_r"], s["kps_r"], s["scale_factors"],
                                                                                 s["n_rows"], s["min_d"], s["max_d"]))
        st["parity"] = bool(np.array_equal(bs, ob) and np.array_equal(bits(ds), bits(od)))
        ff["cpu_oracle_ms"], (onm, oa, odd) = cpu_time(lambda: oracle.search_by_projection(
            f["desc_t"], f["kps_t"], f["grid"], f["queries"], f["desc_q"], f["blocked0"], f["mode"], f["th_accept"], f["best_init"], f["nnratio"],
            f["check_ori"], f["count_twice"]))
        nm, a, d = call2.result()
        ff["parity"] = bool(nm == onm and np.array_equal(a, oa) and np.array_equal(bits(d), bits(odd)))
    for x in (L, R, Cur):
        x.release()
    return {"stereo_row_match_1500": st, "frame_to_frame_projection_1500_th15": ff}


def c4(pkg, ctx, reps, oracle):
    """C4: local-mapping triangulation, 20 covisible pairs x 2000 x 2000 as one batched job (prepare + run + fetch)."""
    syn = pkg.synth
    c = syn.config4()
    kfs = c["kfs"]
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    masks = [k["has_mp"] for k in kfs]
    last = {}

    def call():
        job = pkg.PairJob(ctx, pkg.ASDM_JOB_TRIANGULATION, sets, masks, c["ia"], c["ib"], 0.6, 0, c["F12"], c["epipoles"],
                          c["scale_factors"], c["level_sigma2"])
        job.run()
        last["res"] = job.fetch()
        job.destroy()
    r = {"call_ms": timeit(call, max(20, reps // 4))["median_ms"], "launches": ctx.stat("gpu_launches")}
    r["keyframe_pairs_per_sec"] = 20.0 / (r["call_ms"] * 1e-3)
    if oracle is not None:
        r["cpu_oracle_ms"], (onm, om, od) = cpu_time(lambda: oracle.batch_triangulation(kfs, c["ia"], c["ib"], c["F12"], c["epipoles"], c["scale_factors"],
                                                                                       c["level_sigma2"], 0, threads=1), n=1)
        nm, m, d = last["res"]
        n = len(kfs[0]["kps"])
        r["parity"] = bool(np.array_equal(nm, onm) and np.array_equal(m[:, :n], om) and np.array_equal(bits(d[:, :n]), bits(od)))
    for x in sets:
        x.release()
    return r


def init(pkg, ctx, reps, oracle):
    """Monocular initialisation: SearchForInitialization, 4000 keypoints, window 100."""
    syn = pkg.synth
    c = syn.config_init()
    F1 = pkg.DescriptorSet(ctx, c["desc1"], c["kps1"], c["grid2"])
    F2 = pkg.DescriptorSet(ctx, c["desc2"], c["kps2"], c["grid2"])
    mt = pkg.ORBmatcher(ctx, 0.9, True)
    r = timeit(lambda: mt.SearchForInitialization(F1, F2, c["prev_matched"], c["window"]), max(20, reps // 4))
    r["launches"] = ctx.stat("gpu_launches")
    if oracle is not None:
        r["cpu_oracle_ms"], ores = cpu_time(lambda: oracle.search_for_initialization(c["desc1"], c["kps1"], c["desc2"], c["kps2"], c["grid2"],
                                                                                     c["prev_matched"], c["window"], 0.9, 1), n=1)
        nm, m, d, prev = mt.SearchForInitialization(F1, F2, c["prev_matched"], c["window"])
        r["parity"] = bool(nm == ores[0] and np.array_equal(m, ores[1]))
    F1.release(); F2.release()
    return r


def voc(pkg, ctx, reps, oracle):
    """Frame::ComputeBoW: vocabulary transform of 2000 / 4000 features, k=10 L=5 (DBoW2 default shape), levelsup 4."""
    syn = pkg.synth
    tree = syn.vocab_tree(k=10, L=5, seed_extra=30)
    v = pkg.ORBVocabulary(ctx, tree)
    out = {}
    for n in (2000, 4000):
        qd = syn.vocab_queries(tree, n, seed_extra=31, sigma=0.08)
        S = pkg.DescriptorSet(ctx, qd)
        r = timeit(lambda: v.transform(S, 4), reps)
        r["launches"] = ctx.stat("gpu_launches")
        r["device_only_ms"] = timeit(lambda: (v.transform(S, 4, host=False), ctx.synchronize()), reps)["median_ms"]
        if oracle is not None:
            r["cpu_oracle_ms"], _ = cpu_time(lambda: oracle.voc_transform(tree, qd, 4))
        out[f"{n}_features"] = r
        S.release()
    v.release()
    return out


PARTS = {"C1": ("C1_frame_pair_2000x2000_SearchByBoW", c1), "C2": ("C2_SearchByProjection_2000kp_vs_8000mp", c2),
         "C3": ("C3_euroc_1500", c3), "C4": ("C4_triangulation_20_pairs_2000x2000", c4),
         "init": ("init_SearchForInitialization_4000_w100", init), "voc": ("ComputeBoW_transform_k10_L5", voc)}


def measure(pkg, ctx, reps=200, oracle=None, only=None):
    out = {}
    for key, (name, fn) in PARTS.items():
        if only is None or key in only:
            out[name] = fn(pkg, ctx, reps, oracle)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=200)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    pkg = importlib.import_module("asd-slam_b200")
    oracle = None
    if not a.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_py as oracle  # CPU baseline leg only
        oracle.build()
    ctx = pkg.Context(0)
    res = measure(pkg, ctx, a.reps, oracle, set(a.only.split(",")) if a.only else None)
    ctx.close()
    print(json.dumps(res, indent=1))
