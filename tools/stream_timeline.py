#!/usr/bin/env python
"""GPU timeline of the streamed C5 sweep: CUDA events on the context stream after every wave's sets and kernels."""
import ctypes as C, importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import torch
rt = C.CDLL("libcudart.so.12") if os.path.exists("/usr/local/cuda/lib64/libcudart.so.12") else C.CDLL("libcudart.so")
pkg = importlib.import_module("asd-slam_b200")
ctx = pkg.Context(0)
c = bench.make_inputs(pkg, 0)
kfs = c["kfs"]; n = bench.N_DESC; P = bench.N_PAIRS
masks = [k["valid"] for k in kfs]
h_desc = [ctx.pinned((n, 128), np.float32) for _ in kfs]
for h, k in zip(h_desc, kfs):
    h[...] = k["desc"]
h_match, h_dist, h_nm = ctx.pinned((P, n), np.int32), ctx.pinned((P, n), np.float32), ctx.pinned((P,), np.int32)
stream = C.c_void_p(ctx.stream())
kps_all, fv_all = [k["kps"] for k in kfs], [k["fv"] for k in kfs]
def ev():
    e = C.c_void_p(); rt.cudaEventCreate(C.byref(e)); return e
def rec(e):
    rt.cudaEventRecord(e, stream)
def ms(a, b):
    f = C.c_float(); rt.cudaEventElapsedTime(C.byref(f), a, b); return f.value
ALL = [[64], [32, 64], None] if len(sys.argv) < 2 else [[int(x) for x in a.split(',')] for a in sys.argv[1:]]
for bounds in ALL:
    sw = pkg.StreamedPairSweep(ctx, pkg.ASDM_JOB_BOW_KF_KF, len(kfs), c["ia"], c["ib"], c["nnratio"], c["check_ori"], bounds=bounds)
    for it in range(4):
        ctx.synchronize()
        e0 = ev(); rec(e0); t0 = time.perf_counter()
        marks = []
        sets, jobs, done = [], [], 0
        for (a, b), bound in zip(sw.slices, sw.bounds):
            if bound > done:
                sets.extend(pkg.DescriptorSet.create_batch(ctx, h_desc[done:bound], kps_all[done:bound], c["grid"], fv_all[done:bound]))
                done = bound
            e = ev(); rec(e); marks.append((f"sets<{bound} ready", e, time.perf_counter() - t0))
            sel = sw.order[a:b]
            jb = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets[:bound], masks[:bound], sw.ia[sel], sw.ib[sel], c["nnratio"], c["check_ori"])
            jb.run()
            e = ev(); rec(e); marks.append((f"wave<{bound} done ({b - a} pairs)", e, time.perf_counter() - t0))
            jb.fetch_async(h_match[a:b], h_dist[a:b], h_nm[a:b])
            e = ev(); rec(e); marks.append((f"   d2h<{bound} done", e, time.perf_counter() - t0))
            jobs.append((jb, a, b))
        for jb, a, b in jobs:
            jb.fetch(h_match[a:b], h_dist[a:b], h_nm[a:b]); jb.destroy()
        t1 = time.perf_counter() - t0
        for s_ in sets:
            s_.release()
    print(f"bounds {sw.bounds}: total {1e3 * t1:.3f} ms")
    for name, e, th in marks:
        print(f"   {name:32s} gpu {ms(e0, e):7.3f} ms   host queued at {1e3 * th:7.3f} ms")
