#!/usr/bin/env python
"""Host-side breakdown of the bench's e2e step (C5 through the C ABI with host buffers): where the wall time goes."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

pkg = importlib.import_module("asd-slam_b200")
ctx = pkg.Context(0)
c = bench.make_inputs(pkg, 0)
kfs = c["kfs"]; n = bench.N_DESC; P = bench.N_PAIRS
masks = [k["valid"] for k in kfs]
h_desc = [ctx.pinned((n, 128), np.float32) for _ in kfs]
for h, k in zip(h_desc, kfs):
    h[...] = k["desc"]
S = n
h_match, h_dist, h_nm = ctx.pinned((P, S), np.int32), ctx.pinned((P, S), np.float32), ctx.pinned((P,), np.int32)
T = {}
def tick(name, t0):
    T.setdefault(name, []).append(time.perf_counter() - t0)
for it in range(8):
    ctx.synchronize()
    t = time.perf_counter(); t_all = t
    ss = [pkg.DescriptorSet(ctx, h, k["kps"], c["grid"], None, async_upload=True) for h, k in zip(h_desc, kfs)]
    tick("set_create_async x64", t); t = time.perf_counter()
    for s_, k in zip(ss, kfs):
        s_.set_featvec(k["fv"])
    tick("set_featvec x64", t); t = time.perf_counter()
    jb = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, ss, masks, c["ia"], c["ib"], c["nnratio"], c["check_ori"])
    tick("job_prepare (waits for the uploads)", t); t = time.perf_counter()
    jb.run()
    tick("job_run (launch only)", t); t = time.perf_counter()
    jb.fetch(h_match, h_dist, h_nm)
    tick("job_fetch (kernels + D2H)", t); t = time.perf_counter()
    jb.destroy()
    for s_ in ss:
        s_.release()
    tick("destroy/release", t)
    tick("TOTAL", t_all)
for k, v in T.items():
    print(f"{k:40s} {1e3 * np.median(v[2:]):8.3f} ms   (first {1e3 * v[0]:8.3f})")
