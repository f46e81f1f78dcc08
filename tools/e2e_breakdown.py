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

# ---- streamed sweep (pairs searched in readiness waves while later keyframes upload): host time per phase
print("\nstreamed sweep")
sw = pkg.StreamedPairSweep(ctx, pkg.ASDM_JOB_BOW_KF_KF, len(kfs), c["ia"], c["ib"], c["nnratio"], c["check_ori"])
T.clear()
for it in range(8):
    ctx.synchronize()
    t_all = time.perf_counter()
    sets, jobs, done = [], [], 0
    for (a, b), bound in zip(sw.slices, sw.bounds):
        t = time.perf_counter()
        while done < bound:
            sets.append(pkg.DescriptorSet(ctx, h_desc[done], kfs[done]["kps"], c["grid"], kfs[done]["fv"], async_upload=True))
            done += 1
        tick("sets (create + featvec)", t); t = time.perf_counter()
        sel = sw.order[a:b]
        jb = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets[:bound], masks[:bound], sw.ia[sel], sw.ib[sel], c["nnratio"], c["check_ori"])
        tick(f"prepare wave<{bound}", t); t = time.perf_counter()
        jb.run()
        tick("run", t)
        jobs.append((jb, a, b))
    for jb, a, b in jobs:
        t = time.perf_counter()
        jb.fetch(h_match[a:b], h_dist[a:b], h_nm[a:b])
        tick(f"fetch {a}:{b}", t)
        jb.destroy()
    for s_ in sets:
        s_.release()
    tick("TOTAL", t_all)
for k, v in T.items():
    v = np.asarray(v).reshape(8, -1).sum(axis=1) if len(v) > 8 else np.asarray(v)
    print(f"{k:40s} {1e3 * np.median(v[2:]):8.3f} ms")
