#!/usr/bin/env python
"""How fast do 64 keyframe uploads go through asdm_set_create_async? (PCIe floor: 131 MB / ~50 GB/s = 2.6 ms)"""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("asd-slam_b200")
ctx = pkg.Context(0)
c = bench.make_inputs(pkg, 0)
kfs = c["kfs"]; n = bench.N_DESC
h_desc = [ctx.pinned((n, 128), np.float32) for _ in kfs]
for h, k in zip(h_desc, kfs):
    h[...] = k["desc"]
def run(name, fn, reps=6):
    ts = []
    for _ in range(reps):
        ctx.synchronize(); t = time.perf_counter()
        ss, t_issue = fn()
        ctx.synchronize(); ts.append((time.perf_counter() - t, t_issue - t))
        for s in ss: s.release()
    tot = np.median([a for a, _ in ts[2:]]); iss = np.median([b for _, b in ts[2:]])
    print(f"{name:44s} issue {1e3*iss:6.3f} ms  done {1e3*tot:6.3f} ms  -> {131.072e6*len(kfs)/64/tot/1e9:5.1f} GB/s")
def full():
    ss = [pkg.DescriptorSet(ctx, h, k["kps"], c["grid"], k["fv"], async_upload=True) for h, k in zip(h_desc, kfs)]
    return ss, time.perf_counter()
def nofv():
    ss = [pkg.DescriptorSet(ctx, h, k["kps"], c["grid"], None, async_upload=True) for h, k in zip(h_desc, kfs)]
    return ss, time.perf_counter()
def desc_only():
    ss = [pkg.DescriptorSet(ctx, h, None, None, None, async_upload=True) for h, k in zip(h_desc, kfs)]
    return ss, time.perf_counter()
run("desc + kps + grid + featvec", full)
run("desc + kps + grid", nofv)
run("desc only (pack kernel)", desc_only)
kps_all, fv_all = [k["kps"] for k in kfs], [k["fv"] for k in kfs]
def batch():
    ss = pkg.DescriptorSet.create_batch(ctx, h_desc, kps_all, c["grid"], fv_all)
    return ss, time.perf_counter()
def batch_desc_only():
    ss = pkg.DescriptorSet.create_batch(ctx, h_desc)
    return ss, time.perf_counter()
def batch4x16():
    ss = []
    for a in range(0, 64, 16):
        ss += pkg.DescriptorSet.create_batch(ctx, h_desc[a:a + 16], kps_all[a:a + 16], c["grid"], fv_all[a:a + 16])
    return ss, time.perf_counter()
run("create_batch(64)", batch)
run("create_batch(64) desc only", batch_desc_only)
run("create_batch 4 x 16", batch4x16)
# raw DMA: 64 cudaMemcpyAsync of 2 MB pinned -> device on one torch stream
import torch
dev = [torch.empty((n, 128), dtype=torch.float32, device="cuda") for _ in kfs]
pin = [torch.from_numpy(h) for h in h_desc]
st = torch.cuda.Stream()
for rep in range(5):
    torch.cuda.synchronize(); t = time.perf_counter()
    with torch.cuda.stream(st):
        for d, p in zip(dev, pin):
            d.copy_(p, non_blocking=True)
    ti = time.perf_counter(); torch.cuda.synchronize(); tt = time.perf_counter() - t
print(f"{'raw torch copies (pinned? ' + str(pin[0].is_pinned()) + ')':44s} issue {1e3*(ti-t):6.3f} ms  done {1e3*tt:6.3f} ms  -> {131.072e6/tt/1e9:5.1f} GB/s")
