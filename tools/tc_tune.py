#!/usr/bin/env python
"""Times the tensor-core shortlist kernel on the C5 sweep for each pipeline-shape variant (asdm_set_option
"tc_variant"); variants that drop the epilogue produce no results and are for timing the TMA + MMA pipeline only."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("asd-slam_b200")
ctx = pkg.Context(0)
c = bench.make_inputs(pkg, 0)
kfs = c["kfs"]
sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
job = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, [k["valid"] for k in kfs], c["ia"], c["ib"], c["nnratio"], c["check_ori"])
flops = 2.0 * len(c["ia"]) * 4000 * 4000 * 128
variants = [int(v) for v in sys.argv[1:] if not v.startswith("-")] or [0, 1, 2, 3, 4]
LONG = "--long" in sys.argv  # 1000 steps per variant with nvidia-smi clock / power sampling
for v in variants:
    ctx.set_option("tc_variant", v)
    for _ in range(3):
        job.run()
    ctx.synchronize()
    ctx.set_option("time_kernels", 1)
    ctx.stat("dist_stage_us")
    steps = 1000 if LONG else 30
    sampler = bench.ClockSampler(0) if LONG else None
    import time
    t0 = time.time()
    for _ in range(steps):
        job.run()
    ctx.synchronize()
    t1 = time.time()
    clk = sampler.stop(t0, t1) if sampler else ""
    us = ctx.stat("dist_stage_us") / steps
    ctx.set_option("time_kernels", 0)
    print(f"variant {v}: {us:8.1f} us  {flops / us / 1e6:7.0f} TFLOP/s  {clk}")
