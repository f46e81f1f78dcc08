import importlib, sys, os, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("asd-slam_b200")
ctx = pkg.Context(0)
P = int(sys.argv[1]) if len(sys.argv) > 1 else 512
c, g_ia, g_ib = bench.global_sweep(pkg, 1, P)
kfs = c["kfs"]
sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
masks = [k["valid"] for k in kfs]
for level in (0, 1, 2):
    ctx.set_option("resolve_level", level)
    job = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, masks, g_ia, g_ib, c["nnratio"], c["check_ori"])
    for _ in range(5):
        job.run()
    job.settle()
    ctx.set_option("time_kernels", 2)
    acc = {k: [] for k in ("zero", "distance", "rerank", "resolve")}
    for _ in range(10):
        job.run(); ctx.synchronize()
        for k in acc: acc[k].append(ctx.stat("job_us_" + k))
    ctx.set_option("time_kernels", 0)
    print("pairs", P, "resolve level", level, {k: float(np.median(v)) for k, v in acc.items()})
    job.destroy()
