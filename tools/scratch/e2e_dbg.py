import importlib, sys, os, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("asd-slam_b200")
ctx = pkg.Context(0)
c, g_ia, g_ib = bench.global_sweep(pkg, 1, 512)
kfs = c["kfs"]; n = 4000
masks = [k["valid"] for k in kfs]
kps_all, fv_all = [k["kps"] for k in kfs], [k["fv"] for k in kfs]
h_desc = [ctx.pinned((n, 128), np.float32) for _ in kfs]
for h, k in zip(h_desc, kfs): h[...] = k["desc"]
sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
lvl = int(sys.argv[1]) if len(sys.argv) > 1 else 0
ctx.set_option("resolve_level", lvl)
job = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, masks, g_ia, g_ib, c["nnratio"], c["check_ori"])
job.run(); nm, match, dist = job.fetch()
print("resident retries", ctx.stat("retries"))
ctx.set_option("resolve_level", 2)
job.run(); nm2, match2, dist2 = job.fetch()
print("global-kernel equal:", np.array_equal(nm, nm2), np.array_equal(match, match2), np.array_equal(dist.view(np.uint32), dist2.view(np.uint32)))
ctx.set_option("resolve_level", lvl)
S = job.stride
h_match, h_dist, h_nm = ctx.pinned((512, S), np.int32), ctx.pinned((512, S), np.float32), ctx.pinned((512,), np.int32)
sweep = pkg.StreamedPairSweep(ctx, pkg.ASDM_JOB_BOW_KF_KF, len(kfs), g_ia, g_ib, c["nnratio"], c["check_ori"])
for step in range(6):
    h_match[...] = -7; h_nm[...] = -7
    ss = sweep.run(lambda lo, hi: pkg.DescriptorSet.create_batch(ctx, h_desc[lo:hi], kps_all[lo:hi], c["grid"], fv_all[lo:hi]), masks, h_match, h_dist, h_nm)
    for s_ in ss: s_.release()
    bad = np.nonzero((h_match != match[sweep.order]).any(axis=1))[0]
    badn = np.nonzero(h_nm != nm[sweep.order])[0]
    print("step", step, "bad rows", len(bad), bad[:8], "bad nm", len(badn), "retries(last)", ctx.stat("retries"), "untouched rows", int((h_match[:, 0] == -7).sum()))
