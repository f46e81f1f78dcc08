"""ms per job run (back to back, no events between the kernels) for a sweep of P pairs on one GPU, no exchange."""
import importlib, sys, os, time, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("asd-slam_b200")
ctx = pkg.Context(0)
for P in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "512,64").split(",")]:
    c, g_ia, g_ib = bench.global_sweep(pkg, 1, P)
    kfs = c["kfs"]
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    masks = [k["valid"] for k in kfs]
    job = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, masks, g_ia, g_ib, c["nnratio"], c["check_ori"])
    for _ in range(5):
        job.run()
    job.settle()
    ref = None
    for rep in range(2):
        for pdl in (0,):
            for _ in range(5):
                job.run()
            ctx.synchronize()
            n = 100 if P <= 128 else 30
            t0 = time.perf_counter()
            for _ in range(n):
                job.run()
            ctx.synchronize()
            ms = 1e3 * (time.perf_counter() - t0) / n
            nm, match, dist = job.fetch()
            if ref is None:
                ref = (nm.copy(), match.copy(), dist.copy())
            same = np.array_equal(nm, ref[0]) and np.array_equal(match, ref[1]) and np.array_equal(dist.view(np.uint32), ref[2].view(np.uint32))
            print(f"pairs {P}: {ms:.4f} ms per run, same results {same}")
    job.destroy()
    for s in sets:
        s.release()
