#!/usr/bin/env python
"""C4 (local-mapping triangulation: 20 covisible keyframe pairs x 2000 x 2000 per new keyframe,
ORBmatcher::SearchForTriangulation, LocalMapping.cc:335-370) sharded over 1/2/4/8 GPUs, resident keyframe sets.

Two readings of the config, both printed by rank 0 as one JSON object:
  one_keyframe  STRONG scaling: the 20 pairs of ONE new keyframe, cut into >= 40 (pair, row-block) units
                (SearchForTriangulation has no cross-query state) and dealt round-robin; results of all units end
                up on every rank.  0.1 ms of work: this is a latency number.
  batch         WEAK scaling: `--keyframes` new keyframes' worth of pairs per GPU (default 64 x 20 = 1280 pairs per
                rank), whole pairs dealt round-robin, results published to every rank.

  python tools/scale_c4.py                                                        (1 GPU)
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/scale_c4.py
Timing: CUDA events on the job's stream around `--steps` steps after 5 warm-up steps, barrier + synchronize on both
sides, max over ranks.  Rank 0 checks the merged results of one_keyframe against the CPU oracle (bit-exact).
"""
import argparse
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--keyframes", type=int, default=64, help="batch mode: new keyframes (x 20 pairs) per GPU per step")
    ap.add_argument("--gather", default="push", choices=["push", "nccl"])
    ap.add_argument("--blocks", type=int, default=2, help="one_keyframe mode at N>1: row blocks per pair (2 -> 40 units)")
    ap.add_argument("--no-dist", action="store_true", help="publish index lists + counts only (the reference returns no distances)")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module("asd-slam_b200")
    ctx = pkg.Context(local)
    stream = torch.cuda.Stream()
    ctx.set_stream(stream.cuda_stream)
    c = pkg.synth.config4()
    kfs = c["kfs"]
    n = len(kfs[0]["kps"])
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    masks = [k["has_mp"] for k in kfs]

    def timed(job, steps):
        def step():
            job.run()
            if world > 1:
                job.gather()
        for _ in range(5):
            step()
            job.settle()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0.record(stream)
        for _ in range(steps):
            step()
        e1.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=f"cuda:{local}")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        job.settle()
        assert job.retries == 0
        return ms / steps

    def make_job(ia, ib, F12, ep, blocks):
        shards = pkg.PairShards(np.full(len(ia), n), world, blocks)
        gathered = pkg.GatheredResults(ctx, rank, world, shards.per_rank, n) if world > 1 and args.gather == "push" else None
        job = pkg.ShardedPairJob(ctx, pkg.ASDM_JOB_TRIANGULATION, sets, masks, ia, ib, 0.6, 0, F12, ep, c["scale_factors"],
                                 c["level_sigma2"], rank=rank, world=world, blocks_per_pair=blocks, gathered=gathered,
                                 publish_dist=not args.no_dist)
        return job, gathered

    out = {"config": "C4 SearchForTriangulation, 20 covisible pairs x 2000 x 2000 per new keyframe", "n_gpus": world,
           "steps": args.steps, "gather": args.gather if world > 1 else "none",
           "published": "match indices + counts" if args.no_dist else "match indices + distances + counts"}

    # ---- one new keyframe: 20 pairs as >= 40 row-block units over the ranks
    blocks = 1 if world == 1 else args.blocks
    job, g = make_job(c["ia"], c["ib"], c["F12"], c["epipoles"], blocks)
    ms = timed(job, args.steps)
    nm, match, d = job.fetch()
    one = {"scaling": "strong", "units": job.shards.n_units, "blocks_per_pair": blocks, "ms_per_step": ms,
           "keyframe_pairs_per_sec": 20.0 / (ms * 1e-3), "descriptor_pairs_per_sec": 20.0 * n * n / (ms * 1e-3),
           "launches_per_step": job.launches, "transport": g.transport if g is not None else None}
    if rank == 0 and not args.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_py as oracle  # checker only
        onm, om, od = oracle.batch_triangulation(kfs, c["ia"], c["ib"], c["F12"], c["epipoles"], c["scale_factors"],
                                                 c["level_sigma2"], 0, threads=oracle.max_threads())
        one["bit_exact_vs_oracle"] = bool(np.array_equal(nm, onm) and np.array_equal(match[:, :n], om[:, :n]) and
                                          (args.no_dist and world > 1 or
                                           np.array_equal(d[:, :n].view(np.uint32)[match[:, :n] >= 0],
                                                          od[:, :n].view(np.uint32)[om[:, :n] >= 0])))
    first_nm, first_match = nm.copy(), match.copy()
    job.destroy()
    if g is not None:
        g.close()
    out["one_keyframe"] = one

    # ---- a batch: `keyframes` new keyframes x 20 pairs per GPU, whole pairs round-robin (weak scaling)
    reps = args.keyframes * world
    ia, ib = np.tile(c["ia"], reps), np.tile(c["ib"], reps)
    F12, ep = np.tile(c["F12"], (reps, 1)), np.tile(c["epipoles"], (reps, 1))
    job, g = make_job(ia, ib, F12, ep, 1)
    ms = timed(job, max(20, args.steps // 4))
    nm, match, d = job.fetch()
    same = bool(np.array_equal(nm, np.tile(first_nm, reps)) and np.array_equal(match, np.tile(first_match, (reps, 1))))
    out["batch"] = {"scaling": "weak", "pairs_per_gpu": 20 * args.keyframes, "ms_per_step": ms,
                    "keyframe_pairs_per_sec": len(ia) / (ms * 1e-3), "descriptor_pairs_per_sec": len(ia) * float(n) * n / (ms * 1e-3),
                    "launches_per_step": job.launches, "transport": g.transport if g is not None else None,
                    "equals_one_keyframe_results": same}
    job.destroy()
    if g is not None:
        g.close()
    if rank == 0:
        print(json.dumps(out), flush=True)
    for s in sets:
        s.release()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
