#!/usr/bin/env python
"""bench.py -- headline benchmark of the ASD descriptor-matching hot path on B200.

Workload (BASELINE.json configs[4], "C5"): loop-closure verification sweep -- 512 candidate keyframe pairs, each
4000 x 4000 ASD descriptors (128-d fp32), matched with ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*) semantics
(ratio 0.85, rotation check; reference ORBmatcher.cc:533-666, LoopClosing.cc:291-318) over 64 distinct keyframes.
A "step" is one pass of the whole sweep.  Metric: descriptor pairs / s (= pairs * 4000 * 4000 * steps / time).

  python bench.py [--gpus N] [--steps K] [--warmup W]            our arm (one process per GPU; N>1 under torchrun)
  python bench.py --scaling strong ...                            the same 512 pairs dealt over the N ranks
  python bench.py --impl reference ...                            the reference's own CPU matcher (oracle/_ref: its
                                                                   ORBmatcher.cc compiled unmodified) on the host cores

Default (weak scaling): an N*512-pair sweep over the same 64 keyframes, dealt round-robin over the ranks by the library's
multi-GPU group (asdm_group_*, csrc/group.cu: no torch / NCCL on the data path -- torch.distributed only launches the
processes, exchanges the window handles once and reduces the timings).  Every weak-scaling line also carries
  strong : the fixed 512-pair sweep over the N ranks (ms, pairs/s, speed-up against one GPU measured in the same run)
  c1..c4 : the other numbers of BASELINE.json's metric -- ms per 2000x2000 frame pair (C1), the tracking searches (C2, C3)
           through the C ABI, keyframe pairs / s of the triangulation batch (C4) at this N -- each with the CPU oracle's time
           and a parity flag.
"""
import argparse
import hashlib
import importlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_DESC = 4000
N_KF = 64
N_PAIRS = 512
NNRATIO = 0.85
METRIC = "descriptor_pairs_per_sec"
UNIT = "descriptor pairs/s"
WORKLOAD = ("C5 loop-closure sweep: 512 keyframe pairs x 4000x4000 ASD descriptors (128-d fp32) per GPU, "
            "SearchByBoW(KF,KF) ratio 0.85 + rotation histogram, 64 distinct keyframes")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(bf16=float(d["bf16_tflops_sustained"]), bf16_burst=float(d["bf16_tflops"]), hbm=float(d["hbm_gbs"]),
                    source="measured")
    return dict(bf16=1400.0, bf16_burst=1590.0, hbm=6650.0, source="fallback")


def git_blob(path):
    with open(path, "rb") as f:
        data = f.read()
    return hashlib.sha1(b"blob %d\0" % len(data) + data).hexdigest()


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one k_tc_shortlist launch on this workload, from the committed
    `ncu --set full` summary under profiles/ whose `source_blob:` line names the kernel source this run was built from
    (tools/ncu_summary.py records it). No such summary -> (None, why): a capture of another kernel version is not quoted."""
    import glob
    import re
    src = os.path.join(ROOT, "asd-slam_b200", "csrc", "tc_shortlist.cu")
    blob = git_blob(src)
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_tc_shortlist*.txt")), reverse=True):
        with open(path) as f:
            text = f.read()
        if blob not in text.split("== kernel", 1)[0]:
            continue
        tot = 0.0
        for section in text.split("== kernel")[1:]:  # a summary may hold several kernels: only k_tc_shortlist's counters
            if "k_tc_shortlist" not in section.split("\n", 1)[0]:
                continue
            for m in re.finditer(r"dram__bytes_(read|write)\.sum\s+([0-9.]+)\s+(\w+)", section):
                tot += float(m.group(2)) * mult.get(m.group(3), 1.0)
            break
        return (tot or None), os.path.relpath(path, ROOT)
    return None, f"no ncu summary under profiles/ for tc_shortlist.cu blob {blob[:12]}"


def make_inputs(pkg, rank, n=N_DESC, n_kf=N_KF, n_pairs=N_PAIRS):
    c = pkg.synth.config5(n=n, n_pairs=n_pairs, n_kf=n_kf, seed_extra=0)
    if rank:
        rng = np.random.default_rng(0xC5 + rank)  # this rank's shard of the sweep: its own candidate pairs
        ia = rng.integers(0, n_kf, n_pairs).astype(np.int32)
        c["ia"], c["ib"] = ia, ((ia + 1 + rng.integers(0, n_kf - 1, n_pairs)) % n_kf).astype(np.int32)
    return c


def global_sweep(pkg, world, n_pairs):
    """The world*n_pairs candidate pairs of the whole job, interleaved so that the round-robin deal of the group hands
    rank r exactly make_inputs(rank=r)'s pairs (rank 0's shard is the N=1 workload)."""
    c = make_inputs(pkg, 0, n_pairs=n_pairs)
    ia = np.empty(world * n_pairs, np.int32)
    ib = np.empty(world * n_pairs, np.int32)
    ia[0::world], ib[0::world] = c["ia"], c["ib"]
    for r in range(1, world):
        cr = make_inputs(pkg, r, n_pairs=n_pairs)
        ia[r::world], ib[r::world] = cr["ia"], cr["ib"]
    return c, ia, ib


class ClockSampler:
    """SM clock / throttle reasons sampled every 5 ms during the timed region through NVML (B200_PROFILING.md: the
    clocks line); falls back to nvidia-smi when the binding is missing."""

    def __init__(self, index):
        self.rows, self.stop_flag, self.t, self.h = [], False, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.h = None

    def _pump(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                mx = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
                pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.rows.append((time.time(), sm, mx, pw, rs))
            except Exception:
                pass
            time.sleep(0.005)

    def stop(self, t0, t1):
        if self.h is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"], "samples": 0}
        self.stop_flag = True
        self.t.join(timeout=1.0)
        nv = self.nv
        inside = [r for r in self.rows if t0 <= r[0] <= t1]
        near = inside or sorted(self.rows, key=lambda r: min(abs(r[0] - t0), abs(r[0] - t1)))[:3]
        names = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        reasons = sorted({k for r in near for k, bit in names.items() if r[4] & bit})
        return {"sm_mhz": float(np.median([r[1] for r in near])) if near else None,
                "sm_max_mhz": float(max(r[2] for r in near)) if near else None,
                "power_w_max": float(max(r[3] for r in near)) if near else None, "samples": len(inside),
                "samples_near": len(near), "reasons": reasons}


# --------------------------------------------------------------------------------------------- CPU arms
def host_threads():
    """Host cores this process may use. torchrun exports OMP_NUM_THREADS=1 to every rank, which would throttle the
    CPU arm to one thread at N>1; only rank 0 runs CPU work, so it takes the whole affinity mask."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_sample(oracle, c, pairs, threads):
    """Times the CPU oracle (port of the reference matcher) on the first `pairs` pairs of the workload."""
    ia, ib = c["ia"][:pairs], c["ib"][:pairs]
    t = time.perf_counter()
    nm, match, dist = oracle.batch_bow_kf_kf(c["kfs"], ia, ib, c["nnratio"], c["check_ori"], threads=threads)
    dt = time.perf_counter() - t
    return dt, nm, match


def ref_sample(ref, c, pairs, threads, rows=None):
    """Times the reference's OWN ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*) (oracle/_ref, ORBmatcher.cc compiled
    unmodified) on the first `pairs` pairs of the workload, WHOLE pairs (all 4000 query features x 4000) unless `rows`
    restricts the query features (warm-up only), one pair per host thread at a time (ctypes drops the GIL). The reference
    matcher is serial; threads over pairs is the most it can use."""
    from concurrent.futures import ThreadPoolExecutor
    kfs, nn, ori = c["kfs"], c["nnratio"], c["check_ori"]

    def one(p):
        a, b = kfs[int(c["ia"][p])], kfs[int(c["ib"][p])]
        va = a["valid"]
        if rows is not None:
            va = np.array(va, dtype=np.uint8, copy=True)
            va[rows:] = 0
        return ref.search_by_bow_kf_kf(a["desc"], a["kps"], a["fv"], va, b["desc"], b["kps"], b["fv"], b["valid"], nn, ori)

    t = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        res = list(ex.map(one, range(pairs)))
    return time.perf_counter() - t, res


VIEWS_LIB = os.path.join(ROOT, "oracle", "_ref", "libref_orbmatcher_views.so")


def run_reference(args):
    """The reference arm: the reference's unmodified ORBmatcher.cc on all host threads. A step = one WHOLE 4000 x 4000
    keyframe pair per thread (about 15 s); the warm-up steps run a 64-row slice (a CPU arm has nothing to warm)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pkg = importlib.import_module("asd-slam_b200")
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as oracle  # the one place outside tests/smoke where bench.py executes oracle/: the reference arm
    import ref_py as ref
    oracle.build()
    threads = host_threads()
    c = make_inputs(pkg, 0)
    n = N_DESC
    other = None
    if ref.available():
        kind = "reference"
        pairs = min(threads, N_PAIRS)
        # headline: the cv::Mat stand-in in header mode (row() / reshape() are views, convertTo allocates once -- what
        # OpenCV 3.2 does); the copying stand-in of round 1 is timed once beside it
        views = os.path.exists(VIEWS_LIB)
        if views:
            ref.use(VIEWS_LIB)
        run = lambda: ref_sample(ref, c, pairs, threads)
        warm = lambda: ref_sample(ref, c, pairs, threads, rows=64)
        units = pairs * float(n) * n
        sample = (f"{pairs} WHOLE keyframe pairs of the {N_PAIRS} per step (4000 x 4000 each, one per host thread), the reference's "
                  f"unmodified ORBmatcher.cc (oracle/_ref, g++ -O2) over a from-scratch cv::Mat stand-in in "
                  f"{'header mode (views, one allocation per convertTo, as OpenCV)' if views else 'copy mode'}; "
                  f"warm-up steps run a 64-row slice")
    else:
        kind = "port"
        pairs = min(N_PAIRS, 4 * threads)
        run = lambda: cpu_sample(oracle, c, pairs, threads)
        warm = run
        units = pairs * float(n) * n
        sample = f"{pairs} of the {N_PAIRS} keyframe pairs per step (4000x4000 each), {threads} OpenMP threads over pairs"
    for _ in range(args.warmup):
        warm()
    # K steps as asked (20 whole-pair steps are about five minutes). Safety valve for a slow host: when the first step says
    # that K of them would take more than ten minutes, fewer steps are timed and the line says so (the value is a rate)
    steps_asked = args.steps
    t0 = time.perf_counter()
    done = 0
    while done < args.steps:
        run()
        done += 1
        if done == 1:
            t1 = time.perf_counter() - t0
            if t1 * args.steps > 600.0:
                args.steps = max(2, int(600.0 / t1))
    dt = time.perf_counter() - t0
    value = units * args.steps / dt
    if kind == "reference" and os.path.exists(VIEWS_LIB):
        ref.use(None)
        dtc, _ = ref_sample(ref, c, pairs, threads)
        other = {"value": units / dtc, "unit": UNIT, "what": "the same step with the copying cv::Mat stand-in of round 1 (row() copies, "
                 "convertTo allocates twice)", "share_of_stand_in_copies": 1.0 - (dt / args.steps) / dtc}
    # the conservative anchor: the OpenMP port of the same algorithm on the same threads
    dtp, _, _ = cpu_sample(oracle, c, min(N_PAIRS, 4 * threads), threads)
    port = {"value": min(N_PAIRS, 4 * threads) * float(n) * n / dtp, "unit": UNIT, "cores": threads, "kind": "port"}
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": args.scaling,
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": WORKLOAD, "sample": sample, "steps_asked": steps_asked},
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample,
                            "copy_mode_stand_in": other, "port_same_threads": port},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "keyframe_pairs_per_sec": value / (float(n) * n), "gpu_launches": 0}
    print(json.dumps(out), flush=True)


# ---------------------------------------------------------------------------------------------- our arm
def stage_times(ctx, job, reps=5):
    """Per-stage device times (ms, medians) of one run of `job` on this rank: events around each kernel (time_kernels=2)."""
    ctx.set_option("time_kernels", 2)
    acc = {k: [] for k in ("zero", "distance", "rerank", "resolve")}
    for _ in range(reps):
        for _ in range(3):  # back to back: the events of the LAST run are read, taken on a GPU that is already busy
            job.run()
        job.flush() if hasattr(job, "flush") else None
        ctx.synchronize()
        for k in acc:
            acc[k].append(ctx.stat("job_us_" + k))
    ctx.set_option("time_kernels", 0)
    return {k: float(np.median(v)) * 1e-3 for k, v in acc.items()}


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the matcher has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    pkg = importlib.import_module("asd-slam_b200")
    G = importlib.import_module("asd-slam_b200.group")
    ctx = pkg.Context(local)
    if args.epi_warps:
        ctx.set_option("tc_epi_warps", args.epi_warps)
    if args.tc_variant:
        ctx.set_option("tc_variant", args.tc_variant)
    # the multi-GPU group of the library: device windows, peer stores, device-side barriers. torch.distributed only
    # carries the handle exchange at set-up (and the reduction of the timings below)
    grp = G.Group.join(ctx, rank, world, G.torch_allgather() if world > 1 else None, window_bytes=args.window_mb << 20)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    strong = args.scaling == "strong"
    n, P = N_DESC, args.pairs

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def time_steps(step, steps, warm, flush=None):
        """ms per step of `step()` (queues work on the context stream), device-timed, max over ranks."""
        for _ in range(warm):
            step()
        ctx.synchronize()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            step()
        if flush is not None:
            flush()  # the exchange of the last step closes inside the timed region
        e1.record(stream)
        ctx.synchronize()
        return max_over_ranks(e0.elapsed_time(e1)) / steps

    # ---- inputs: the whole job on every rank (the library deals the units); weak: world*P pairs, strong: P pairs
    c, g_ia, g_ib = global_sweep(pkg, world, P)
    ia1, ib1 = c["ia"], c["ib"]  # the N=1 workload (rank 0's shard of the weak sweep)
    if strong:
        g_ia, g_ib = ia1, ib1
    kfs = c["kfs"]
    masks = [k["valid"] for k in kfs]
    kps_all, fv_all = [k["kps"] for k in kfs], [k["fv"] for k in kfs]
    h_desc = [ctx.pinned((n, 128), np.float32) for _ in kfs]  # the SLAM process holds its keyframes in page-locked memory
    for h, k in zip(h_desc, kfs):
        h[...] = k["desc"]
    prepared = grp.prepare_sets(h_desc, kps_all, c["grid"], fv_all)

    # ---- resident arm: the keyframes are uploaded (once over PCIe, replicated over NVLink by the pack kernel) before
    # the timed region, as the SLAM system does at keyframe creation; a step = the sweep + the exchange of the results
    sets = grp.sets_create(prepared=prepared)
    job = G.GroupJob(grp, pkg.ASDM_JOB_BOW_KF_KF, sets, masks, g_ia, g_ib, c["nnratio"], c["check_ori"], gather_device=world > 1,
                     gather_dist=args.gather_dist)
    S = job.stride
    for _ in range(max(args.warmup, 3)):
        job.run()
        job.wait()  # grows candidate pools if the first runs overflowed
    tc_used = ctx.stat("tc_used")
    ctx.set_option("time_kernels", 1)
    ctx.stat("dist_stage_us")  # drain
    barrier()
    ctx.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = ctx.stat("total_launches")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record(stream)
    for _ in range(args.steps):
        job.run()
    job.flush()  # the exchange of the last step closes inside the timed region
    e1.record(stream)
    ctx.synchronize()
    barrier()
    t1 = time.time()
    clocks = sampler.stop(t0, t1) if sampler else None
    launches = ctx.stat("total_launches") - l0
    ms = max_over_ranks(e0.elapsed_time(e1))
    dist_us = ctx.stat("dist_stage_us")
    ctx.set_option("time_kernels", 0)
    job.wait()
    assert ctx.stat("retries") == 0, "a candidate pool overflowed inside the timed region"
    shortlist_hits = ctx.stat("tc_pairs")
    nm, match, _ = job.fetch()  # the whole batch on every process, in global pair order
    n_job_pairs = len(g_ia)

    # ---- strong scaling of the FIXED 512-pair sweep (LoopClosing.cc:291-318 is one batch per keyframe): the same pairs
    # over the N ranks, against one GPU measured in this very run (rank 0 alone, through the single-GPU asdm_job_*)
    strong_rec = None
    if not strong:
        k_steps = max(10, args.steps)
        if world > 1:
            job_s = G.GroupJob(grp, pkg.ASDM_JOB_BOW_KF_KF, sets, masks, ia1, ib1, c["nnratio"], c["check_ori"], gather_device=True,
                               gather_dist=args.gather_dist)
            job_s.run(); job_s.wait()
            ms_n = time_steps(job_s.run, k_steps, 3, job_s.flush)
            job_s.wait()
            stage_ms = stage_times(ctx, job_s)
            nm_s, match_s, _ = job_s.fetch()
            job_s.destroy()
            ms_1 = None
            if rank == 0:
                one = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, [row[0] for row in sets], masks, ia1, ib1, c["nnratio"], c["check_ori"])
                one.run(); one.settle()
                for _ in range(3):
                    one.run()
                ctx.synchronize()
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record(stream)
                for _ in range(k_steps):
                    one.run()
                a1.record(stream)
                ctx.synchronize()
                ms_1 = a0.elapsed_time(a1) / k_steps
                nm_1, match_1, _ = one.fetch()
                one.destroy()
                strong_same = bool(np.array_equal(nm_s, nm_1) and np.array_equal(match_s, match_1))
            barrier()
        else:
            ms_n = ms_1 = ms / args.steps
            strong_same = True
            stage_ms = stage_times(ctx, job)
        if rank == 0:
            strong_rec = {"workload": f"the fixed {P}-pair sweep dealt over {world} ranks (strong scaling), results gathered on every GPU",
                          "ms_per_step": ms_n, "one_gpu_ms_per_step": ms_1, "speedup_vs_one_gpu": ms_1 / ms_n,
                          "descriptor_pairs_per_sec": P * float(n) * n / (ms_n * 1e-3), "same_results_as_one_gpu": strong_same,
                          "floor_terms_ms": dict(stage_ms, other=max(0.0, ms_n - sum(stage_ms.values())),
                                                 note="rank 0's share of the sweep, stage by stage (CUDA events around each kernel, "
                                                      "separate runs): zero = clearing the counters, distance = k_tc_shortlist, rerank = "
                                                      "exact distances of the short list, resolve = greedy replay + publishing the rows "
                                                      "to every GPU; other = input pull, row bounds, launch gaps and the exchange barrier (0 when the "
                                                      "event-bracketed stages, which cannot overlap, add up to more than the step). "
                                                      "Only `distance` shrinks with N; the rest is the floor of the step")}

    # ---- e2e arm: the reference-facing call with HOST buffers, every step: keyframes from page-locked host memory,
    # FeatureVectors, job description, search, match lists back in host memory
    e2e_steps = max(3, min(args.steps, 10))
    h2d_side = sum(k["kps"].nbytes + 8 * len(k["kps"]) for k in kfs)
    if world == 1:
        # the keyframes stream in over PCIe and the pairs are searched in readiness order (StreamedPairSweep), so the
        # upload of the later keyframes overlaps the search of the earlier pairs; row r of the outputs = pair order[r]
        h_match, h_dist, h_nm = ctx.pinned((P, S), np.int32), ctx.pinned((P, S), np.float32), ctx.pinned((P,), np.int32)
        sweep = pkg.StreamedPairSweep(ctx, pkg.ASDM_JOB_BOW_KF_KF, len(kfs), g_ia, g_ib, c["nnratio"], c["check_ori"])

        def e2e_step():
            ss = sweep.run(lambda lo, hi: pkg.DescriptorSet.create_batch(ctx, h_desc[lo:hi], kps_all[lo:hi], c["grid"], fv_all[lo:hi]),
                           masks, h_match, h_dist, h_nm)
            for s_ in ss:
                s_.release()
        what = (f"per step: 64 keyframes uploaded from page-locked host memory (asdm_set_create_batch per wave), pairs searched in "
                f"{len(sweep.bounds)} readiness waves (keyframe bounds {sweep.bounds}) that overlap the uploads, all match lists "
                "and distances read back")
        h2d = sum(k["desc"].nbytes for k in kfs) + h2d_side + sum(len(k["valid"]) for k in kfs) + 8 * P
        h2d += sum(sum(len(k["valid"]) for k in kfs[:b]) for b in sweep.bounds[:-1])  # every wave re-sends the masks it may touch
        d2h = P * S * 8 + P * 4
        own_h2d = h2d
    else:
        # every keyframe crosses PCIe ONCE (rank k % N uploads keyframe k); the owner's pack kernel replicates rows, bf16
        # copy, norms, keypoints and FeatureVector arrays into every GPU over NVLink; the units are dealt round-robin; each
        # rank's resolve kernel stores its finished rows straight into ONE host block shared by the processes (its own
        # PCIe link); a host barrier closes the step. Rank 0 -- the SLAM process -- then holds the whole batch.
        try:
            h_match = grp.host_array((n_job_pairs, S), np.int32)
            h_nm = grp.host_array((job.n_units,), np.int32)
            shared_host = True
        except pkg.AsdmError:  # no shared mapping on this box: every process keeps the rows of its own units
            h_match, h_nm, shared_host = ctx.pinned((n_job_pairs, S), np.int32), ctx.pinned((job.n_units,), np.int32), False

        waves = [int(x) for x in args.waves.split(",")] if args.waves else None
        if waves is None and world >= 4:  # a rank's upload share is short next to its searches: three waves measured best at 8
            nk = len(kfs)
            waves = sorted({max(2, int(round(nk * f))) for f in (0.375, 0.69, 1.0)})
        gsweep = G.StreamedGroupSweep(grp, pkg.ASDM_JOB_BOW_KF_KF, h_desc, kps_all, c["grid"], fv_all, g_ia, g_ib, c["nnratio"],
                                      c["check_ori"], bounds=waves)

        e2e_timers = {}

        def e2e_step():
            ss = gsweep.run(masks, h_match, h_nm, e2e_timers)
            tr = time.perf_counter()
            G.Group.release_sets(ss)
            e2e_timers["release_sets"] = e2e_timers.get("release_sets", 0.0) + time.perf_counter() - tr
        what = (f"per step, in {len(gsweep.bounds)} readiness waves (keyframe bounds {gsweep.bounds}) whose searches overlap the later "
                "uploads: each keyframe uploaded once (rank k%N, page-locked host memory) and replicated to every GPU by the "
                "pack kernel (peer stores over NVLink), pairs dealt round-robin by the library, every rank's resolve kernel stores "
                "its rows into " + ("one host block shared by the processes (memfd)" if shared_host else "its own page-locked block") +
                " over its own PCIe link (indices + counts: what SearchByBoW returns), host barrier")
        own_kf = len(range(rank, len(kfs), world))
        own_h2d = own_kf * (n * 512 + n * 24) + sum(len(k["valid"]) for k in kfs) + 8 * (n_job_pairs // world)
        h2d = sum(k["desc"].nbytes for k in kfs) + h2d_side + world * (sum(len(k["valid"]) for k in kfs) + 8 * (n_job_pairs // world))
        resent = sum(sum(len(k["valid"]) for k in kfs[:bd]) for bd in gsweep.bounds[:-1])  # every wave re-sends the masks it may touch
        h2d += world * resent
        own_h2d += resent
        d2h = n_job_pairs * S * 4 + job.n_units * 4
    e2e_step()
    e2e_step()
    if world > 1:
        e2e_timers.clear()
    barrier()
    ctx.synchronize()
    te = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    ctx.synchronize()
    e2e_ms = max_over_ranks(1e3 * (time.perf_counter() - te) / e2e_steps)
    if world == 1:
        e2e_same = bool(np.array_equal(h_match, match[sweep.order]) and np.array_equal(h_nm, nm[sweep.order]))
    elif shared_host:  # the shared block holds the whole batch in every process, rows in readiness order
        e2e_same = bool(np.array_equal(h_match, match[gsweep.order]) and np.array_equal(h_nm, nm[gsweep.order]))
    else:  # private blocks: the rows of this rank's units (unit u of a wave runs on rank u % world)
        ok = True
        for a, b_ in gsweep.slices:
            rows = np.arange(a + rank, b_, world)
            ok &= bool(np.array_equal(h_match[rows], match[gsweep.order[rows]]) and np.array_equal(h_nm[rows], nm[gsweep.order[rows]]))
        e2e_same = ok
    assert e2e_same, "e2e results differ from the resident-input results"
    # The floor of a cold sweep, measured on this box through the same calls: the step's keyframe uploads ALONE (every
    # rank its own keyframes, the pack kernel behind them, at N > 1 the replication over NVLink), no search, no read-back.
    # (A 64 MB torch copy as PCIe probe read 25 GB/s on a box whose sweep demonstrably moved 31 GB/s: other host memory.)
    def upload_only():
        if world == 1:
            ss = pkg.DescriptorSet.create_batch(ctx, h_desc, kps_all, c["grid"], fv_all)
        else:
            ss = grp.sets_create(prepared=prepared)
        ctx.synchronize()
        if world == 1:
            for s_ in ss:
                s_.release()
        else:
            G.Group.release_sets(ss)
    upload_only()
    up_ms = []
    for _ in range(3):
        barrier()
        tu = time.perf_counter()
        upload_only()
        up_ms.append(1e3 * (time.perf_counter() - tu))
    e2e_floor_ms = max_over_ranks(min(up_ms))
    pcie_gbs = own_h2d / (min(up_ms) * 1e-3) / 1e9

    # ---- the other numbers of BASELINE.json's metric. C4 at this N: one new keyframe (20 covisible pairs x 2000 x 2000,
    # LocalMapping.cc:335-370) and a batch of 64 of them (1280 pairs) dealt over the ranks, 128-row-aligned row blocks
    c4 = pkg.synth.config4()
    k4 = c4["kfs"]
    sets4 = grp.sets_create([np.ascontiguousarray(k["desc"], np.float32) for k in k4], [k["kps"] for k in k4], c4["grid"],
                            [k["fv"] for k in k4])
    m4 = [k["has_mp"] for k in k4]
    tri = dict(nnratio=0.6, check_ori=0, F12=c4["F12"], epipoles=c4["epipoles"], scale_factors=c4["scale_factors"], level_sigma2=c4["level_sigma2"])
    blocks = 1 if world == 1 else 2
    j1 = G.GroupJob(grp, pkg.ASDM_JOB_TRIANGULATION, sets4, m4, c4["ia"], c4["ib"], blocks_per_pair=blocks, gather_device=world > 1, **tri)
    j1.run(); j1.wait()
    ms_c4_one = time_steps(j1.run, 100, 5, j1.flush)
    j1.wait()
    nm4, match4, dist4 = j1.fetch()
    j1.destroy()
    reps = 64
    triB = dict(tri, F12=np.tile(c4["F12"], (reps, 1)), epipoles=np.tile(c4["epipoles"], (reps, 1)))
    jB = G.GroupJob(grp, pkg.ASDM_JOB_TRIANGULATION, sets4, m4, np.tile(c4["ia"], reps), np.tile(c4["ib"], reps), blocks_per_pair=1,
                    gather_device=world > 1, **triB)
    jB.run(); jB.wait()
    ms_c4_batch = time_steps(jB.run, 20, 3, jB.flush)
    jB.wait()
    jB.destroy()
    G.Group.release_sets(sets4)
    barrier()

    # ---- last GPU measurement of all ranks (it leaves the GPUs power-capped): the same step over a LONG region (400 steps, ~0.6 s): the K-step figure above is a burst at boost clocks; this one
    # is what a sweep that keeps coming sustains (power cap, clocks sampled over the region)
    sus_steps = 400
    sampler2 = ClockSampler(local) if rank == 0 else None
    ts0 = time.time()
    sus_ms = time_steps(job.run, sus_steps, 0, job.flush)
    barrier()
    ts1 = time.time()
    sus_clocks = sampler2.stop(ts0, ts1) if sampler2 else None
    job.wait()
    sustained_rec = {"steps": sus_steps, "ms_per_step": sus_ms, "clocks": sus_clocks,
                     "descriptor_pairs_per_sec": (P if strong else P * world) * float(n) * n / (sus_ms * 1e-3),
                     "what": "the timed step repeated back to back over a long region; `value` above is the K-step burst"}

    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import oracle_py as oracle  # cpu_baseline leg only: the checker, never the thing measured
        import latency
        oracle.build()
        pk = peaks()
        threads = host_threads()
        units = float(P if not strong else P / world) * n * n  # descriptor pairs per step per GPU
        total_units = float(n_job_pairs) * n * n
        value = total_units * args.steps / (ms * 1e-3)
        flops = 2.0 * units * 128  # 2*N*M*128 per keyframe pair, counted once
        dist_ms = dist_us * 1e-3 / max(1, args.steps)
        timed_s = ms * 1e-3
        if tc_used:
            achieved = flops / (dist_ms * 1e-3) / 1e12
            long_run = timed_s > 2.0
            peak = pk["bf16"] if long_run else pk["bf16_burst"]
            roof = {"bound": "tensor", "kernel": "k_tc_shortlist", "achieved": achieved, "peak": peak,
                    "unit": "TFLOP/s", "frac": achieved / peak, "traffic": None,
                    "peak_source": f"{pk['source']} dense bf16 cuBLAS, " + ("sustained (timed region > 2 s)" if long_run else
                                   f"BURST figure: the timed region is {timed_s * 1e3:.0f} ms at boost clocks") +
                                   " (the denominator is what cuBLAS reached on this pool, not a hardware limit: frac may touch 1)",
                    "frac_of_sustained_peak": achieved / pk["bf16"],
                    "frac_of_nominal_2250": achieved / 2250.0}
        else:
            achieved = 3.0 * units * 128 / (dist_ms * 1e-3) / 1e12  # sub, mul, add per element on the fp32 pipes
            roof = {"bound": "fp32", "kernel": "k_cand_exact", "achieved": achieved, "peak": None, "unit": "TFLOP/s",
                    "frac": None, "traffic": None, "note": "exact fp32 CUDA-core path (tensor-core shortlist disabled)"}
        if tc_used and P == N_PAIRS and not strong:
            roof["traffic"], roof["traffic_source"] = ncu_traffic()
            roof["traffic_unit"] = "bytes of DRAM read+write per launch (algorithmic: 64 keyframes x 1.024 MB bf16 read once)"
        roof["ms_per_launch"] = dist_ms
        roof["share_of_step"] = dist_ms / (ms / args.steps)

        # CPU baseline: the oracle port of the reference matcher on this box's host cores, bounded sample
        c0 = dict(c, ia=g_ia, ib=g_ib)  # the first pairs of the GLOBAL sweep: at N>1 they belong to different ranks
        cpu_pairs = int(min(n_job_pairs, max(threads, 8)))
        dtc, onm, omatch = cpu_sample(oracle, c0, cpu_pairs, threads)
        if dtc < 5.0:
            cpu_pairs = int(min(n_job_pairs, cpu_pairs * max(2, int(10.0 / max(dtc, 1e-3)))))
            dtc, onm, omatch = cpu_sample(oracle, c0, cpu_pairs, threads)
        parity = bool(np.array_equal(omatch, match[:cpu_pairs, :n]) and np.array_equal(onm, nm[:cpu_pairs]))
        cpu_val = cpu_pairs * float(n) * n / dtc
        # the reference's unmodified ORBmatcher.cc (oracle/_ref) on WHOLE pairs, one per thread, compared with the GPU's rows
        ref_note = None
        import ref_py as ref  # cpu_baseline leg only
        if n == N_DESC and ref.available() and not args.no_ref:
            if os.path.exists(VIEWS_LIB):
                ref.use(VIEWS_LIB)
            rp = min(threads, n_job_pairs)
            dtr, rres = ref_sample(ref, c0, rp, threads)
            rpar = all(rres[p][0] == nm[p] and bool(np.array_equal(rres[p][1], match[p, :n])) for p in range(rp))
            ref_note = {"value": rp * float(n) * n / dtr, "unit": UNIT, "cores": threads, "kind": "reference",
                        "sample": f"the first {rp} keyframe pairs, WHOLE (4000 x 4000), one per host thread, unmodified ORBmatcher.cc "
                                  "(oracle/_ref, cv::Mat stand-in in header mode)", "parity_vs_gpu_rows": bool(rpar)}
        # C1 - C3 through the C ABI on this GPU (single-pair tracking: replicas only), C4 from above
        lat = latency.measure(pkg, ctx, reps=100, oracle=oracle, only={"C1", "C2", "C3"})
        c2r = lat["C2_SearchByProjection_2000kp_vs_8000mp"]
        c3r = lat["C3_euroc_1500"]
        for rec, key in ((c2r, "abi_call_resident_pool_ms"), (c3r["stereo_row_match_1500"], "abi_call_ms"),
                         (c3r["frame_to_frame_projection_1500_th15"], "abi_call_ms")):
            rec["achieved_gbs"] = rec["algorithmic_bytes"] / (rec[key] * 1e-3) / 1e9
            rec["frac_of_hbm_peak"] = rec["achieved_gbs"] / pk["hbm"]
        o4 = oracle.batch_triangulation(k4, c4["ia"], c4["ib"], c4["F12"], c4["epipoles"], c4["scale_factors"], c4["level_sigma2"], 0,
                                        threads=threads)
        n4 = len(k4[0]["kps"])
        c4_rec = {"one_keyframe_20_pairs": {"ms": ms_c4_one, "keyframe_pairs_per_sec": 20.0 / (ms_c4_one * 1e-3), "units": 20 * blocks,
                                            "parity": bool(np.array_equal(nm4, o4[0]) and np.array_equal(match4[:, :n4], o4[1]) and
                                                           np.array_equal(dist4[:, :n4].view(np.uint32), o4[2].view(np.uint32)))},
                  "batch_1280_pairs": {"ms": ms_c4_batch, "keyframe_pairs_per_sec": 1280.0 / (ms_c4_batch * 1e-3)},
                  "n_gpus": world, "what": "SearchForTriangulation jobs dealt over the ranks by asdm_group_job_*, records gathered on every GPU; "
                                           "device-timed, inputs resident"}
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
               "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
               "dtype": "f32", "data": "synthetic",
               "config": {"workload": WORKLOAD if not strong else WORKLOAD.replace("per GPU", "in total, dealt over the GPUs"),
                          "pairs_per_gpu": P if not strong else P / world, "pairs_total": n_job_pairs, "descriptors_per_keyframe": n,
                          "l2": "inputs larger than L2 (197 MB of descriptors + candidate pools vs 126 MB); no flush",
                          "timed_region_ms": ms, "tensor_core_shortlist": bool(tc_used), "shortlist_hits_per_step": int(shortlist_hits),
                          "sharding": ("asdm_group_* (csrc/group.cu): units dealt round-robin; every rank's resolve kernel stores its finished "
                                       + ("match / distance / count" if args.gather_dist else "match / count (what SearchByBoW returns; distances stay "
                                          "with the rank that computed them)") + " records into every GPU's window (peer stores over NVLink), "
                                       "device-side flag barrier on its own stream -- no NCCL on the data path") if world > 1 else "single GPU"},
               "keyframe_pairs_per_sec": n_job_pairs * args.steps / (ms * 1e-3),
               "e2e": {"value": total_units / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                       "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms, "floor_ms": e2e_floor_ms, "pcie_h2d_gbs": pcie_gbs,
                       "host_ms_per_step_by_phase": ({k: 1e3 * v / e2e_steps for k, v in e2e_timers.items()} if world > 1 else None),
                       "floor_what": "the step's keyframe uploads ALONE, timed in this run through the same calls (best of 3, busiest rank; "
                                     "at N > 1 incl. the NVLink replication): the searches and the read-back can hide behind "
                                     "it, nothing else can shrink it. pcie_h2d_gbs = this rank's H2D bytes / that time",
                       "what": what},
               "gpu_launches": int(launches),
               "roofline": roof,
               "cpu_baseline": {"value": cpu_val, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"first {cpu_pairs} of the {n_job_pairs} keyframe pairs, {threads} OpenMP threads over pairs",
                                "parity_on_sample": parity, "reference_unmodified": ref_note},
               "strong": strong_rec, "sustained": sustained_rec,
               "c1": lat["C1_frame_pair_2000x2000_SearchByBoW"], "c2": c2r, "c3": c3r, "c4": c4_rec,
               "clocks": clocks}
        print(json.dumps(out), flush=True)
    barrier()
    job.destroy()
    G.Group.release_sets(sets)
    grp.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: 512 pairs per GPU (default); strong: the same 512 pairs dealt over the GPUs")
    ap.add_argument("--pairs", type=int, default=N_PAIRS, help="keyframe pairs per step (per GPU under weak scaling)")
    ap.add_argument("--window-mb", type=int, default=3072, help="device window of the multi-GPU group per rank")
    ap.add_argument("--waves", default="", help="e2e at N>1: keyframe bounds of the readiness waves, e.g. 16,32,48,64 (default: 6 waves at N=2, 3 at N>=4)")
    ap.add_argument("--gather-dist", action="store_true", help="N>1: the distance rows cross NVLink too (default: index lists + counts)")
    ap.add_argument("--no-ref", action="store_true", help="skip the whole-pair run of the reference's own matcher (about 15 s)")
    ap.add_argument("--tc-variant", type=int, default=0, help="tuning: pipeline shape of the tensor-core kernel")
    ap.add_argument("--epi-warps", type=int, default=0, help="tuning: epilogue warps of the tensor-core kernel (4 or 8)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
