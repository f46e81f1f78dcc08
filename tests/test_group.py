"""The native multi-GPU group (include/asdmatch.h "multi-GPU groups", csrc/group.cu).

  * C++: tests/cpp/test_group.cc drives asdm_group_* from plain C++ -- no Python, torch or NCCL in the data path -- and
    compares C5 / C4 jobs over N ranks with the single-GPU asdm_job_* results bit for bit.
  * Python mirror, one process hosting every rank (Group.create), against the CPU oracle.
  * One process per rank (Group.join: cudaIpc windows, a memfd host block shared by the processes, gloo only for the
    handle exchange), against the CPU oracle.
On a box with one GPU the ranks share it: everything but NVLink itself is exercised.
"""
import importlib
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "test_group.bin")

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("ranks,n,pairs", [(2, 1200, 24), (3, 700, 10), (1, 500, 6)])
def test_cpp_group_equals_single_gpu(ctx, ranks, n, pairs):
    if not os.path.exists(BIN):
        subprocess.check_call(["make", "-C", os.path.dirname(BIN)])
    r = subprocess.run([BIN, str(ranks), str(n), str(pairs)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all comparisons identical" in r.stdout and "DIFFERENT" not in r.stdout


def _case(pkg, kind, n):
    if kind == "tri":
        c = pkg.synth.config4(n=n, n_pairs=7)
        return c, [k["has_mp"] for k in c["kfs"]], pkg.ASDM_JOB_TRIANGULATION, dict(
            nnratio=0.6, check_ori=0, F12=c["F12"], epipoles=c["epipoles"], scale_factors=c["scale_factors"], level_sigma2=c["level_sigma2"])
    c = pkg.synth.config5(n=n, n_pairs=11, n_kf=5, n_nodes=6)
    return c, [k["valid"] for k in c["kfs"]], pkg.ASDM_JOB_BOW_KF_KF, dict(nnratio=c["nnratio"], check_ori=c["check_ori"])


def _oracle(oracle, kind, c):
    if kind == "tri":
        return oracle.batch_triangulation(c["kfs"], c["ia"], c["ib"], c["F12"], c["epipoles"], c["scale_factors"], c["level_sigma2"], 0,
                                          threads=oracle.max_threads())
    return oracle.batch_bow_kf_kf(c["kfs"], c["ia"], c["ib"], c["nnratio"], c["check_ori"], threads=oracle.max_threads())


@pytest.mark.parametrize("kind,n,blocks", [("bow", 900, 1), ("bow", 701, 1), ("tri", 1000, 1), ("tri", 1000, 3)])
@pytest.mark.parametrize("world", [2, 3])
def test_group_in_one_process_matches_oracle(pkg, ctx, oracle, kind, n, blocks, world):
    G = importlib.import_module("asd-slam_b200.group")
    ng = _n_gpus()
    g = G.Group.create([r % ng for r in range(world)], window_bytes=256 << 20)
    try:
        c, masks, job_kind, kw = _case(pkg, kind, n)
        ref = _oracle(oracle, kind, c)
        descs = [np.ascontiguousarray(k["desc"], np.float32) for k in c["kfs"]]
        for step in range(2):  # the second batch recycles the window range of the first
            sets = g.sets_create(descs, [k["kps"] for k in c["kfs"]], c["grid"], [k["fv"] for k in c["kfs"]])
            job = G.GroupJob(g, job_kind, sets, masks, c["ia"], c["ib"], blocks_per_pair=blocks, gather_device=(step == 1),
                             gather_dist=(world == 2), **kw)
            nm, match, dist = job.fetch()
            assert np.array_equal(nm, ref[0]) and ref[0].sum() > 0
            assert np.array_equal(match[:, :n], ref[1])
            assert np.array_equal(dist[:, :n].view(np.uint32), ref[2].view(np.uint32))
            job.destroy()
            # the replicated sets are ordinary sets: a single-pair search on the LAST member sees the same keyframes
            a, b = int(c["ia"][0]), int(c["ib"][0])
            if kind == "bow":
                mt = pkg.ORBmatcher(g.ctxs[-1], c["nnratio"], True)
                n1, m1, _ = mt.SearchByBoW_KF_KF(sets[a][-1], masks[a], sets[b][-1], masks[b])
                assert n1 == ref[0][0] and np.array_equal(m1, ref[1][0])
            for row in sets:
                for s in row:
                    s.release()
    finally:
        g.close()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _join_worker(rank, world, port, kind, n, blocks, q):
    try:
        import torch
        import torch.distributed as dist
        sys.path.insert(0, ROOT)
        pkg = importlib.import_module("asd-slam_b200")
        G = importlib.import_module("asd-slam_b200.group")
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        dev = rank % torch.cuda.device_count()
        ctx = pkg.Context(dev)
        g = G.Group.join(ctx, rank, world, G.torch_allgather(), window_bytes=256 << 20)
        c, masks, job_kind, kw = _case(pkg, kind, n)
        descs = [np.ascontiguousarray(k["desc"], np.float32) for k in c["kfs"]]
        outs = []
        # a host block shared by the processes: every rank's kernels store their rows into it
        sets = g.sets_create(descs, [k["kps"] for k in c["kfs"]], c["grid"], [k["fv"] for k in c["kfs"]])
        job = G.GroupJob(g, job_kind, sets, masks, c["ia"], c["ib"], blocks_per_pair=blocks, gather_device=True, **kw)
        h_match = g.host_array((job.n_pairs, job.stride), np.int32)
        h_dist = g.host_array((job.n_pairs, job.stride), np.float32)
        h_nm = g.host_array((job.n_units,), np.int32)
        for step in range(3):
            job.run(h_match, h_dist, h_nm)
            job.wait()
            outs.append((job.counts(h_nm), h_match.copy(), h_dist.copy()))
            g.host_barrier()  # nobody re-runs into the block while a slower process still copies out of it
        outs.append(job.fetch())
        job.destroy()
        for row in sets:
            for s in row:
                s.release()
        for a in (h_match, h_dist, h_nm):
            g.host_free(a)
        g.close()
        ctx.close()
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, True, outs))
    except Exception:  # pragma: no cover
        import traceback
        q.put((rank, False, traceback.format_exc()))


@pytest.mark.parametrize("kind,n,blocks", [("bow", 800, 1), ("tri", 900, 2)])
def test_group_one_process_per_rank(pkg, oracle, kind, n, blocks):
    """SURVEY 8e as bench.py runs it: one process per rank. Windows mapped through cudaIpc, results stored by every rank's
    kernels into ONE host block shared by the processes (memfd), gloo only exchanges the handles."""
    import torch.multiprocessing as mp
    mctx = mp.get_context("spawn")
    q = mctx.Queue()
    port = _free_port()
    procs = [mctx.Process(target=_join_worker, args=(r, 2, port, kind, n, blocks, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), [r for r in res if not r[1]]
    c, _, _, _ = _case(pkg, kind, n)
    ref = _oracle(oracle, kind, c)
    assert ref[0].sum() > 0
    for _, _, outs in res:
        for nm, match, dist_ in outs:
            assert np.array_equal(nm, ref[0])
            assert np.array_equal(match[:, :n], ref[1])
            assert np.array_equal(dist_[:, :n].view(np.uint32), ref[2].view(np.uint32))
