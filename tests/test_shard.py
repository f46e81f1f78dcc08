"""Sharding of batched keyframe-pair searches over ranks (asd-slam_b200/shard.py).

CPU: the partition / merge host logic, and a world_size-2 gloo run in which every rank computes its own units (with the
oracle standing in for the GPU, as the checker) and the all-gathered, merged batch equals the un-sharded result.
GPU: row-block units through the C ABI (asdm_job_desc.row_begin / row_end) and ShardedPairJob at world size 1.
"""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _shard_mod():
    importlib.import_module("asd-slam_b200")
    return importlib.import_module("asd-slam_b200.shard")


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("blocks", [1, 2, 5])
def test_partition_covers_every_row_once(world, blocks):
    sh = _shard_mod()
    n_rows = [2000, 1999, 128, 1, 0, 4000, 257]
    ps = sh.PairShards(n_rows, world, blocks)
    seen = [np.zeros(max(n, 1), np.int32) for n in n_rows]
    owned = np.zeros(ps.n_units, np.int32)
    for r in range(world):
        mine = ps.local_units(r)
        assert len(mine) <= ps.per_rank
        owned[mine] += 1
        for u in mine:
            p, a, b = ps.unit_pair[u], ps.unit_r0[u], ps.unit_r1[u]
            assert 0 <= a <= b <= n_rows[p]
            assert a % 128 == 0 and (b % 128 == 0 or b == n_rows[p])  # tile-aligned cuts
            seen[p][a:b] += 1
    assert (owned == 1).all()
    for p, n in enumerate(n_rows):
        assert (seen[p][:n] == 1).all()
    # loads differ by at most one unit
    loads = [len(ps.local_units(r)) for r in range(world)]
    assert max(loads) - min(loads) <= 1
    if blocks == 2 and world == 8:  # C4: 20 pairs -> 40 units -> 5 per GPU
        c4 = sh.PairShards([2000] * 20, 8, 2)
        assert c4.n_units == 40 and all(len(c4.local_units(r)) == 5 for r in range(8))


def _fake_unit_outputs(ps, rank, full_match, full_dist, stride):
    """What a rank's job returns for its units, derived from the un-sharded result (rows outside the unit are -1)."""
    mine = ps.local_units(rank)
    m = np.full((len(mine), stride), -1, np.int32)
    d = np.full((len(mine), stride), -1.0, np.float32)
    nm = np.zeros(len(mine), np.int32)
    for s, u in enumerate(mine):
        p, a, b = ps.unit_pair[u], ps.unit_r0[u], ps.unit_r1[u]
        m[s, a:b] = full_match[p, a:b]
        d[s, a:b] = full_dist[p, a:b]
        nm[s] = (full_match[p, a:b] >= 0).sum()
    return m, d, nm


@pytest.mark.parametrize("world,blocks", [(1, 1), (2, 1), (3, 2), (8, 2), (4, 3)])
def test_merge_inverts_the_partition(world, blocks):
    sh = _shard_mod()
    rng = np.random.default_rng(7)
    P, stride = 11, 700
    full_match = rng.integers(-1, stride, (P, stride)).astype(np.int32)
    full_dist = np.where(full_match >= 0, rng.random((P, stride)), -1.0).astype(np.float32)
    ps = sh.PairShards([stride] * P, world, blocks)
    gm = np.full((world, ps.per_rank, stride), -1, np.int32)
    gd = np.full((world, ps.per_rank, stride), -1.0, np.float32)
    gn = np.full((world, ps.per_rank), -1, np.int32)  # ragged tails stay -1 and must be ignored
    for r in range(world):
        m, d, n = _fake_unit_outputs(ps, r, full_match, full_dist, stride)
        gm[r, :len(m)], gd[r, :len(m)], gn[r, :len(m)] = m, d, n
    nm, match, dist = ps.merge(gm, gd, gn)
    assert np.array_equal(match, full_match) and np.array_equal(dist, full_dist)
    assert np.array_equal(nm, (full_match >= 0).sum(axis=1))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _gloo_worker(rank, world, port, blocks, q):
    try:
        import torch
        import torch.distributed as dist
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        pkg = importlib.import_module("asd-slam_b200")
        sh = importlib.import_module("asd-slam_b200.shard")
        import oracle_py as oracle  # the checker stands in for the GPU here
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        n, pairs = 300, 7
        c = pkg.synth.config4(n=n, n_pairs=pairs, n_nodes=3)
        ps = sh.PairShards([n] * pairs, world, blocks)
        mine = ps.local_units(rank)
        up = ps.unit_pair[mine]
        # this rank's own units only
        onm, om, od = oracle.batch_triangulation(c["kfs"], c["ia"][up], c["ib"][up], c["F12"][up], c["epipoles"][up],
                                                 c["scale_factors"], c["level_sigma2"], 0)
        for s, u in enumerate(mine):  # a row-block unit reports only its own query rows
            a, b = ps.unit_r0[u], ps.unit_r1[u]
            om[s, :a], om[s, b:] = -1, -1
            od[s, :a], od[s, b:] = -1.0, -1.0
            onm[s] = (om[s] >= 0).sum()
        g_m = sh.all_gather_results(torch.from_numpy(om), ps.per_rank)
        g_d = sh.all_gather_results(torch.from_numpy(od), ps.per_rank)
        g_n = sh.all_gather_results(torch.from_numpy(onm), ps.per_rank)
        nm, match, dist_ = ps.merge(g_m.numpy(), g_d.numpy(), g_n.numpy())
        fnm, fm, fd = oracle.batch_triangulation(c["kfs"], c["ia"], c["ib"], c["F12"], c["epipoles"], c["scale_factors"],
                                                 c["level_sigma2"], 0)
        ok = bool(np.array_equal(match, fm) and np.array_equal(dist_.view(np.uint32), fd.view(np.uint32)) and
                  np.array_equal(nm, fnm) and fnm.sum() > 0)
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, ok, int(fnm.sum())))
    except Exception as e:  # pragma: no cover
        q.put((rank, False, repr(e)))


@pytest.mark.parametrize("blocks", [1, 2])
def test_world2_gloo_allgather_matches_unsharded(blocks, oracle):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, blocks, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res


# ------------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("force_exact", [0, 1])
@pytest.mark.parametrize("world,blocks", [(2, 2), (3, 3)])
def test_row_block_units_match_oracle(pkg, ctx, oracle, world, blocks, force_exact):
    sh = _shard_mod()
    n, pairs = 1100, 5
    c = pkg.synth.config4(n=n, n_pairs=pairs, n_nodes=2)
    kfs = c["kfs"]
    ctx.set_option("force_exact", force_exact)
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    masks = [k["has_mp"] for k in kfs]
    ps = sh.PairShards([n] * pairs, world, blocks)
    gm = np.full((world, ps.per_rank, n), -1, np.int32)
    gd = np.full((world, ps.per_rank, n), -1.0, np.float32)
    gn = np.zeros((world, ps.per_rank), np.int32)
    for r in range(world):  # the ranks one after the other on this GPU
        mine = ps.local_units(r)
        up = ps.unit_pair[mine]
        job = pkg.PairJob(ctx, pkg.ASDM_JOB_TRIANGULATION, sets, masks, c["ia"][up], c["ib"][up], 0.6, 0, c["F12"][up],
                          c["epipoles"][up], c["scale_factors"], c["level_sigma2"], row_begin=ps.unit_r0[mine],
                          row_end=ps.unit_r1[mine])
        job.run()
        nm, m, d = job.fetch()
        for s, u in enumerate(mine):  # nothing outside the unit's rows
            a, b = ps.unit_r0[u], ps.unit_r1[u]
            assert (m[s, :a] == -1).all() and (m[s, b:n] == -1).all()
        gm[r, :len(mine)], gd[r, :len(mine)], gn[r, :len(mine)] = m[:, :n], d[:, :n], nm
        job.destroy()
    ctx.set_option("force_exact", 0)
    nm, match, dist = ps.merge(gm, gd, gn)
    onm, om, od = oracle.batch_triangulation(kfs, c["ia"], c["ib"], c["F12"], c["epipoles"], c["scale_factors"],
                                             c["level_sigma2"], 0, threads=oracle.max_threads())
    assert np.array_equal(match, om) and np.array_equal(dist.view(np.uint32), od.view(np.uint32))
    assert np.array_equal(nm, onm) and onm.sum() > 0
    for s in sets:
        s.release()


@pytest.mark.gpu
def test_row_ranges_rejected_for_greedy_searches(pkg, ctx):
    c = pkg.synth.config5(n=300, n_pairs=2, n_kf=3)
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in c["kfs"]]
    with pytest.raises(pkg.AsdmError):
        pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, [k["valid"] for k in c["kfs"]], c["ia"], c["ib"], 0.85, 1,
                    row_begin=[0, 0], row_end=[128, 128])
    for s in sets:
        s.release()


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["tri", "bow"])
def test_sharded_job_world1(pkg, ctx, oracle, kind):
    import torch
    sh = _shard_mod()
    stream = torch.cuda.Stream()
    ctx.set_stream(stream.cuda_stream)
    try:
        if kind == "tri":
            n, pairs = 900, 6
            c = pkg.synth.config4(n=n, n_pairs=pairs)
            sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in c["kfs"]]
            job = sh.ShardedPairJob(ctx, pkg.ASDM_JOB_TRIANGULATION, sets, [k["has_mp"] for k in c["kfs"]], c["ia"], c["ib"],
                                    0.6, 0, c["F12"], c["epipoles"], c["scale_factors"], c["level_sigma2"], blocks_per_pair=2)
            assert job.shards.n_units == 2 * pairs
            ref = oracle.batch_triangulation(c["kfs"], c["ia"], c["ib"], c["F12"], c["epipoles"], c["scale_factors"],
                                             c["level_sigma2"], 0, threads=oracle.max_threads())
        else:
            n = 700
            c = pkg.synth.config5(n=n, n_pairs=9, n_kf=5)
            sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in c["kfs"]]
            job = sh.ShardedPairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, [k["valid"] for k in c["kfs"]], c["ia"], c["ib"],
                                    c["nnratio"], c["check_ori"])
            ref = oracle.batch_bow_kf_kf(c["kfs"], c["ia"], c["ib"], c["nnratio"], c["check_ori"], threads=oracle.max_threads())
        job.run()
        nm, match, dist = job.fetch()
        assert np.array_equal(match[:, :n], ref[1]) and np.array_equal(dist[:, :n].view(np.uint32), ref[2].view(np.uint32))
        assert np.array_equal(nm, ref[0])
        job.destroy()
        for s in sets:
            s.release()
    finally:
        ctx.synchronize()
        ctx.set_stream(None)


def _publish_worker(rank, world, port, kind, n, q):
    """One rank of a two-process group that shares GPU 0: the gathered arrays of each process are mapped into the other
    (cudaIpc works between processes on one device too), the job publishes its rows from the resolve kernel."""
    try:
        import torch
        import torch.distributed as dist
        sys.path.insert(0, ROOT)
        pkg = importlib.import_module("asd-slam_b200")
        sh = importlib.import_module("asd-slam_b200.shard")
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        torch.cuda.set_device(0)
        ctx = pkg.Context(0)
        stream = torch.cuda.Stream()
        ctx.set_stream(stream.cuda_stream)
        if kind == "tri":
            pairs, blocks = 5, 2
            c = pkg.synth.config4(n=n, n_pairs=pairs)
            sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in c["kfs"]]
            masks = [k["has_mp"] for k in c["kfs"]]
            args = (0.6, 0, c["F12"], c["epipoles"], c["scale_factors"], c["level_sigma2"])
            job_kind = pkg.ASDM_JOB_TRIANGULATION
        else:
            pairs, blocks = 9, 1
            c = pkg.synth.config5(n=n, n_pairs=pairs, n_kf=5)
            sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in c["kfs"]]
            masks = [k["valid"] for k in c["kfs"]]
            args = (c["nnratio"], c["check_ori"])
            job_kind = pkg.ASDM_JOB_BOW_KF_KF
        per_rank = sh.PairShards([n] * pairs, world, blocks).per_rank
        gathered = sh.GatheredResults(ctx, rank, world, per_rank, n)
        outs = []
        for step in range(3):  # a new job per step on the same arrays: the generations alternate
            job = sh.ShardedPairJob(ctx, job_kind, sets, masks, c["ia"], c["ib"], *args, rank=rank, world=world,
                                    blocks_per_pair=blocks, gathered=gathered, chunks=3 if kind == "tri" else 1)
            job.run()
            outs.append(job.fetch())
            job.destroy()
        gathered.close()
        for s in sets:
            s.release()
        ctx.close()
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, True, [tuple(np.ascontiguousarray(x) for x in o) for o in outs]))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, False, traceback.format_exc()))


@pytest.mark.gpu
@pytest.mark.parametrize("kind,n", [("bow", 700), ("bow", 701), ("tri", 900)])
def test_world2_rows_published_by_the_resolve_kernel(pkg, oracle, kind, n):
    """SURVEY 8e on the GPU: two ranks (two processes sharing GPU 0, gloo for the handle exchange and the barrier), the
    match lists reach both ranks through asdm_job_set_publish -- no all-gather call -- and equal the unsharded oracle.
    n = 701: rows that cannot take 16-byte stores; 9 pairs over 2 ranks: a rank with an empty last slot."""
    import torch.multiprocessing as mp
    mctx = mp.get_context("spawn")
    q = mctx.Queue()
    port = _free_port()
    procs = [mctx.Process(target=_publish_worker, args=(r, 2, port, kind, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), [r for r in res if not r[1]]
    if kind == "tri":
        c = pkg.synth.config4(n=n, n_pairs=5)
        ref = oracle.batch_triangulation(c["kfs"], c["ia"], c["ib"], c["F12"], c["epipoles"], c["scale_factors"],
                                         c["level_sigma2"], 0, threads=oracle.max_threads())
    else:
        c = pkg.synth.config5(n=n, n_pairs=9, n_kf=5)
        ref = oracle.batch_bow_kf_kf(c["kfs"], c["ia"], c["ib"], c["nnratio"], c["check_ori"], threads=oracle.max_threads())
    assert ref[0].sum() > 0
    for _, _, outs in res:  # every rank, every step
        for nm, match, dist_ in outs:
            assert np.array_equal(nm, ref[0])
            assert np.array_equal(match[:, :n], ref[1])
            assert np.array_equal(dist_[:, :n].view(np.uint32), ref[2].view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("batch", [False, True])
def test_sets_from_replicated_device_rows(pkg, ctx, oracle, batch):
    """Keyframes uploaded into a device slab (ReplicatedDescriptors, world 1) and turned into sets device-to-device
    behave exactly like sets uploaded from the host."""
    import torch
    stream = torch.cuda.Stream()
    ctx.set_stream(stream.cuda_stream)
    try:
        n = 600
        c = pkg.synth.config5(n=n, n_pairs=4, n_kf=3)
        kfs = c["kfs"]
        h = [torch.from_numpy(k["desc"]).pin_memory() for k in kfs]
        rep = pkg.ReplicatedDescriptors(ctx, len(kfs), n)
        rep.upload(h)
        if batch:  # asdm_set_create_batch on device rows: one round of copies, ordered behind the all-gather
            sets = pkg.DescriptorSet.create_batch(ctx, [rep.rows(i, n) for i in range(len(kfs))], [k["kps"] for k in kfs],
                                                  c["grid"], [k["fv"] for k in kfs])
        else:
            sets = [pkg.DescriptorSet(ctx, rep.rows(i, n), k["kps"], c["grid"], k["fv"]) for i, k in enumerate(kfs)]
        for s, k in zip(sets, kfs):
            assert np.array_equal(s.read_desc(), k["desc"])
        job = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, [k["valid"] for k in kfs], c["ia"], c["ib"], c["nnratio"], 1)
        job.run()
        nm, match, dist = job.fetch()
        onm, om, od = oracle.batch_bow_kf_kf(kfs, c["ia"], c["ib"], c["nnratio"], 1)
        assert np.array_equal(match[:, :n], om) and np.array_equal(nm, onm)
        job.destroy()
        for s in sets:
            s.release()
    finally:
        ctx.synchronize()
        ctx.set_stream(None)
