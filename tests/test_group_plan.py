"""Host logic of the native multi-GPU group (csrc/group.cu) that needs no device: the unit plan of a group job
(asdm_group_plan_units -- the table asdm_group_job_prepare deals the work by) -- on CPU.

  * properties of the plan for ragged, empty and tile-straddling pairs at world 1 / 2 / 3 / 8;
  * a world_size-2 gloo run: every rank derives the plan itself, computes ONLY its own units (the CPU oracle stands in
    for the GPU, as the checker) and writes them where the library's publish epilogue would (row of the unit's pair,
    columns of its row block; unit counts in unit order); the summed arrays equal the un-sharded oracle result and every
    element was written by exactly one rank.
The GPU side of the same path is tests/test_group.py and tests/cpp/test_group.cc.
"""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _grp():
    importlib.import_module("asd-slam_b200")
    return importlib.import_module("asd-slam_b200.group")


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("blocks", [1, 2, 5, 40])
def test_plan_covers_every_row_once(world, blocks):
    G = _grp()
    n_rows = [2000, 1999, 128, 1, 0, 4000, 257, 129]
    up, r0, r1, rk = G.plan_units(n_rows, blocks, world)
    assert len(up) == len(r0) == len(r1) == len(rk)
    assert list(up) == sorted(up)                         # units in pair order, blocks of a pair ascending
    assert np.array_equal(rk, np.arange(len(up)) % world)  # round-robin deal
    seen = [np.zeros(max(n, 1), np.int32) for n in n_rows]
    for p, a, b in zip(up, r0, r1):
        assert 0 <= a <= b <= n_rows[p]
        assert a % 128 == 0 and (b % 128 == 0 or b == n_rows[p])  # cuts on tile boundaries of the tensor-core kernel
        seen[p][a:b] += 1
    for p, n in enumerate(n_rows):
        assert (seen[p][:n] == 1).all()
        n_units = int((up == p).sum())
        assert n_units == max(1, min(blocks, (n + 127) // 128))  # an empty pair still is one unit (its count is 0)
    # the ranks' loads differ by at most one unit
    per = np.bincount(rk, minlength=world)
    assert per.max() - per.min() <= 1


def test_plan_rejects_bad_arguments():
    G = _grp()
    with pytest.raises(Exception):
        G.plan_units([100, -1], 1, 2)
    with pytest.raises(Exception):
        G.plan_units([100], 0, 2)
    with pytest.raises(Exception):
        G.plan_units([100], 1, 0)
    up, r0, r1, rk = G.plan_units([], 3, 4)
    assert len(up) == 0


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
        G = importlib.import_module("asd-slam_b200.group")
        import oracle_py as oracle  # the checker stands in for the GPU here
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        n, pairs = 300, 7
        c = pkg.synth.config4(n=n, n_pairs=pairs, n_nodes=3)
        up, r0, r1, rk = G.plan_units([n] * pairs, blocks, world)
        plans = [None] * world
        dist.all_gather_object(plans, (up.tolist(), r0.tolist(), r1.tolist(), rk.tolist()))
        same_plan = all(p == plans[0] for p in plans)
        mine = np.nonzero(rk == rank)[0]
        mp_ = up[mine]
        onm, om, od = oracle.batch_triangulation(c["kfs"], c["ia"][mp_], c["ib"][mp_], c["F12"][mp_], c["epipoles"][mp_],
                                                 c["scale_factors"], c["level_sigma2"], 0)
        # what the publish epilogue does: row = the unit's pair, columns = its row block, the count in unit order
        match = torch.zeros((pairs, n), dtype=torch.int64)
        dbits = torch.zeros((pairs, n), dtype=torch.int64)
        written = torch.zeros((pairs, n), dtype=torch.int64)
        unit_nm = torch.zeros(len(up), dtype=torch.int64)
        for s, u in enumerate(mine):
            p, a, b = int(up[u]), int(r0[u]), int(r1[u])
            match[p, a:b] = torch.from_numpy(om[s, a:b].astype(np.int64))
            dbits[p, a:b] = torch.from_numpy(od[s, a:b].view(np.uint32).astype(np.int64))
            written[p, a:b] += 1
            unit_nm[u] = int((om[s, a:b] >= 0).sum())
        for t in (match, dbits, written, unit_nm):
            dist.all_reduce(t)
        nm = np.zeros(pairs, np.int64)
        np.add.at(nm, up, unit_nm.numpy())  # asdm_group_job_counts
        fnm, fm, fd = oracle.batch_triangulation(c["kfs"], c["ia"], c["ib"], c["F12"], c["epipoles"], c["scale_factors"],
                                                 c["level_sigma2"], 0)
        ok = bool(same_plan and (written.numpy() == 1).all() and np.array_equal(match.numpy(), fm.astype(np.int64)) and
                  np.array_equal(dbits.numpy(), fd.view(np.uint32).astype(np.int64)) and np.array_equal(nm, fnm) and
                  fnm.sum() > 0)
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, ok, int(fnm.sum())))
    except Exception as e:  # pragma: no cover
        q.put((rank, False, repr(e)))


@pytest.mark.parametrize("blocks", [1, 3])
def test_world2_gloo_units_merge_to_the_unsharded_result(blocks, oracle):
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
