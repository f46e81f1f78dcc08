"""Sharding of batched keyframe-pair searches over GPUs.

The units of the two batched workloads are independent: the loop bodies of LocalMapping.cc:335-370 (one
SearchForTriangulation per covisible keyframe, "C4") and LoopClosing.cc:291-318 (one SearchByBoW per loop candidate,
"C5").  Keyframe descriptor sets are replicated on every GPU (about 3 MB per 4000-keypoint keyframe), the units are
dealt round-robin to the ranks, every rank writes fixed-stride records (int32 index + fp32 distance per query row, one
count per unit) and ONE all-gather per result array makes the whole batch visible on every rank.  There is no other
data-path collective.

SearchForTriangulation has no cross-query state (vbMatched2 is never set, ORBmatcher.cc:669-822) and its only caller
switches the rotation check off (LocalMapping.cc:309), so a pair may also be split into 128-row-aligned query
row-blocks: 20 covisible pairs become >= 40 units and divide evenly over 8 GPUs.  The BoW searches are greedy and are
sharded by whole pairs only.

`PairShards` is pure host logic (numpy); `all_gather_results` works on CPU tensors over gloo and on CUDA tensors over
NCCL; `ShardedPairJob` ties both to the device-resident `PairJob`.  On GPUs the exchange is fused into the job:
`GatheredResults` maps every rank's gathered arrays into every other rank (cudaIpc over NVLink) and the job's resolve
kernel stores each finished row into all of them (asdm_job_set_publish), so what remains behind the kernels is a
one-word barrier instead of three all-gathers of 16 MB per rank.
"""
import numpy as np

from ._lib import ASDM_JOB_TRIANGULATION


class PairShards:
    """Static round-robin partition of (pair, row-block) units over `world` ranks.

    n_rows: number of query rows of each pair (len n_pairs) -- keyframe sizes of sets[ia[p]].
    blocks_per_pair > 1 splits every pair into that many row blocks whose boundaries are multiples of `align` rows
    (the tensor-core tile height), so no tile is computed twice.
    """

    def __init__(self, n_rows, world, blocks_per_pair=1, align=128):
        n_rows = np.asarray(n_rows, dtype=np.int64).reshape(-1)
        self.n_pairs = len(n_rows)
        self.world = int(world)
        assert self.world >= 1 and blocks_per_pair >= 1
        # vectorised (a loop-closure sweep over 8 GPUs deals 4096 pairs per step): unit b of pair p covers the rows
        # [cut_b, cut_{b+1}) with cut_b = (tiles * b // nb) * align, cut_nb = n
        tiles = -(-n_rows // align)
        nb = np.maximum(1, np.minimum(blocks_per_pair, tiles))
        up = np.repeat(np.arange(self.n_pairs, dtype=np.int64), nb)
        b = np.arange(len(up), dtype=np.int64) - np.repeat(np.cumsum(nb) - nb, nb)
        t_u, nb_u, n_u = tiles[up], nb[up], n_rows[up]
        c0 = (t_u * b // nb_u) * align
        c1 = np.where(b + 1 == nb_u, n_u, (t_u * (b + 1) // nb_u) * align)
        self.unit_pair = up.astype(np.int32)
        self.unit_r0 = np.minimum(c0, n_u).astype(np.int32)
        self.unit_r1 = np.minimum(c1, n_u).astype(np.int32)
        self.n_units = int(len(up))
        self.per_rank = -(-self.n_units // self.world) if self.n_units else 0
        self.split_rows = blocks_per_pair > 1

    def local_units(self, rank):
        """Global unit ids owned by `rank`, in local order (unit u lives at rank u % world, slot u // world)."""
        return np.arange(rank, self.n_units, self.world, dtype=np.int64)

    def merge(self, g_match, g_dist, g_nm):
        """Rank-major gathered records -> per-pair results.

        g_match, g_dist: [world, per_rank, stride]; g_nm: [world, per_rank].  Returns (nm[n_pairs], match[n_pairs,
        stride], dist[n_pairs, stride]) exactly as a single un-sharded job over all pairs would report them.
        """
        g_match = np.asarray(g_match).reshape(self.world, self.per_rank, -1)
        g_dist = np.asarray(g_dist).reshape(self.world, self.per_rank, -1)
        g_nm = np.asarray(g_nm).reshape(self.world, self.per_rank)
        stride = g_match.shape[2]
        match = np.full((self.n_pairs, stride), -1, np.int32)
        dist = np.full((self.n_pairs, stride), -1.0, np.float32)
        nm = np.zeros(self.n_pairs, np.int32)
        u = np.arange(self.n_units)
        r, sl = u % self.world, u // self.world
        if not self.split_rows:  # one unit per pair: a pure permutation of the records
            match[self.unit_pair] = g_match[r, sl]
            dist[self.unit_pair] = g_dist[r, sl]
            nm[self.unit_pair] = g_nm[r, sl]
            return nm, match, dist
        for k in range(self.n_units):
            p, a, b = self.unit_pair[k], int(self.unit_r0[k]), int(self.unit_r1[k])
            match[p, a:b] = g_match[r[k], sl[k], a:b]
            dist[p, a:b] = g_dist[r[k], sl[k], a:b]
            nm[p] += g_nm[r[k], sl[k]]
        return nm, match, dist


def all_gather_results(local, per_rank, group=None, out=None):
    """All-gathers one fixed-stride record array.  local: torch tensor [n_local, ...] (CPU over gloo, CUDA over NCCL),
    n_local <= per_rank (ragged tails are padded with -1).  Returns a tensor [world, per_rank, ...] on every rank."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if local.shape[0] != per_rank:
        padded = torch.full((per_rank,) + tuple(local.shape[1:]), -1, dtype=local.dtype, device=local.device)
        padded[:local.shape[0]] = local
        local = padded
    local = local.contiguous()
    if out is None:
        out = torch.empty((world,) + tuple(local.shape), dtype=local.dtype, device=local.device)
    if world == 1:
        out[0].copy_(local)
    else:
        dist.all_gather_into_tensor(out.view(-1), local.view(-1), group=group)
    return out


class ReplicatedDescriptors:
    """Keyframe descriptors replicated on every GPU of the group ("replicate each keyframe's device
    descriptor set to all GPUs once at keyframe creation") without sending them N times over PCIe: rank r uploads the
    keyframes k = r (mod world) from its pinned host buffers, one all-gather over NVLink hands every rank all of them.
    `rows(k)` is then a device view to create the local DescriptorSet from (device-to-device, stream-ordered).
    """

    def __init__(self, ctx, n_kf, n_max, rank=0, world=1, group=None):
        import torch
        self._torch, self.ctx, self.rank, self.world, self.group = torch, ctx, rank, world, group
        self.n_kf, self.n_max = n_kf, n_max
        self.per_rank = -(-n_kf // world)
        dev = torch.device("cuda", ctx.device)
        self.slab = torch.zeros((world, self.per_rank, n_max, 128), dtype=torch.float32, device=dev)
        self.mine = torch.zeros((self.per_rank, n_max, 128), dtype=torch.float32, device=dev) if world > 1 else self.slab[0]
        self._ext = None

    def _stream(self):
        p = self.ctx.stream()
        if self._ext is None or self._ext[0] != p:
            self._ext = (p, self._torch.cuda.ExternalStream(p, device=self.slab.device))
        return self._ext[1]

    def upload(self, h_desc):
        """h_desc[k]: pinned host torch tensor [n_k, 128] fp32 of keyframe k (only this rank's are read).  Queued on
        the context stream: H2D of the local keyframes, then the NVLink all-gather."""
        torch = self._torch
        with torch.cuda.stream(self._stream()):
            for k in range(self.rank, self.n_kf, self.world):
                self.mine[k // self.world, :h_desc[k].shape[0]].copy_(h_desc[k], non_blocking=True)
            if self.world > 1:
                import torch.distributed as dist
                dist.all_gather_into_tensor(self.slab.view(-1), self.mine.view(-1), group=self.group)

    def rows(self, k, n):
        return self.slab[k % self.world, k // self.world, :n]


class GatheredResults:
    """Rank-major gathered record arrays ([world, per_rank, stride] match / dist, [world, per_rank] counts) of a
    sharded job on every rank, writable by every other rank of the node.  Two transports:

      * "multicast": torch symmetric memory (CUDA VMM allocations bound to one NVSwitch multicast object, NVLS); the
        job stores a row ONCE to the multicast address and the switch delivers it to all ranks, its own included;
      * "peer": asdm_shared_alloc + cudaIpc mappings; the job stores the row to every rank's array over NVLink.

    "auto" takes multicast when the group is NCCL, the rows allow 16-byte stores and the rendezvous yields a multicast
    mapping, peer otherwise.  Two generations alternate from run to run: a rank that is one step ahead (it cannot be
    more, every step ends with a barrier of the group) never overwrites what a slower rank is still reading.  Create
    once, hand to every ShardedPairJob of the sweep; collective (all ranks must construct it).
    """

    GENERATIONS = 2

    def __init__(self, ctx, rank, world, per_rank, stride, group=None, transport="auto"):
        import torch
        import torch.distributed as dist

        self.ctx, self.rank, self.world, self.group = ctx, rank, world, group
        self.per_rank, self.stride = per_rank, stride
        rows = world * per_rank
        al = lambda x: (x + 255) & ~255
        self._o_dist = al(rows * stride * 4)
        self._o_nm = self._o_dist + al(rows * stride * 4)
        self._gen_bytes = self._o_nm + al(rows * 4)
        total = self.GENERATIONS * self._gen_bytes
        dev = torch.device("cuda", ctx.device)
        self._torch = torch
        self.base = self.mc_base = self._symm = None
        want_mc = transport == "multicast" or (transport == "auto" and world > 1 and stride % 4 == 0 and
                                               dist.get_backend(group) == "nccl")
        if want_mc:
            try:
                import torch.distributed._symmetric_memory as symm_mem
                t = symm_mem.empty(total // 4, dtype=torch.int32, device=dev)
                h = symm_mem.rendezvous(t, group=group if group is not None else dist.group.WORLD)
                ok = torch.tensor([1 if h.multicast_ptr else 0], dtype=torch.int32, device=dev)
                dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
                if int(ok.item()):
                    self._symm, self.mc_base, whole = (t, h), int(h.multicast_ptr), t
            except Exception:
                if transport == "multicast":
                    raise
        if self.mc_base is None:
            if transport == "multicast":
                raise RuntimeError("no NVLink multicast mapping on this system")
            self.base, handle = ctx.shared_alloc(total)
            handles = [None] * world
            if world > 1:
                dist.all_gather_object(handles, handle, group=group)
            self.bases = [self.base if r == rank else ctx.shared_open(handles[r]) for r in range(world)]
            whole = torch.as_tensor(_DevArray(self.base, (total // 4,), "<i4"), device=dev)
        self.transport = "multicast" if self.mc_base is not None else "peer"
        whole.fill_(-1)  # rows no unit is dealt to stay "no match"; their counts are zeroed below
        self.views = []
        n_md = rows * stride
        for g in range(self.GENERATIONS):
            o = g * self._gen_bytes // 4
            m = whole[o:o + n_md].view(world, per_rank, stride)
            d = whole[o + self._o_dist // 4:o + self._o_dist // 4 + n_md].view(torch.float32).view(world, per_rank, stride)
            n = whole[o + self._o_nm // 4:o + self._o_nm // 4 + rows].view(world, per_rank)
            d.fill_(-1.0)
            n.zero_()
            self.views.append((m, d, n))
        self._token = torch.zeros(1, dtype=torch.int32, device=dev)
        self.gen = self.GENERATIONS - 1
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier(group=group)  # every mapping exists and is initialised before anyone publishes

    def next_generation(self):
        self.gen = (self.gen + 1) % self.GENERATIONS
        return self.gen

    def ptrs(self, gen):
        """(match, dist, nm) destination pointers of generation `gen`: one multicast address each, or one per rank."""
        o = gen * self._gen_bytes
        bases = [self.mc_base] if self.mc_base is not None else self.bases
        return ([b + o for b in bases], [b + o + self._o_dist for b in bases], [b + o + self._o_nm for b in bases])

    def current(self):
        return self.views[self.gen]

    def barrier(self, stream):
        """Stream-ordered barrier of the group: once it has run here, every rank's kernels queued before its own
        barrier -- hence all their stores into this rank's arrays -- are complete."""
        if self.world > 1:
            import torch.distributed as dist
            if dist.get_backend(self.group) == "nccl":
                with self._torch.cuda.stream(stream):
                    dist.all_reduce(self._token, group=self.group)
            else:  # host-side group (tests: two ranks on one GPU over gloo)
                self.ctx.synchronize()
                dist.barrier(group=self.group)

    def close(self):
        if self.base is None and self._symm is None:
            return
        self.ctx.synchronize()
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=self.group)  # nobody is still storing into a mapping that is about to go away
        self.views = []
        if self._symm is not None:
            self._symm = self.mc_base = None
            return
        for r, b in enumerate(self.bases):
            if r != self.rank:
                self.ctx.shared_close(b)
        self.ctx.shared_free(self.base)
        self.base = None


class _DevArray:
    """Exposes a raw device pointer of the library to torch (zero copy)."""

    def __init__(self, p, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(p), False), "version": 2}


class ShardedPairJob:
    """A batched keyframe-pair search whose units are dealt over the ranks of a torch.distributed group.

    Every rank passes the SAME global description (sets replicated on its own GPU, the same ia / ib / F12 / epipoles);
    the rank prepares a device-resident PairJob for its own units only.  run() launches the local kernels, gather()
    queues the all-gathers behind them on the context's stream (NCCL over NVLink), fetch() returns the merged per-pair
    results on the host.
    """

    def __init__(self, ctx, kind, sets, masks, ia, ib, nnratio=0.6, check_ori=True, F12=None, epipoles=None,
                 scale_factors=None, level_sigma2=None, rank=0, world=1, group=None, blocks_per_pair=1, gathered=None,
                 chunks=1, publish_dist=True):
        """gathered: a GatheredResults of matching shape -- the job then publishes its rows into every rank's arrays
        from its own resolve kernel (in `chunks` sub-jobs, see below) and gather() is a barrier; None: NCCL / gloo
        all-gathers of the local outputs."""
        import torch
        from .matcher import PairJob

        ia, ib = np.asarray(ia, np.int32), np.asarray(ib, np.int32)
        if blocks_per_pair > 1 and (kind != ASDM_JOB_TRIANGULATION or check_ori):
            raise ValueError("row-block units need a triangulation job without rotation check")
        self.ctx, self.rank, self.world, self.group = ctx, rank, world, group
        self.publish_dist = bool(publish_dist)  # False: only index lists + counts cross NVLink (g_dist rows are stale)
        set_n = np.array([s.n for s in sets], np.int64)
        self.shards = PairShards(set_n[ia], world, blocks_per_pair)
        mine = self.shards.local_units(rank)
        up = self.shards.unit_pair[mine]
        # A publishing job may run as `chunks` sub-jobs over contiguous slices of the local units, back to back on one
        # stream: the rows of slice k cross NVLink while slice k+1 is being searched, so only the last slice's transfer
        # is left in front of the barrier.  Measured on 2 x B200 (C5, 512 pairs per rank): 4 chunks cost 0.15 ms of
        # kernel tails and launch gaps per step and can save at most the 0.1 ms the transfer takes -- default 1.
        n_chunks = max(1, min(int(chunks), len(mine))) if gathered is not None else 1
        cuts = [len(mine) * k // n_chunks for k in range(n_chunks + 1)]
        self.jobs, self.chunk_row0 = [], cuts[:-1]
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            kw = {}
            if kind == ASDM_JOB_TRIANGULATION:
                kw = dict(F12=np.asarray(F12, np.float32).reshape(-1, 9)[up[lo:hi]],
                          epipoles=np.asarray(epipoles, np.float32).reshape(-1, 2)[up[lo:hi]],
                          scale_factors=scale_factors, level_sigma2=level_sigma2)
                if blocks_per_pair > 1:
                    kw.update(row_begin=self.shards.unit_r0[mine[lo:hi]], row_end=self.shards.unit_r1[mine[lo:hi]])
            self.jobs.append(PairJob(ctx, kind, sets, masks, ia[up[lo:hi]], ib[up[lo:hi]], nnratio, check_ori, **kw))
        self.job = self.jobs[0]
        self.n_local = len(mine)
        # one stride for every rank: the largest query set of the batch (jobs size theirs from their own units)
        self.stride = int(set_n[ia].max()) if len(ia) else 1
        assert all(j.stride <= self.stride for j in self.jobs)
        dev = torch.device("cuda", ctx.device)
        self._torch = torch
        pr = self.shards.per_rank
        self.gathered = gathered
        if gathered is not None:
            if (gathered.world, gathered.per_rank) != (world, pr) or gathered.stride < self.stride:
                raise ValueError("GatheredResults shaped for another job")
            self.stride = gathered.stride
            self.g_match, self.g_dist, self.g_nm = gathered.current()
            return
        self.g_match = torch.empty((world, pr, self.stride), dtype=torch.int32, device=dev)
        self.g_dist = torch.empty((world, pr, self.stride), dtype=torch.float32, device=dev)
        self.g_nm = torch.empty((world, pr), dtype=torch.int32, device=dev)
        self._direct = self.n_local == pr and self.job.stride == self.stride and pr > 0
        if self._direct:
            pm, pd_, pn = self.job.device_outputs()
            self.l_match = torch.as_tensor(_DevArray(pm, (pr, self.stride), "<i4"), device=dev)
            self.l_dist = torch.as_tensor(_DevArray(pd_, (pr, self.stride), "<f4"), device=dev)
            self.l_nm = torch.as_tensor(_DevArray(pn, (pr,), "<i4"), device=dev)

    def run(self):
        g = self.gathered
        self.launches = 0
        if g is not None:  # the rows go out from the resolve kernel itself, into the next generation of arrays
            gen = g.next_generation()
            self.g_match, self.g_dist, self.g_nm = g.current()
            pm, pd_, pn = g.ptrs(gen)
        for jb, r0 in zip(self.jobs, self.chunk_row0):
            if g is not None and self.n_local:
                jb.set_publish(pm, pd_, pn, self.rank * g.per_rank + r0, g.stride, multicast=g.transport == "multicast",
                               dist=self.publish_dist)
            jb.run()
            self.launches += self.ctx.stat("gpu_launches")

    def settle(self):
        self.retries = self.tc_pairs = 0
        for jb in self.jobs:
            jb.settle()
            self.retries += self.ctx.stat("retries")
            self.tc_pairs += self.ctx.stat("tc_pairs")

    def _locals(self):
        if self._direct:
            return self.l_match, self.l_dist, self.l_nm
        torch = self._torch
        dev = self.g_match.device
        m = torch.full((self.n_local, self.stride), -1, dtype=torch.int32, device=dev)
        d = torch.full((self.n_local, self.stride), -1.0, dtype=torch.float32, device=dev)
        n = torch.zeros((self.n_local,), dtype=torch.int32, device=dev)
        if self.n_local:
            pm, pd_, pn = self.job.device_outputs()
            js = self.job.stride
            m[:, :js] = torch.as_tensor(_DevArray(pm, (self.n_local, js), "<i4"), device=dev)
            d[:, :js] = torch.as_tensor(_DevArray(pd_, (self.n_local, js), "<f4"), device=dev)
            n.copy_(torch.as_tensor(_DevArray(pn, (self.n_local,), "<i4"), device=dev))
        return m, d, n

    def _stream(self):
        p = self.ctx.stream()
        if getattr(self, "_ext", None) is None or self._ext[0] != p:
            self._ext = (p, self._torch.cuda.ExternalStream(p, device=self.g_match.device))
        return self._ext[1]

    def gather(self):
        """Queues the exchange on the stream the job ran on (set with Context.set_stream to a torch stream): a barrier
        when the job publishes its rows itself, the all-gathers of the local outputs otherwise."""
        torch = self._torch
        if self.gathered is not None:
            self.gathered.barrier(self._stream())
            return
        with torch.cuda.stream(self._stream()):
            m, d, n = self._locals()
            pr = self.shards.per_rank
            all_gather_results(m, pr, self.group, self.g_match)
            all_gather_results(d, pr, self.group, self.g_dist)
            all_gather_results(n, pr, self.group, self.g_nm)

    def fetch(self):
        """Merged per-pair results on the host (every rank gets the whole batch)."""
        self.settle()
        self.gather()
        self.ctx.synchronize()
        return self.shards.merge(self.g_match.cpu().numpy(), self.g_dist.cpu().numpy(), self.g_nm.cpu().numpy())

    def fetch_gathered(self, h_match, h_dist, h_nm):
        """gather() + device->host copies of the rank-major records into caller-provided pinned torch tensors
        ([world, per_rank, stride] / [world, per_rank]); `shards.merge` turns them into per-pair results.
        h_match / h_dist may be None on ranks whose host only needs the per-pair counts (worker ranks: the match lists
        stay on the GPU, the consumer process reads them on its own rank)."""
        self.settle()
        self.gather()
        with self._torch.cuda.stream(self._stream()):
            if h_match is not None:
                h_match.copy_(self.g_match, non_blocking=True)
            if h_dist is not None:
                h_dist.copy_(self.g_dist, non_blocking=True)
            h_nm.copy_(self.g_nm, non_blocking=True)
        self.ctx.synchronize()

    def destroy(self):
        for jb in self.jobs:
            jb.destroy()
