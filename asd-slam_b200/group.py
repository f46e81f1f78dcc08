"""Thin mirror of the multi-GPU group API of libasdmatch_cuda.so (include/asdmatch.h "multi-GPU groups", csrc/group.cu).

Everything that moves data lives in the library: keyframes replicated over NVLink by the kernel that packs them, units
dealt round-robin over the ranks, finished rows stored by the resolve kernels into host memory (every rank over its own
PCIe link) and / or into every rank's GPU, device-side barriers.  This module only converts arguments.

  Group.create(devices)                     one process drives every GPU (what a C++ SLAM process would do)
  Group.join(ctx, rank, world, allgather)   one process per GPU (bench.py under torchrun); `allgather(bytes) -> [bytes]`
                                            is any all-gather of small host blobs -- torch_allgather() wraps
                                            torch.distributed for it.  It is called a few times while the group and its
                                            shared host blocks are set up, never on the search path.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import ASDM_GATHER_DEVICE, ASDM_GATHER_NO_DIST, ASDM_JOB_TRIANGULATION, KP_DTYPE, AsdmError, FeatVec, GJobDesc, Grid, SetDesc, ptr
from .matcher import Context, DescriptorSet, _f32, _i32, _u8


def torch_allgather(group=None):
    """allgather(blob) over a torch.distributed process group (NCCL: through a CUDA byte tensor; gloo: CPU)."""
    import torch
    import torch.distributed as dist

    def fn(blob):
        world = dist.get_world_size(group)
        if world == 1:
            return [blob]
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
        send = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev)
        recv = torch.empty(world * len(blob), dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(recv, send, group=group)
        raw = recv.cpu().numpy().tobytes()
        return [raw[r * len(blob):(r + 1) * len(blob)] for r in range(world)]
    return fn


def plan_units(query_rows, blocks_per_pair, world):
    """asdm_group_plan_units: the unit table of a group job -- (unit_pair, row0, row1, rank) arrays, host arithmetic only
    (works without a GPU; every rank computes the same table)."""
    lib = _lib.load()
    rows = np.ascontiguousarray(query_rows, np.int32)
    cap = max(1, len(rows)) * max(1, int(blocks_per_pair))
    up, r0, r1, rk = (np.empty(cap, np.int32) for _ in range(4))
    n = C.c_int(0)
    rc = lib.asdm_group_plan_units(_lib.ptr(rows), len(rows), int(blocks_per_pair), int(world), _lib.ptr(up), _lib.ptr(r0),
                                   _lib.ptr(r1), _lib.ptr(rk), cap, C.byref(n))
    if rc:
        raise _lib.AsdmError(f"asdm_group_plan_units: error {rc}")
    return up[:n.value].copy(), r0[:n.value].copy(), r1[:n.value].copy(), rk[:n.value].copy()


class Group:
    def __init__(self, lib, h, cb=None):
        self.lib, self.h, self._cb = lib, h, cb
        self.world = lib.asdm_group_world(h)
        self.local_count = lib.asdm_group_local_count(h)
        self.ranks = [lib.asdm_group_local_rank(h, l) for l in range(self.local_count)]
        self.ctxs = []
        for l in range(self.local_count):  # non-owning views of the members' contexts (streams, options, counters)
            c = Context.__new__(Context)
            c.lib, c.h, c.device = lib, C.c_void_p(lib.asdm_group_ctx(h, l)), None
            self.ctxs.append(c)

    @classmethod
    def create(cls, devices, window_bytes=0):
        lib = _lib.load()
        arr = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        rc = lib.asdm_group_create(arr, len(devices), window_bytes, C.byref(h))
        if rc != 0:
            raise AsdmError(f"asdm_group_create({list(devices)}) failed with {rc} (no B200, or no peer access between the GPUs)")
        return cls(lib, h)

    @classmethod
    def join(cls, ctx, rank, world, allgather, window_bytes=0):
        lib = ctx.lib

        def cb(user, send, recv, nbytes):
            try:
                parts = allgather(C.string_at(send, nbytes))
                C.memmove(recv, b"".join(parts), nbytes * len(parts))
                return 0
            except Exception:  # the library turns this into an error code; never let it cross the C frame
                return 1
        cfn = _lib.ALLGATHER_FN(cb)
        h = C.c_void_p()
        rc = lib.asdm_group_join(ctx.h, rank, world, window_bytes, cfn, None, C.byref(h))
        if rc != 0:
            raise AsdmError(f"asdm_group_join(rank {rank} of {world}) failed with {rc}")
        return cls(lib, h, cfn)

    def check(self, rc):
        if rc != 0:
            raise AsdmError(f"libasdmatch_cuda group error {rc}: {self.lib.asdm_group_last_error(self.h).decode()}")

    def host_array(self, shape, dtype):
        """numpy array over host memory every rank's GPU can store into and every process of the group can read."""
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        self.check(self.lib.asdm_group_host_alloc(self.h, max(n, 1), C.byref(p)))
        buf = (C.c_char * max(n, 1)).from_address(p.value)
        a = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
        return a

    def host_free(self, a):
        self.check(self.lib.asdm_group_host_free(self.h, C.c_void_p(a.ctypes.data)))

    def host_barrier(self):
        self.check(self.lib.asdm_group_host_barrier(self.h))

    def prepare_sets(self, descs, kps=None, grid=None, featvecs=None):
        """The argument block of sets_create, converted once (a keyframe batch that is uploaded again and again: the e2e
        arm of bench.py). Pass the result to sets_create(prepared=...)."""
        m = len(descs)
        g = None
        if grid is not None:
            g = grid if isinstance(grid, Grid) else Grid(*grid)
        arr = (SetDesc * m)()
        keep = []
        for k in range(m):
            d, a = descs[k], arr[k]
            assert d.dtype == np.float32 and d.ndim == 2 and d.shape[1] == 128 and d.flags["C_CONTIGUOUS"]
            a.desc, a.n, a.row_stride_bytes = d.ctypes.data, d.shape[0], 0
            if kps is not None:
                kp = np.ascontiguousarray(kps[k], dtype=KP_DTYPE)
                assert len(kp) == a.n
                a.kps = kp.ctypes.data
                keep.append(kp)
                if g is not None:
                    a.grid = C.pointer(g)
            if featvecs is not None:
                node_ids, off, idx = featvecs[k]
                node_ids, off, idx = np.ascontiguousarray(node_ids, dtype=np.uint32), _i32(off), _i32(idx)
                fv = FeatVec(len(node_ids), ptr(node_ids), ptr(off), ptr(idx))
                a.featvec = C.pointer(fv)
                keep.append((node_ids, off, idx, fv))
        return (arr, m, keep, list(descs))

    def sets_create(self, descs=None, kps=None, grid=None, featvecs=None, prepared=None):
        """asdm_group_sets_create: keyframe k crosses PCIe once (rank k % world uploads descs[k]) and is replicated by
        the owner's pack kernel. Returns sets[k][l] = DescriptorSet of keyframe k on local member l. descs[k] must stay
        alive and unmodified until the next synchronising call."""
        arr, m, _keep, descs = prepared if prepared is not None else self.prepare_sets(descs, kps, grid, featvecs)
        L = self.local_count
        hs = (C.c_void_p * (m * L))()
        self.check(self.lib.asdm_group_sets_create(self.h, arr, m, hs))
        out = []
        for k in range(m):
            row = []
            for l in range(L):
                s = DescriptorSet.__new__(DescriptorSet)
                s.ctx, s.h, s.n, s._desc_keepalive = self.ctxs[l], C.c_void_p(hs[k * L + l]), int(arr[k].n), descs[k]
                row.append(s)
            out.append(row)
        return out

    @staticmethod
    def release_sets(sets):
        for row in sets:
            for s in row:
                s.release()

    def close(self):
        if self.h:
            self.lib.asdm_group_destroy(self.h)
            self.h = None


class GroupJob:
    """A batched keyframe-pair search dealt over the ranks of a Group (asdm_group_job_*): unit u -- a pair, or one of
    `blocks_per_pair` row blocks of a pair -- runs on rank u % world."""

    def __init__(self, group, kind, sets, masks, ia, ib, nnratio=0.6, check_ori=True, F12=None, epipoles=None, scale_factors=None,
                 level_sigma2=None, blocks_per_pair=1, gather_device=False, gather_dist=True):
        self.g = group
        L, n_kf = group.local_count, len(sets)
        self.n_pairs = len(ia)
        self._keep = k = dict(masks=[_u8(m) for m in masks], ia=_i32(ia), ib=_i32(ib))
        set_arr = (C.c_void_p * (n_kf * L))(*[sets[i][l].h for i in range(n_kf) for l in range(L)])
        mask_arr = (C.c_void_p * n_kf)(*[m.ctypes.data for m in k["masks"]])
        d = GJobDesc()
        j = d.job
        j.kind, j.n_kf, j.sets, j.masks = kind, n_kf, set_arr, mask_arr
        j.n_pairs, j.ia, j.ib = self.n_pairs, ptr(k["ia"]), ptr(k["ib"])
        j.nnratio, j.check_ori = float(nnratio), int(bool(check_ori))
        if kind == ASDM_JOB_TRIANGULATION:
            k["F12"], k["ep"] = _f32(F12).reshape(-1), _f32(epipoles).reshape(-1)
            k["sf"], k["ls"] = _f32(scale_factors), _f32(level_sigma2)
            j.F12, j.epipoles, j.scale_factors, j.level_sigma2 = ptr(k["F12"]), ptr(k["ep"]), ptr(k["sf"]), ptr(k["ls"])
            j.n_levels = len(k["sf"])
        d.blocks_per_pair = int(blocks_per_pair)
        d.gather = (ASDM_GATHER_DEVICE | (0 if gather_dist else ASDM_GATHER_NO_DIST)) if gather_device else 0
        h = C.c_void_p()
        group.check(group.lib.asdm_group_job_prepare(group.h, C.byref(d), C.byref(h)))
        self.h = h
        self.n_units = group.lib.asdm_group_job_units(h)
        self.stride = group.lib.asdm_group_job_stride(h)

    def run(self, h_match=None, h_dist=None, h_unit_nm=None):
        """Launches the local units; the kernels store finished rows into the given host arrays (Group.host_array):
        h_match / h_dist [n_pairs, stride] in pair order, h_unit_nm [n_units]."""
        self.g.check(self.g.lib.asdm_group_job_run(self.h, ptr(h_match), ptr(h_dist), ptr(h_unit_nm)))

    def wait(self):
        self.g.check(self.g.lib.asdm_group_job_wait(self.h))

    def flush(self):
        """Orders what is queued next on the members' streams behind the exchange of the last run (no host wait)."""
        self.g.check(self.g.lib.asdm_group_job_flush(self.h))

    def counts(self, h_unit_nm):
        nm = np.empty(self.n_pairs, np.int32)
        self.g.check(self.g.lib.asdm_group_job_counts(self.h, ptr(h_unit_nm), ptr(nm)))
        return nm

    def fetch(self):
        """The whole batch on this process: (nm[n_pairs], match[n_pairs, stride], dist[n_pairs, stride])."""
        match = np.empty((self.n_pairs, self.stride), np.int32)
        dist = np.empty((self.n_pairs, self.stride), np.float32)
        nm = np.empty(self.n_pairs, np.int32)
        self.g.check(self.g.lib.asdm_group_job_fetch(self.h, ptr(match), ptr(dist), ptr(nm)))
        return nm, match, dist

    def gathered(self, l=0):
        """Device pointers (match, dist, unit_nm) of the records of all units on local member l (gather_device=True)."""
        m, d, n = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self.g.check(self.g.lib.asdm_group_job_gathered(self.h, l, C.byref(m), C.byref(d), C.byref(n)))
        return m.value, d.value, n.value

    def destroy(self):
        if self.h:
            self.g.lib.asdm_group_job_destroy(self.h)
            self.h = None


class StreamedGroupSweep:
    """A keyframe-pair sweep over keyframes that are NOT yet resident, on a Group: the keyframes are uploaded in index
    order, in waves (each keyframe once over PCIe by its owner, replicated by the pack kernel), and the pairs are searched
    in readiness order -- wave w holds the pairs whose later keyframe arrived with wave w and is dealt over the ranks as
    soon as that wave is queued, so the searches overlap the uploads of the later waves (matcher.StreamedPairSweep is the
    single-GPU form). Finished rows are stored by the resolve kernels into the caller's host arrays; row r belongs to pair
    `order[r]`."""

    def __init__(self, group, kind, descs, kps, grid, featvecs, ia, ib, nnratio=0.6, check_ori=True, bounds=None):
        self.g, self.kind = group, kind
        self.n_kf = len(descs)
        self.ia, self.ib = _i32(ia), _i32(ib)
        self.nnratio, self.check_ori = nnratio, check_ori
        if bounds is None:
            bounds = sorted({max(2, int(round(self.n_kf * f))) for f in (0.375, 0.5625, 0.72, 0.845, 0.94, 1.0)})
        self.bounds = [int(b) for b in bounds]
        assert self.bounds[-1] == self.n_kf
        later = np.maximum(self.ia, self.ib)
        self.order = np.argsort(later, kind="stable").astype(np.int64)
        cuts = np.searchsorted(later[self.order], self.bounds, side="left")
        self.slices = [(int(a), int(b)) for a, b in zip(np.concatenate(([0], cuts[:-1])), cuts)]
        lo = [0] + self.bounds[:-1]
        # argument blocks converted once: the same keyframes are uploaded step after step (bench.py e2e arm)
        self.prepared = [group.prepare_sets(descs[a:b], kps[a:b], grid, featvecs[a:b]) for a, b in zip(lo, self.bounds)]

    def run(self, masks, h_match, h_unit_nm, timers=None):
        """h_match [n_pairs, stride] / h_unit_nm [n_pairs]: host arrays of Group.host_array. Returns the sets (the caller
        releases them with Group.release_sets). timers: dict that receives the host seconds spent per phase."""
        import time
        sets, jobs = [], []
        tick = time.perf_counter
        t = tick()

        def lap(key):
            nonlocal t
            if timers is not None:
                now = tick()
                timers[key] = timers.get(key, 0.0) + now - t
                t = now
        for (a, b), prep, bound in zip(self.slices, self.prepared, self.bounds):
            sets.extend(self.g.sets_create(prepared=prep))
            lap("sets_create")
            if b <= a:
                continue
            sel = self.order[a:b]
            jb = GroupJob(self.g, self.kind, sets[:bound], masks[:bound], self.ia[sel], self.ib[sel], self.nnratio, self.check_ori)
            lap("job_prepare")
            jb.run(h_match[a:b], None, h_unit_nm[a:b])
            lap("job_run")
            jobs.append(jb)
        for jb in jobs:
            jb.wait()
            lap("job_wait")
            jb.destroy()
            lap("job_destroy")
        return sets
