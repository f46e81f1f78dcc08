"""Host-side mirror of the reference matcher interface over the C ABI (include/asdmatch.h).

Names and argument meaning follow ORB_SLAM2::ORBmatcher (reference src/vslam/include/ORBmatcher.h:48-90):
`ORBmatcher(nnratio, checkOri)` with SearchByBoW / SearchForTriangulation / SearchByProjection /
SearchForInitialization.  Frame / KeyFrame state is passed as a `DescriptorSet` (device-resident descriptors,
keypoints, grid, FeatureVector -- uploaded once, reference Frame.h:147 / KeyFrame.h:199) plus the small per-call
masks the reference reads from MapPoint pointers.  MapPoint* never crosses the boundary: results are index arrays.

Every call goes to libasdmatch_cuda.so; there is no CPU path in this package.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import (ASDM_JOB_BOW_KF_KF, ASDM_JOB_TRIANGULATION, ASDM_PROJ_BEST_ONLY, ASDM_PROJ_RATIO_LEVEL, KP_DTYPE,
                   PROJ_QUERY_DTYPE, TH_HIGH, TH_LOW, AsdmError, BowdbGates, FeatVec, Grid, JobDesc, ProjParams, SetDesc, VocabDesc,
                   ptr)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class Context:
    """One per GPU (asdm_create). Owns a CUDA stream; calls are serialised inside the library."""

    def __init__(self, device=0):
        self.lib = _lib.load()
        h = C.c_void_p()
        rc = self.lib.asdm_create(int(device), C.byref(h))
        if rc != 0:
            raise AsdmError(f"asdm_create(device={device}) failed with {rc}: no usable B200 "
                            "(the matcher has no CPU fallback)")
        self.h = h
        self.device = int(device)

    def fork(self):
        """asdm_fork: a context for another thread on the same GPU that may search this context's sets."""
        c = Context.__new__(Context)
        c.lib, c.device = self.lib, self.device
        h = C.c_void_p()
        self.check(self.lib.asdm_fork(self.h, C.byref(h)))
        c.h = h
        return c

    def check(self, rc):
        if rc != 0:
            raise AsdmError(f"libasdmatch_cuda error {rc}: {self.lib.asdm_last_error(self.h).decode()}")

    def set_option(self, key, value):
        self.check(self.lib.asdm_set_option(self.h, key.encode(), int(value)))

    def stat(self, key):
        v = C.c_int64()
        self.check(self.lib.asdm_get_stat(self.h, key.encode(), C.byref(v)))
        return v.value

    def stream(self):
        return self.lib.asdm_stream(self.h)

    def set_stream(self, stream_ptr):
        self.check(self.lib.asdm_set_stream(self.h, C.c_void_p(stream_ptr)))

    def synchronize(self):
        self.check(self.lib.asdm_synchronize(self.h))

    def close(self):
        if self.h:
            self.lib.asdm_destroy(self.h)
            self.h = None

    def shared_alloc(self, nbytes):
        """Device memory other processes of the node can map (asdm_shared_alloc): (pointer, 64-byte handle)."""
        p, h = C.c_void_p(), (C.c_uint8 * 64)()
        self.check(self.lib.asdm_shared_alloc(self.h, nbytes, C.byref(p), h))
        return p.value, bytes(h)

    def shared_open(self, handle):
        p, h = C.c_void_p(), (C.c_uint8 * 64).from_buffer_copy(handle)
        self.check(self.lib.asdm_shared_open(self.h, h, C.byref(p)))
        return p.value

    def shared_close(self, p):
        self.check(self.lib.asdm_shared_close(self.h, C.c_void_p(p)))

    def shared_free(self, p):
        self.check(self.lib.asdm_shared_free(self.h, C.c_void_p(p)))

    def pinned(self, shape, dtype):
        """numpy array backed by page-locked memory (asdm_host_alloc); keep the returned array alive."""
        return PinnedArray(self.lib, shape, dtype).a


class PinnedArray:
    def __init__(self, lib, shape, dtype):
        self.lib = lib
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        if lib.asdm_host_alloc(C.byref(p), max(n, 1)) != 0:
            raise AsdmError("asdm_host_alloc failed")
        self.p = p
        buf = (C.c_char * max(n, 1)).from_address(p.value)
        self.a = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
        _PINNED.append(self)


_PINNED = []  # pinned blocks live for the process (bench/test helper)


class DescriptorSet:
    """Device-resident descriptors (+ keypoints, grid, FeatureVector) of one Frame / KeyFrame / map-point pool."""

    def __init__(self, ctx, desc, kps=None, grid=None, featvec=None, async_upload=False):
        """async_upload: asdm_set_create_async -- `desc` must stay alive and unmodified until the next synchronising
        call on the context (use Context.pinned() memory for a true async DMA)."""
        self.ctx = ctx
        dev_ptr = None
        if hasattr(desc, "__cuda_array_interface__"):  # device-resident rows (e.g. a torch CUDA tensor): D2D, stream-ordered
            cai = desc.__cuda_array_interface__
            assert cai["typestr"] == "<f4" and len(cai["shape"]) == 2 and cai["shape"][1] == 128 and not cai.get("strides")
            dev_ptr, self.n = C.c_void_p(cai["data"][0]), int(cai["shape"][0])
            async_upload = True
        else:
            desc = _f32(desc)
            assert desc.ndim == 2 and desc.shape[1] == 128
            self.n = desc.shape[0]
        g = None
        if grid is not None:
            g = grid if isinstance(grid, Grid) else Grid(*grid)
        if kps is not None:
            kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
            assert len(kps) == self.n
        h = C.c_void_p()
        create = ctx.lib.asdm_set_create_async if async_upload else ctx.lib.asdm_set_create
        self._desc_keepalive = desc if async_upload else None
        ctx.check(create(ctx.h, dev_ptr if dev_ptr is not None else ptr(desc), self.n, 0, ptr(kps) if kps is not None else None,
                         C.byref(g) if g is not None else None, C.byref(h)))
        self.h = h
        if featvec is not None:
            self.set_featvec(featvec)

    def set_featvec(self, fv):
        node_ids, off, idx = fv
        node_ids = np.ascontiguousarray(node_ids, dtype=np.uint32)
        off, idx = _i32(off), _i32(idx)
        c = FeatVec(len(node_ids), ptr(node_ids), ptr(off), ptr(idx))
        self.ctx.check(self.ctx.lib.asdm_set_featvec(self.h, C.byref(c)))

    def read_desc(self, row0=0, n_rows=None):
        n_rows = self.n - row0 if n_rows is None else n_rows
        out = np.empty((n_rows, 128), np.float32)
        self.ctx.check(self.ctx.lib.asdm_set_read_desc(self.h, row0, n_rows, ptr(out)))
        return out

    @classmethod
    def create_batch(cls, ctx, descs, kps=None, grid=None, featvecs=None):
        """asdm_set_create_batch: the sets of a batch of keyframes in one call (asynchronous uploads: every `descs[k]`
        must stay alive and unmodified until the next synchronising call; page-locked memory gives a true async DMA).
        kps: list of KP_DTYPE arrays, featvecs: list of (node_ids, off, idx) -- or None."""
        m = len(descs)
        g = None
        if grid is not None:
            g = grid if isinstance(grid, Grid) else Grid(*grid)
        arr = (SetDesc * m)()
        keep = []
        for k in range(m):
            d = descs[k]
            a = arr[k]
            if hasattr(d, "__cuda_array_interface__"):  # rows already on this GPU (e.g. a torch CUDA tensor): device-to-device
                cai = d.__cuda_array_interface__
                assert cai["typestr"] == "<f4" and len(cai["shape"]) == 2 and cai["shape"][1] == 128 and not cai.get("strides")
                a.desc, a.n, a.row_stride_bytes = cai["data"][0], int(cai["shape"][0]), 0
            else:
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
        hs = (C.c_void_p * m)()
        ctx.check(ctx.lib.asdm_set_create_batch(ctx.h, arr, m, hs))
        out = []
        for k in range(m):
            self = cls.__new__(cls)
            self.ctx, self.h, self.n, self._desc_keepalive = ctx, C.c_void_p(hs[k]), int(arr[k].n), descs[k]
            out.append(self)
        return out

    @classmethod
    def from_map_block(cls, ctx, block, kps=None, grid=None):
        """Set from the descriptor block of a chamo.map frame (visual_map_seri.cc:272-285). Returns (set, bytes used).
        `block`: uint8 array starting at the block; it must stay alive until the next synchronising call."""
        self = cls.__new__(cls)
        self.ctx = ctx
        block = np.ascontiguousarray(block, np.uint8)
        g = None
        if grid is not None:
            g = grid if isinstance(grid, Grid) else Grid(*grid)
        if kps is not None:
            kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
        h, used = C.c_void_p(), C.c_size_t()
        ctx.check(ctx.lib.asdm_set_load_block(ctx.h, ptr(block), block.nbytes, ptr(kps) if kps is not None else None,
                                              C.byref(g) if g is not None else None, C.byref(h), C.byref(used)))
        self.h, self._desc_keepalive = h, block
        self.n = ctx.lib.asdm_set_size(h)
        return self, used.value

    def retain(self):
        self.ctx.lib.asdm_set_retain(self.h)

    def release(self):
        if self.h:
            self.ctx.lib.asdm_set_release(self.h)
            self.h = None


class PairJob:
    """Batched keyframe-pair search whose inputs are resident in HBM (asdm_job_*): the loop bodies of
    LoopClosing.cc:291-318 (kind BOW_KF_KF) and LocalMapping.cc:335-370 (kind TRIANGULATION)."""

    def __init__(self, ctx, kind, sets, masks, ia, ib, nnratio=0.6, check_ori=True, F12=None, epipoles=None,
                 scale_factors=None, level_sigma2=None, row_begin=None, row_end=None):
        self.ctx = ctx
        self.n_pairs = len(ia)
        n_kf = len(sets)
        self._keep = dict(masks=[_u8(m) for m in masks], ia=_i32(ia), ib=_i32(ib))
        k = self._keep
        set_arr = (C.c_void_p * n_kf)(*[s.h for s in sets])
        mask_arr = (C.c_void_p * n_kf)(*[m.ctypes.data for m in k["masks"]])
        d = JobDesc()
        d.kind, d.n_kf, d.sets, d.masks = kind, n_kf, set_arr, mask_arr
        d.n_pairs, d.ia, d.ib = self.n_pairs, ptr(k["ia"]), ptr(k["ib"])
        d.nnratio, d.check_ori = float(nnratio), int(bool(check_ori))
        if kind == ASDM_JOB_TRIANGULATION:
            k["F12"], k["ep"] = _f32(F12).reshape(-1), _f32(epipoles).reshape(-1)
            k["sf"], k["ls"] = _f32(scale_factors), _f32(level_sigma2)
            d.F12, d.epipoles, d.scale_factors, d.level_sigma2 = ptr(k["F12"]), ptr(k["ep"]), ptr(k["sf"]), ptr(k["ls"])
            d.n_levels = len(k["sf"])
        if row_begin is not None:  # row-block units of a pair (triangulation without rotation check only)
            k["rb"], k["re"] = _i32(row_begin), _i32(row_end)
            d.row_begin, d.row_end = ptr(k["rb"]), ptr(k["re"])
        h = C.c_void_p()
        ctx.check(ctx.lib.asdm_job_prepare(ctx.h, C.byref(d), C.byref(h)))
        self.h = h
        self.stride = ctx.lib.asdm_job_out_stride(h)

    def run(self):
        self.ctx.check(self.ctx.lib.asdm_job_run(self.h))

    def settle(self):
        self.ctx.check(self.ctx.lib.asdm_job_settle(self.h))

    def fetch(self, match=None, dist=None, nm=None):
        P, S = self.n_pairs, self.stride
        match = np.empty((P, S), np.int32) if match is None else match
        dist = np.empty((P, S), np.float32) if dist is None else dist
        nm = np.empty(P, np.int32) if nm is None else nm
        self.ctx.check(self.ctx.lib.asdm_job_fetch(self.h, ptr(match), ptr(dist), ptr(nm)))
        return nm, match, dist

    def fetch_async(self, match, dist, nm):
        """Queues the result copies behind the run; `fetch` on the same arrays later validates (and finishes) them."""
        self.ctx.check(self.ctx.lib.asdm_job_fetch_async(self.h, ptr(match), ptr(dist), ptr(nm)))

    def device_outputs(self):
        m, d, n = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self.ctx.check(self.ctx.lib.asdm_job_device_outputs(self.h, C.byref(m), C.byref(d), C.byref(n)))
        return m.value, d.value, n.value

    def set_publish(self, match_ptrs, dist_ptrs, nm_ptrs, row0, stride, multicast=False, dist=True):
        """The job's resolve kernel also stores every finished row into these rank-major gathered arrays (device
        pointers: this GPU's and its peers', or one NVSwitch multicast mapping of them all -- see
        shard.GatheredResults); an empty list switches it off. dist=False publishes index lists and counts only (the
        reference's calls return no distances); the distance rows then stay in the job's local outputs."""
        n = len(match_ptrs)
        arr = lambda ps: (C.c_void_p * max(n, 1))(*ps)
        self.ctx.check(self.ctx.lib.asdm_job_set_publish(self.h, n, arr(match_ptrs), arr(dist_ptrs), arr(nm_ptrs), row0, stride,
                                                         (1 if multicast else 0) | (0 if dist else 2)))

    def destroy(self):
        if self.h:
            self.ctx.lib.asdm_job_destroy(self.h)
            self.h = None


class StreamedPairSweep:
    """Keyframe-pair sweep over keyframes that are NOT yet resident (a cold loop-closure sweep, or a map being loaded):
    keyframes are uploaded in index order and pairs are searched in readiness order, so the search of the pairs whose
    keyframes have arrived overlaps the PCIe upload of the later ones. Pairs are dealt into waves by max(ia, ib):
    wave w holds the pairs whose later keyframe lies in [bounds[w-1], bounds[w]) and is launched as soon as those
    keyframes are queued (asdm_job_prepare never waits for uploads in flight). Results come back in readiness order:
    row r of the outputs belongs to pair `order[r]` (the caller indexes, nothing is permuted on the host)."""

    def __init__(self, ctx, kind, n_kf, ia, ib, nnratio=0.6, check_ori=True, bounds=None, **job_kw):
        self.ctx, self.kind, self.n_kf = ctx, kind, int(n_kf)
        self.ia, self.ib = _i32(ia), _i32(ib)
        self.nnratio, self.check_ori, self.job_kw = nnratio, check_ori, job_kw
        if bounds is None:
            # waves shrink towards the end: the last one (after the last upload) is all that is not hidden.
            # pairs with max < b are a (b/n)^2 share of a uniform sweep
            # (measured on C5, B200 with tools/stream_timeline.py: 4 to 7 waves all land within 5 %; every wave costs
            # ~0.1 ms of fixed kernel tails, so more waves stop paying once the searches, not the uploads, are the long pole)
            bounds = sorted({max(2, int(round(self.n_kf * f))) for f in (0.375, 0.5625, 0.72, 0.845, 0.94, 1.0)})
        self.bounds = [int(b) for b in bounds]
        assert self.bounds[-1] == self.n_kf
        later = np.maximum(self.ia, self.ib)
        self.order = np.argsort(later, kind="stable").astype(np.int64)
        cuts = np.searchsorted(later[self.order], self.bounds, side="left")
        self.slices = [(int(a), int(b)) for a, b in zip(np.concatenate(([0], cuts[:-1])), cuts)]
        self._per_pair = {k: (np.asarray(v).reshape(len(self.ia), -1) if k in ("F12", "epipoles", "row_begin", "row_end") else v)
                          for k, v in job_kw.items()}

    def run(self, make_sets, masks, match, dist, nm):
        """make_sets(lo, hi) -> list of DescriptorSets of keyframes lo..hi-1 with their FeatureVectors attached, uploaded
        asynchronously (DescriptorSet.create_batch). match / dist: (n_pairs, stride) arrays, nm: (n_pairs,), page-locked.
        Returns the sets (the caller releases them)."""
        sets, jobs, done = [], [], 0
        for (a, b), bound in zip(self.slices, self.bounds):
            if bound > done:
                sets.extend(make_sets(done, bound))
                done = bound
            if b <= a:
                continue
            sel = self.order[a:b]
            kw = {k: (v[sel] if k in ("F12", "epipoles", "row_begin", "row_end") else v) for k, v in self._per_pair.items()}
            jb = PairJob(self.ctx, self.kind, sets[:bound], masks[:bound], self.ia[sel], self.ib[sel], self.nnratio,
                         self.check_ori, **kw)
            jb.run()
            jb.fetch_async(match[a:b], dist[a:b], nm[a:b])  # the copies of this wave overlap the next wave's kernels
            jobs.append((jb, a, b))
        for jb, a, b in jobs:  # validates each wave; copies again only if a wave had to be rerun
            jb.fetch(match[a:b], dist[a:b], nm[a:b])
            jb.destroy()
        return sets


class ORBmatcher:
    """ORB_SLAM2::ORBmatcher(nnratio=0.6, checkOri=true) (reference ORBmatcher.h:45)."""

    TH_LOW = TH_LOW
    TH_HIGH = TH_HIGH

    def __init__(self, ctx, nnratio=0.6, checkOri=True):
        self.ctx = ctx
        self.mfNNratio = float(nnratio)
        self.mbCheckOrientation = bool(checkOri)

    # -- SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches)            ORBmatcher.cc:156-297
    def SearchByBoW_KF_F(self, kf, valid_kf, f):
        valid_kf = _u8(valid_kf)
        match = np.empty(f.n, np.int32)
        dist = np.empty(f.n, np.float32)
        nm = C.c_int32()
        c = self.ctx
        c.check(c.lib.asdm_search_by_bow_kf_f(c.h, kf.h, ptr(valid_kf), f.h, self.mfNNratio,
                                              int(self.mbCheckOrientation), ptr(match), ptr(dist), C.byref(nm)))
        return nm.value, match, dist

    # -- SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)               ORBmatcher.cc:533-666
    def SearchByBoW_KF_KF(self, kf1, valid1, kf2, valid2):
        valid1, valid2 = _u8(valid1), _u8(valid2)
        match = np.empty(kf1.n, np.int32)
        dist = np.empty(kf1.n, np.float32)
        nm = C.c_int32()
        c = self.ctx
        c.check(c.lib.asdm_search_by_bow_kf_kf(c.h, kf1.h, ptr(valid1), kf2.h, ptr(valid2), self.mfNNratio,
                                               int(self.mbCheckOrientation), ptr(match), ptr(dist), C.byref(nm)))
        return nm.value, match, dist

    # -- SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo)   ORBmatcher.cc:669-822
    def SearchForTriangulation(self, kf1, has_mp1, kf2, has_mp2, F12, ex, ey, scale_factors2, level_sigma2_2):
        has_mp1, has_mp2 = _u8(has_mp1), _u8(has_mp2)
        F12, sf, ls = _f32(F12).reshape(9), _f32(scale_factors2), _f32(level_sigma2_2)
        match = np.empty(kf1.n, np.int32)
        dist = np.empty(kf1.n, np.float32)
        nm = C.c_int32()
        c = self.ctx
        c.check(c.lib.asdm_search_for_triangulation(c.h, kf1.h, ptr(has_mp1), kf2.h, ptr(has_mp2), ptr(F12),
                                                    float(ex), float(ey), ptr(sf), ptr(ls), len(sf),
                                                    int(self.mbCheckOrientation), ptr(match), ptr(dist),
                                                    C.byref(nm)))
        return nm.value, match, dist

    # -- the descriptor loop shared by the four SearchByProjection overloads
    def SearchByProjection(self, target, queries, desc_q=None, q_set=None, q_rows=None, blocked0=None,
                           mode=ASDM_PROJ_BEST_ONLY, th_accept=TH_HIGH, best_init=256.0, count_twice=False,
                           check_ori=None, chi2_max=0.0, inv_level_sigma2=None):
        queries = np.ascontiguousarray(queries, dtype=PROJ_QUERY_DTYPE)
        nq = len(queries)
        dq = _f32(desc_q) if desc_q is not None else None
        rows = _i32(q_rows) if q_rows is not None else None
        b0 = _u8(blocked0) if blocked0 is not None else None
        co = self.mbCheckOrientation if check_ori is None else check_ori
        p = ProjParams(int(mode), float(th_accept), float(best_init), self.mfNNratio, int(bool(co)),
                       int(bool(count_twice)))
        if chi2_max > 0:  # Fuse's reprojection gate (ORBmatcher.cc:919-926)
            inv = _f32(inv_level_sigma2)
            p.chi2_max, p.n_levels = float(chi2_max), len(inv)
            for l, v in enumerate(inv):
                p.inv_level_sigma2[l] = float(v)
        assign = np.empty(target.n, np.int32)
        dist = np.empty(target.n, np.float32)
        nm = C.c_int32()
        c = self.ctx
        c.check(c.lib.asdm_search_by_projection(c.h, target.h, ptr(queries), nq, ptr(dq),
                                                q_set.h if q_set is not None else None, ptr(rows), ptr(b0),
                                                C.byref(p), ptr(assign), ptr(dist), C.byref(nm)))
        return nm.value, assign, dist

    # -- per-query view of the same search: the match phase of Fuse (ORBmatcher.cc:825-1086) and each direction of
    #    SearchBySim3 (:1090-1314), where several queries may choose one keypoint
    def SearchByProjection_PerQuery(self, target, queries, desc_q, blocked0=None, th_accept=TH_LOW, best_init=256.0,
                                    chi2_max=0.0, inv_level_sigma2=None):
        queries = np.ascontiguousarray(queries, dtype=PROJ_QUERY_DTYPE)
        dq = _f32(desc_q)
        b0 = _u8(blocked0) if blocked0 is not None else None
        p = ProjParams(ASDM_PROJ_BEST_ONLY, float(th_accept), float(best_init), self.mfNNratio, 0, 0)
        if chi2_max > 0:
            inv = _f32(inv_level_sigma2)
            p.chi2_max, p.n_levels = float(chi2_max), len(inv)
            for l, v in enumerate(inv):
                p.inv_level_sigma2[l] = float(v)
        best = np.empty(len(queries), np.int32)
        dist = np.empty(len(queries), np.float32)
        nm = C.c_int32()
        c = self.ctx
        c.check(c.lib.asdm_search_by_projection_q(c.h, target.h, ptr(queries), len(queries), ptr(dq), None, None, ptr(b0),
                                                  C.byref(p), ptr(best), ptr(dist), C.byref(nm)))
        return best, dist

    # -- SearchByProjection(Frame&, vector<MapPoint*>&, th)           ORBmatcher.cc:44-122
    @staticmethod
    def RadiusByViewingCos(view_cos):  # ORBmatcher.cc:126-132
        return np.where(np.asarray(view_cos, np.float32).astype(np.float64) > 0.998, np.float32(2.5), np.float32(4.0))

    def SearchByProjection_LocalMap(self, F, track_in_view, proj_x, proj_y, scale_level, view_cos, blocks, desc_mp,
                                    scale_factors, th=1.0, blocked0=None):
        """F: DescriptorSet of the frame; per map point the MapPoint::mTrack* fields (MapPoint.h:95-100)."""
        n = len(proj_x)
        q = np.zeros(n, PROJ_QUERY_DTYPE)
        lvl = _i32(scale_level)
        r = self.RadiusByViewingCos(view_cos).astype(np.float32)
        if th != 1.0:
            r = (r * np.float32(th)).astype(np.float32)
        q["valid"] = _u8(track_in_view)
        q["blocks"] = _u8(blocks)
        q["u"], q["v"] = _f32(proj_x), _f32(proj_y)
        q["r"] = (r * _f32(scale_factors)[lvl]).astype(np.float32)
        q["min_level"], q["max_level"] = lvl - 1, lvl
        return self.SearchByProjection(F, q, desc_q=desc_mp, blocked0=blocked0, mode=ASDM_PROJ_RATIO_LEVEL,
                                       th_accept=TH_HIGH, best_init=256.0, count_twice=True, check_ori=False)

    # -- SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)   ORBmatcher.cc:416-531
    def SearchForInitialization(self, F1, F2, vbPrevMatched, windowSize=10):
        prev = np.array(vbPrevMatched, dtype=np.float32, copy=True).reshape(-1)
        match = np.empty(F1.n, np.int32)
        dist = np.empty(F1.n, np.float32)
        nm = C.c_int32()
        c = self.ctx
        c.check(c.lib.asdm_search_for_initialization(c.h, F1.h, F2.h, ptr(prev), int(windowSize), self.mfNNratio,
                                                     int(self.mbCheckOrientation), ptr(match), ptr(dist),
                                                     C.byref(nm)))
        return nm.value, match, dist, prev.reshape(-1, 2)

    # -- descriptor stage of Frame::ComputeStereoMatches              Frame.cc:360-444
    def StereoRowMatch(self, left, right, scale_factors, n_rows, min_d, max_d):
        sf = _f32(scale_factors)
        best = np.empty(left.n, np.int32)
        dist = np.empty(left.n, np.float32)
        c = self.ctx
        c.check(c.lib.asdm_stereo_row_match(c.h, left.h, right.h, ptr(sf), len(sf), int(n_rows), float(min_d),
                                            float(max_d), ptr(best), ptr(dist)))
        return best, dist


def ComputeDistinctiveDescriptors(ctx, sets, obs_off, obs_set, obs_row):
    """MapPoint::ComputeDistinctiveDescriptors (reference MapPoint.cc:271-338) for a batch of map points.

    sets: DescriptorSets of the keyframes; point p is observed at rows (sets[obs_set[e]], obs_row[e]) for e in
    [obs_off[p], obs_off[p+1]), in the reference's iteration order.  Returns (best, median): position within each
    point's list of the descriptor that becomes MapPoint::mDescriptor (-1: no observation) and its median distance."""
    off, st, row = _i32(obs_off), _i32(obs_set), _i32(obs_row)
    n = len(off) - 1
    best = np.empty(n, np.int32)
    med = np.empty(n, np.float32)
    arr = (C.c_void_p * max(len(sets), 1))(*[s.h for s in sets])
    ctx.check(ctx.lib.asdm_distinctive_descriptors(ctx.h, arr, len(sets), n, ptr(off), ptr(st), ptr(row), ptr(best), ptr(med)))
    return best, med


def save_map_blocks(ctx, sets):
    """Descriptor blocks of `sets`, back to back, as vm::save_visual_map writes them (visual_map_seri.cc:134-146).
    Returns (uint8 array, offsets)."""
    total = sum(ctx.lib.asdm_desc_block_bytes(s.n) for s in sets)
    out = np.empty(total, np.uint8)
    offs = np.zeros(len(sets) + 1, np.uint64)
    hs = (C.c_void_p * len(sets))(*[s.h for s in sets])
    ctx.check(ctx.lib.asdm_set_save_blocks(ctx.h, hs, len(sets), ptr(out), total, ptr(offs)))
    return out, offs.astype(np.int64)


# DBoW2 enums (src/dbow2/include/BowVector.h:36-53)
TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3
L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = 0, 1, 2, 3, 4, 5


class ORBVocabulary:
    """DBoW2::TemplatedVocabulary<FSift::TDescriptor, FSift> (reference src/vslam/include/ORBVocabulary.h:34) resident
    on the device. `tree` holds the flat arrays of include/asdmatch.h asdm_vocab_desc: k, L, weighting, scoring,
    child_off, child_idx, desc (n_nodes x 128), weight (float64), word_id."""

    def __init__(self, ctx, tree):
        self.ctx = ctx
        self._keep = (_i32(tree["child_off"]), _i32(tree["child_idx"]), _f32(tree["desc"]),
                      np.ascontiguousarray(tree["weight"], np.float64), _i32(tree["word_id"]))
        off, idx, desc, w, word = self._keep
        d = VocabDesc(int(tree["k"]), int(tree["L"]), int(tree["weighting"]), int(tree["scoring"]), len(w), ptr(off),
                      ptr(idx), ptr(desc), ptr(w), ptr(word))
        h = C.c_void_p()
        ctx.check(ctx.lib.asdm_vocab_create(ctx.h, C.byref(d), C.byref(h)))
        self.h = h

    def transform(self, dset, levelsup=4, host=True):
        """transform(features, BowVector&, FeatureVector&, levelsup) (TemplatedVocabulary.h:1125-1192) over the rows of
        `dset`; attaches the FeatureVector to the set on the device. Returns dict(word_id, node_id, in_fv, bow_id,
        bow_val), or None with host=False (nothing comes back, the call stays asynchronous)."""
        ctx = self.ctx
        if not host:
            ctx.check(ctx.lib.asdm_vocab_transform(ctx.h, self.h, dset.h, int(levelsup), None, None, None, None, None, None))
            return None
        n = dset.n
        word = np.empty(n, np.int32)
        node = np.empty(n, np.int32)
        inc = np.empty(n, np.uint8)
        bid = np.empty(max(n, 1), np.uint32)
        bval = np.empty(max(n, 1), np.float64)
        nb = C.c_int32()
        ctx.check(ctx.lib.asdm_vocab_transform(ctx.h, self.h, dset.h, int(levelsup), ptr(word), ptr(node), ptr(inc),
                                               C.byref(nb), ptr(bid), ptr(bval)))
        return dict(word_id=word, node_id=node, in_fv=inc, bow_id=bid[:nb.value].copy(), bow_val=bval[:nb.value].copy())

    def release(self):
        if self.h:
            self.ctx.lib.asdm_vocab_release(self.h)
            self.h = None


class KeyFrameDatabase:
    """ORB_SLAM2::KeyFrameDatabase (reference src/vslam/include/KeyFrameDatabase.h:40-62, KeyFrameDatabase.cc:34-320) with
    the BowVectors of all keyframes resident on the device. Keyframes are named by slot = the value add() returns (add
    order = their order inside the reference's inverted lists). BowVectors are (ascending word ids, values) arrays."""

    def __init__(self, ctx):
        self.ctx = ctx
        h = C.c_void_p()
        ctx.check(ctx.lib.asdm_bowdb_create(ctx.h, C.byref(h)))
        self.h = h

    def __len__(self):
        return int(self.ctx.lib.asdm_bowdb_size(self.h))

    @staticmethod
    def _bow(word, val):
        return np.ascontiguousarray(word, np.uint32), np.ascontiguousarray(val, np.float64)

    def add(self, word, val):
        """KeyFrameDatabase::add(pKF) (:41-47); returns the keyframe's slot."""
        w, v = self._bow(word, val)
        slot = C.c_int32(-1)
        self.ctx.check(self.ctx.lib.asdm_bowdb_add(self.h, ptr(w), ptr(v), len(w), C.byref(slot)))
        return slot.value

    def erase(self, slot):
        """KeyFrameDatabase::erase(pKF) (:49-70)."""
        self.ctx.check(self.ctx.lib.asdm_bowdb_erase(self.h, int(slot)))

    def clear(self):
        """KeyFrameDatabase::clear() (:72-76)."""
        self.ctx.check(self.ctx.lib.asdm_bowdb_clear(self.h))

    def query(self, word, val):
        """The device pass alone: (common words, smallest common word, L1 score as float) for every slot."""
        w, v = self._bow(word, val)
        n = len(self)
        common, first, score = np.zeros(n, np.int32), np.zeros(n, np.uint32), np.zeros(n, np.float32)
        self.ctx.check(self.ctx.lib.asdm_bowdb_query(self.h, ptr(w), ptr(v), len(w), ptr(common), ptr(first), ptr(score)))
        return common, first, score

    def _detect(self, fn, word, val, connected, global_flag, neigh_off, neigh_slot, min_score, only_global_map):
        w, v = self._bow(word, val)
        n = len(self)
        keep = [None if a is None else np.ascontiguousarray(a, dt) for a, dt in
                ((connected, np.uint8), (global_flag, np.uint8), (neigh_off, np.int32), (neigh_slot, np.int32))]
        for a, want in zip(keep[:3], (n, n, n + 1)):
            if a is not None and len(a) != want:
                raise ValueError("gate arrays must cover every slot of the database")
        g = BowdbGates(ptr(keep[0]), ptr(keep[1]), ptr(keep[2]), ptr(keep[3]), float(min_score), int(bool(only_global_map)))
        out = np.empty(max(n, 1), np.int32)
        k = C.c_int32(0)
        self.ctx.check(fn(self.h, ptr(w), ptr(v), len(w), C.byref(g), ptr(out), n, C.byref(k)))
        return out[:k.value].copy()

    def DetectLoopCandidates(self, word, val, connected, min_score, only_global_map=False, global_flag=None,
                             neigh_off=None, neigh_slot=None):
        """DetectLoopCandidates(pKF, minScore, only_global_map) (:80-216): word/val = pKF->mBowVec, connected[slot] =
        membership in pKF->GetConnectedKeyFrames(), neigh_* = CSR of GetBestCovisibilityKeyFrames(10) per slot.
        Returns the candidate slots in the reference's order."""
        return self._detect(self.ctx.lib.asdm_bowdb_detect_loop, word, val, connected, global_flag, neigh_off, neigh_slot,
                            min_score, only_global_map)

    def DetectRelocalizationCandidates(self, word, val, only_global_map=False, global_flag=None, neigh_off=None,
                                       neigh_slot=None):
        """DetectRelocalizationCandidates(F, only_global_map) (:219-320); word/val = F->mBowVec."""
        return self._detect(self.ctx.lib.asdm_bowdb_detect_reloc, word, val, None, global_flag, neigh_off, neigh_slot, 0.0,
                            only_global_map)

    def reloc_scores(self):
        """KeyFrame::mRelocScore of every slot as the queries so far left it."""
        s = np.zeros(len(self), np.float32)
        self.ctx.check(self.ctx.lib.asdm_bowdb_reloc_scores(self.h, ptr(s)))
        return s

    def release(self):
        if self.h:
            self.ctx.lib.asdm_bowdb_release(self.h)
            self.h = None
