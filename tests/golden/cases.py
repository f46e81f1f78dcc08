"""The golden cases: small seeded inputs for every entry point of the hot path.  Shared by make_golden.py (which runs
the reference's own ORBmatcher.cc on them, in the container that has /root/reference), by the CPU test that checks the
oracle against the stored outputs, and by the GPU test that checks libasdmatch_cuda.so against them.

Each case is (name, kind, inputs-dict).  Inputs are regenerated from seeds; their SHA-256 is stored beside the
outputs so that a drifting generator is noticed rather than silently comparing different problems."""
import hashlib
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def synth():
    return importlib.import_module("asd-slam_b200").synth


def digest(obj):
    h = hashlib.sha256()

    def walk(o):
        if isinstance(o, dict):
            for k in sorted(o):
                h.update(str(k).encode())
                walk(o[k])
        elif isinstance(o, (list, tuple)):
            for x in o:
                walk(x)
        elif isinstance(o, np.ndarray):
            h.update(str(o.dtype).encode() + str(o.shape).encode())
            h.update(np.ascontiguousarray(o).tobytes())
        else:
            h.update(repr(o).encode())
    walk(obj)
    return h.hexdigest()


def with_duplicates(c, seed):
    """Exact duplicate descriptors -> exactly equal distances -> the scan-order tie rule decides."""
    rng = np.random.default_rng(seed)
    m = len(c["desc_b"])
    db = c["desc_b"].copy()
    db[rng.integers(0, m, m // 6)] = db[rng.integers(0, m, m // 6)]
    da = c["desc_a"].copy()
    da[rng.integers(0, len(da), len(da) // 10)] = db[rng.integers(0, m, len(da) // 10)]
    c = dict(c)
    c["desc_a"], c["desc_b"] = da, db
    rngv = np.random.default_rng(seed + 1)
    c["valid_a"] = (rngv.random(len(da)) < 0.85).astype(np.uint8)
    c["valid_b"] = (rngv.random(m) < 0.85).astype(np.uint8)
    return c


def cases():
    S = synth()
    out = []
    # BoW-gated dense searches (a8, a9)
    for name, n, m, nodes, ratio, ori, dup in [("bow_dense", 500, 600, 1, 0.7, 1, False), ("bow_nodes", 400, 350, 11, 0.75, 1, False),
                                               ("bow_ties", 300, 300, 3, 0.9, 1, True), ("bow_noori", 256, 128, 1, 0.6, 0, True),
                                               ("bow_tiny", 1, 3, 1, 0.7, 1, False)]:
        c = S.config1(n, m, n_nodes=nodes, seed_extra=len(out))
        c = with_duplicates(c, 100 + len(out)) if dup else dict(c, valid_a=np.ones(n, np.uint8), valid_b=np.ones(m, np.uint8))
        out.append((name, "bow", dict(c=c, ratio=ratio, ori=ori)))
    # triangulation (a11)
    for name, n, pairs, nodes, ori in [("tri_dense", 400, 3, 1, 0), ("tri_nodes", 300, 2, 7, 1)]:
        out.append((name, "tri", dict(c=S.config4(n=n, n_pairs=pairs, n_nodes=nodes, seed_extra=len(out)), ori=ori)))
    # initialization (a10)
    for name, n, window in [("init_100", 600, 100), ("init_40", 250, 40)]:
        out.append((name, "init", dict(c=S.config_init(n=n, window=window, seed_extra=len(out)), window=window)))
    # projection-gated (a4, a5, a6, a7)
    for name, th in [("proj_localmap_th1", 1.0), ("proj_localmap_th5", 5.0)]:
        out.append((name, "proj_localmap", dict(c=S.config2(n_kp=500, n_mp=2000, n_in_view=700, th=th, seed_extra=len(out)), th=th)))
    out.append(("proj_f2f", "proj_f2f", dict(c=S.config3(n=600, seed_extra=len(out))["f2f"], th=15.0)))
    for name, kind, th, acc in [("proj_reloc", "proj_reloc", 10.0, 0.7), ("proj_sim3", "proj_sim3", 10.0, 0.5)]:
        f = S.config3(n=500, seed_extra=len(out))["f2f"]
        q = f["queries"].copy()
        rng = np.random.default_rng(len(out))
        level = rng.integers(0, 8, len(q)).astype(np.int32)
        q["r"] = (np.float32(th) * S.scale_factors()[level]).astype(np.float32)
        q["min_level"] = level - 1
        q["max_level"] = level + 1 if kind == "proj_reloc" else level
        q["blocks"] = 1
        if kind == "proj_sim3":
            W, H = S.EUROC["width"], S.EUROC["height"]
            q["valid"] &= ((q["u"] >= 0) & (q["u"] < W) & (q["v"] >= 0) & (q["v"] < H)).astype(np.uint8)
        blocked0 = (rng.random(len(f["kps_t"])) < 0.1).astype(np.uint8)
        out.append((name, kind, dict(f=f, q=q, level=level, blocked0=blocked0, th=th, acc=acc)))
    # Fuse x2 and SearchBySim3: candidates project near keypoints that hold noisy copies of them
    W, H = S.EUROC["width"], S.EUROC["height"]
    for name, kind, th in [("fuse", "fuse", 3.0), ("fuse_sim3", "fuse_sim3", 4.0)]:
        rng = np.random.default_rng(700 + len(out))
        n = 500
        desc_q = S.make_descriptors(rng, n)
        kq = S.make_keypoints(rng, n, W, H)
        desc_t, kps_t, gt = S.make_matched_set(rng, desc_q, kq, n, W, H, match_frac=0.7, d_lo=0.05, d_hi=0.8, pos_sigma=2.0)
        q = np.zeros(n, S.PROJ_QUERY_DTYPE)
        q["valid"] = (rng.random(n) < 0.9).astype(np.uint8)
        q["u"] = np.clip(kq["x"] + rng.normal(0, 1.0, n), 0, W - 1e-2).astype(np.float32)
        q["v"] = np.clip(kq["y"] + rng.normal(0, 1.0, n), 0, H - 1e-2).astype(np.float32)
        level = np.minimum(kq["octave"] + rng.integers(0, 2, n), 7).astype(np.int32)
        occ_obs = np.where(rng.random(n) < 0.5, rng.integers(1, 6, n), 0).astype(np.int32)
        mp_obs = rng.integers(1, 6, n).astype(np.int32)
        # two candidates aimed at the same keypoint: the second sees what the first one changed
        q["u"][1], q["v"][1], level[1] = q["u"][0], q["v"][0], level[0]
        desc_q[1] = desc_q[0]
        out.append((name, kind, dict(desc_t=desc_t, kps_t=kps_t, grid=S.image_grid(W, H), q=q, level=level, desc_q=desc_q,
                                     occ_obs=occ_obs, mp_obs=mp_obs, th=th)))
    rng = np.random.default_rng(900)
    n = 400
    d1 = S.make_descriptors(rng, n)
    k1 = S.make_keypoints(rng, n, W, H)
    d2, k2, gt = S.make_matched_set(rng, d1, k1, n, W, H, match_frac=0.7, d_lo=0.05, d_hi=1.2, pos_sigma=3.0)
    inv = np.full(n, -1, np.int64)
    inv[gt[gt >= 0]] = np.nonzero(gt >= 0)[0]
    uv1 = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)  # where KF1's map points land in KF2
    has = inv >= 0
    uv1[has, 0], uv1[has, 1] = k2["x"][inv[has]], k2["y"][inv[has]]
    uv2 = np.stack([k2["x"], k2["y"]], 1).astype(np.float32)
    has2 = gt >= 0
    uv2[has2, 0], uv2[has2, 1] = k1["x"][gt[has2]], k1["y"][gt[has2]]
    uv1 += rng.normal(0, 1.0, uv1.shape).astype(np.float32)
    uv2 += rng.normal(0, 1.0, uv2.shape).astype(np.float32)
    uv1 = np.clip(uv1, 0, [W - 1e-2, H - 1e-2]).astype(np.float32)
    uv2 = np.clip(uv2, 0, [W - 1e-2, H - 1e-2]).astype(np.float32)
    out.append(("sim3", "sim3", dict(desc1=d1, kps1=k1, desc2=d2, kps2=k2, grid=S.image_grid(W, H),
                                     valid1=(rng.random(n) < 0.85).astype(np.uint8), valid2=(rng.random(n) < 0.85).astype(np.uint8),
                                     uv1=uv1, uv2=uv2, level1=np.minimum(k1["octave"] + rng.integers(0, 2, n), 7).astype(np.int32),
                                     level2=np.minimum(k2["octave"] + rng.integers(0, 2, n), 7).astype(np.int32),
                                     already1=(rng.random(n) < 0.1).astype(np.uint8), th=7.5)))
    # MapPoint::ComputeDistinctiveDescriptors: few observations per point, and points seen from > 32 keyframes
    out.append(("distinct", "distinct", dict(c=S.config_distinctive(n_kf=10, n=200, n_points=300, seed_extra=1))))
    out.append(("distinct_many_obs", "distinct", dict(c=S.config_distinctive(n_kf=44, n=60, n_points=50, seed_extra=2, dup_frac=0.1))))
    # ---- round 2 (appended: the cases above keep their seeds) -----------------------------------------------------
    # Descriptor stage of Frame::ComputeStereoMatches (Frame.cc:366-447, a14 / BASELINE config 3). The reference indexes
    # its row table without a bounds check, so right keypoints stay 8 px off the top and bottom (its extractor keeps 19).
    # stereo_a: clean matches, many best distances in [0.5, 1.0) -- accepted only because thOrbDist is int((1.5+0.5)/2) = 1;
    # stereo_b: noisier right descriptors, best distances on both sides of 1.0 and of TH_HIGH.
    for name, n, noise in [("stereo_a", 700, 0.0), ("stereo_b", 500, 0.06)]:
        s = dict(S.config3(n=n, seed_extra=40 + len(out))["stereo"])
        kr = s["kps_r"].copy()
        kr["y"] = np.clip(kr["y"], 8.0, s["n_rows"] - 9.0).astype(np.float32)
        s["kps_r"] = kr
        if noise:
            rng = np.random.default_rng(41 + len(out))
            s["desc_r"] = (s["desc_r"] + rng.normal(0, noise, s["desc_r"].shape)).astype(np.float32)
        out.append((name, "stereo", dict(s=s)))
    # One keyframe pair of each batched workload at FULL size, run through the reference itself (15-40 s each at
    # generation time): C5 = SearchByBoW(KF, KF) 4000 x 4000 (0.85, rotation check), C4 = SearchForTriangulation 2000 x 2000.
    c5 = S.config5(n=4000, n_pairs=1, n_kf=2, seed_extra=50)
    out.append(("c5_full_pair", "bow_kfkf_full", dict(a=c5["kfs"][0], b=c5["kfs"][1], ratio=0.85, ori=1)))
    c4 = S.config4(n=2000, n_pairs=1, seed_extra=51)
    out.append(("c4_full_pair", "tri", dict(c=c4, ori=0)))
    # The pose-dependent overloads once more with REAL poses (oracle/ref_driver.cc `Scene`, mode 1): EuRoC intrinsics,
    # random SE3 / Sim3, ~6 % of the points behind the camera, ~15 % outside the distance-invariance band, ~15 % of the
    # normals beyond the 60 degree gate. These exercise the host-side gates of the adaptor (Rcw*x3Dw+tcw, PredictScale,
    # decompose_sim3 ...: ORBmatcher.cc:331-368, 852-888, 1459-1510); the flat C ABI and the oracle take windows, not
    # poses, so only tests/test_adaptor.py runs them (reference build for the golden file, adaptor build on the GPU).
    base = {n: (k, a) for n, k, a in out}
    for n in ("proj_f2f", "proj_reloc", "proj_sim3", "fuse", "fuse_sim3", "sim3"):
        k, a = base[n]
        out.append((n + "_posed", k, dict(a, scene=7 + len(out))))
    return out


def first_equal_observation(c, chosen):
    """Position, within each point's observation list, of the first observation whose descriptor equals `chosen[p]`
    bit for bit (-1 for points without observations: the reference leaves their descriptor untouched)."""
    off, okf, orow = c["obs_off"], c["obs_kf"], c["obs_row"]
    out = np.full(len(off) - 1, -1, np.int32)
    for p in range(len(off) - 1):
        for e in range(off[p], off[p + 1]):
            if np.array_equal(c["descs"][okf[e]][orow[e]].view(np.uint32), chosen[p].view(np.uint32)):
                out[p] = e - off[p]
                break
        assert out[p] >= 0 or off[p] == off[p + 1], "the reference chose a descriptor that is not among the observations"
    return out


def chosen_descriptors(c, best):
    """Rows picked by `best` (positions within each point's list); NaN rows where best < 0."""
    out = np.full((len(best), 128), np.nan, np.float32)
    for p, b in enumerate(best):
        if b >= 0:
            e = c["obs_off"][p] + b
            out[p] = c["descs"][c["obs_kf"][e]][c["obs_row"][e]]
    return out


def pair_dists(ref, da, db, ia, ib):
    return np.array([ref.descriptor_distance(da[i], db[j]) for i, j in zip(ia, ib)], np.float32)


def run_case(ref, name, kind, a, sha=None):
    """Runs one case through the flat ref_* harness API (oracle/ref_py.py) -- the reference build for the golden file,
    or the C++ adaptor build under test -- and returns {name/key: array}."""
    if "scene" in a:  # real poses (see cases()): the harness keeps the scene in a global, reset whatever happens
        ref.set_scene(1, a["scene"])
        try:
            return run_case(ref, name, kind, {k: v for k, v in a.items() if k != "scene"}, sha=digest(a))
        finally:
            ref.set_scene(0, 0)
    S = synth()
    out = {}
    out[name + "/sha"] = np.frombuffer(bytes.fromhex(sha or digest(a)), np.uint8)
    if kind == "bow":
        c = a["c"]
        nm, m = ref.search_by_bow_kf_f(c["desc_a"], c["kps_a"], c["fv_a"], c["valid_a"], c["desc_b"], c["kps_b"], c["fv_b"], a["ratio"], a["ori"])
        j = np.nonzero(m >= 0)[0]
        out[name + "/kf_f_nm"], out[name + "/kf_f_match"] = np.int32(nm), m
        out[name + "/kf_f_dist"] = pair_dists(ref, c["desc_a"], c["desc_b"], m[j], j)
        nm, m = ref.search_by_bow_kf_kf(c["desc_a"], c["kps_a"], c["fv_a"], c["valid_a"], c["desc_b"], c["kps_b"], c["fv_b"], c["valid_b"], a["ratio"], a["ori"])
        i = np.nonzero(m >= 0)[0]
        out[name + "/kf_kf_nm"], out[name + "/kf_kf_match"] = np.int32(nm), m
        out[name + "/kf_kf_dist"] = pair_dists(ref, c["desc_a"], c["desc_b"], i, m[i])
    elif kind == "tri":
        c = a["c"]
        kfs = c["kfs"]
        ms, nms = [], []
        for p in range(len(c["ia"])):
            x, y = kfs[c["ia"][p]], kfs[c["ib"][p]]
            nm, m = ref.search_for_triangulation(x["desc"], x["kps"], x["fv"], x["has_mp"], y["desc"], y["kps"], y["fv"], y["has_mp"],
                                                 c["F12"][p], c["epipoles"][p][0], c["epipoles"][p][1], c["scale_factors"], c["level_sigma2"], a["ori"])
            ms.append(m); nms.append(nm)
        out[name + "/match"], out[name + "/nm"] = np.stack(ms), np.array(nms, np.int32)
    elif kind == "init":
        c = a["c"]
        nm, m, prev = ref.search_for_initialization(c["desc1"], c["kps1"], c["desc2"], c["kps2"], c["grid2"], c["prev_matched"], a["window"], c["nnratio"], 1)
        out[name + "/nm"], out[name + "/match"], out[name + "/prev"] = np.int32(nm), m, prev
    elif kind == "proj_localmap":
        c = a["c"]
        nm, asg = ref.sbp_local_map(c["desc_t"], c["kps_t"], c["grid"], c["queries"], c["level"], c["view_cos"], c["desc_q"], c["blocked0"], S.scale_factors(), a["th"], 0.8)
        out[name + "/nm"], out[name + "/assign"] = np.int32(nm), asg
    elif kind == "proj_f2f":
        f = a["c"]
        q = f["queries"]
        nm, asg = ref.sbp_frame_frame(f["desc_t"], f["kps_t"], f["grid"], q, (q["min_level"] + 1).astype(np.int32), f["desc_q"], f["blocked0"], S.scale_factors(), a["th"], 0.8, 1)
        out[name + "/nm"], out[name + "/assign"] = np.int32(nm), asg
    elif kind == "proj_reloc":
        f = a["f"]
        nm, asg = ref.sbp_frame_keyframe(f["desc_t"], f["kps_t"], f["grid"], a["q"], a["level"], f["desc_q"], a["blocked0"], S.scale_factors(), a["th"], a["acc"], 0.9, 1)
        out[name + "/nm"], out[name + "/assign"] = np.int32(nm), asg
    elif kind == "proj_sim3":
        f = a["f"]
        nm, asg = ref.sbp_keyframe_sim3(f["desc_t"], f["kps_t"], f["grid"], a["q"], a["level"], f["desc_q"], a["blocked0"], S.scale_factors(), int(a["th"]), 0.85)
        out[name + "/nm"], out[name + "/assign"] = np.int32(nm), asg
    elif kind == "fuse":
        nf, act = ref.fuse(a["desc_t"], a["kps_t"], a["grid"], a["q"], a["level"], a["desc_q"], a["occ_obs"], a["mp_obs"], S.scale_factors(), a["th"])
        out[name + "/nf"], out[name + "/action"] = np.int32(nf), act
    elif kind == "fuse_sim3":
        nf, rep, add = ref.fuse_sim3(a["desc_t"], a["kps_t"], a["grid"], a["q"], a["level"], a["desc_q"], a["occ_obs"], a["mp_obs"], S.scale_factors(), a["th"])
        out[name + "/nf"], out[name + "/replace"], out[name + "/added"] = np.int32(nf), rep, add
    elif kind == "distinct":
        c = a["c"]
        out[name + "/best"] = first_equal_observation(c, ref.distinctive(c["descs"], c["obs_off"], c["obs_kf"], c["obs_row"]))
    elif kind == "stereo":
        s = a["s"]
        best, dist = ref.stereo_row_match(s["desc_l"], s["kps_l"], s["desc_r"], s["kps_r"], s["scale_factors"], s["n_rows"], s["max_d"])
        out[name + "/best"], out[name + "/dist"] = best, dist
    elif kind == "bow_kfkf_full":
        x, y = a["a"], a["b"]
        nm, m = ref.search_by_bow_kf_kf(x["desc"], x["kps"], x["fv"], x["valid"], y["desc"], y["kps"], y["fv"], y["valid"], a["ratio"], a["ori"])
        i = np.nonzero(m >= 0)[0]
        out[name + "/nm"], out[name + "/match"] = np.int32(nm), m
        out[name + "/dist"] = pair_dists(ref, x["desc"], y["desc"], i, m[i])
    elif kind == "sim3":
        nf, m = ref.search_by_sim3(a["desc1"], a["kps1"], a["desc2"], a["kps2"], a["grid"], a["valid1"], a["valid2"], a["uv1"], a["uv2"],
                                   a["level1"], a["level2"], a["desc1"], a["desc2"], a["already1"], S.scale_factors(), a["th"])
        out[name + "/nf"], out[name + "/match"] = np.int32(nf), m
    return out
