"""Seeded synthetic inputs for the five BASELINE configs.

Descriptors are i.i.d. N(0,1)^128, L2-normalised in fp32 (ASD outputs are signed and unit-norm:
reference ASDNet/ASDNet/ASDNet.py:354-370).  A "matched" second set is a permuted 60 % subset with
additive noise giving true-match squared distances in [0.05, 0.45] plus 40 % fresh outliers.
Keypoints are uniform in the image, octaves follow the extractor's geometric per-level budget
(reference src/vslam/src/ORBextractor.cc:481-492; scale 1.2, 8 levels) and matched pairs share a
global rotation offset + N(0, 5 deg) so the rotation histogram prunes something.
"""
import numpy as np

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("angle", "<f4"), ("octave", "<i4")])
PROJ_QUERY_DTYPE = np.dtype(
    [("valid", "u1"), ("blocks", "u1"), ("pad", "<i2"), ("u", "<f4"), ("v", "<f4"), ("r", "<f4"),
     ("min_level", "<i4"), ("max_level", "<i4"), ("angle", "<f4")])

DESC_DIM = 128
N_LEVELS = 8
SCALE = np.float32(1.2)
GRID_COLS, GRID_ROWS = 64, 48  # reference src/vslam/include/Frame.h:37-38
SEED_BASE = 0xA5D00000

KITTI = dict(width=1241, height=376, fx=718.856, fy=718.856, cx=607.1928, cy=185.2157)
EUROC = dict(width=752, height=480, fx=458.654, fy=457.296, cx=367.215, cy=248.375)


def scale_factors():
    sf = np.ones(N_LEVELS, np.float32)
    for i in range(1, N_LEVELS):
        sf[i] = np.float32(sf[i - 1] * SCALE)
    return sf


def level_sigma2():
    sf = scale_factors()
    return (sf * sf).astype(np.float32)


def image_grid(width, height):
    """(min_x, min_y, max_x, max_y, inv_w, inv_h, cols, rows) as in reference Frame.cc:100-112."""
    min_x, min_y, max_x, max_y = np.float32(0), np.float32(0), np.float32(width), np.float32(height)
    inv_w = np.float32(np.float32(GRID_COLS) / (max_x - min_x))
    inv_h = np.float32(np.float32(GRID_ROWS) / (max_y - min_y))
    return (float(min_x), float(min_y), float(max_x), float(max_y), float(inv_w), float(inv_h), GRID_COLS, GRID_ROWS)


def make_descriptors(rng, n):
    d = rng.standard_normal((n, DESC_DIM)).astype(np.float32)
    d /= np.sqrt((d * d).sum(axis=1, keepdims=True, dtype=np.float32))
    return np.ascontiguousarray(d, dtype=np.float32)


def octave_probs():
    f = 1.0 / float(SCALE)
    p = np.array([f ** l for l in range(N_LEVELS)])
    return p / p.sum()


def make_keypoints(rng, n, width, height):
    kps = np.zeros(n, KP_DTYPE)
    kps["x"] = rng.uniform(0, width, n).astype(np.float32)
    kps["y"] = rng.uniform(0, height, n).astype(np.float32)
    kps["angle"] = rng.uniform(0, 360, n).astype(np.float32)
    kps["octave"] = rng.choice(N_LEVELS, size=n, p=octave_probs()).astype(np.int32)
    return kps


def dense_fv(n):
    """FeatureVector of the `L <= levelsup` regime: every feature in node 0 (SURVEY headline)."""
    return (np.zeros(1, np.uint32), np.array([0, n], np.int32), np.arange(n, dtype=np.int32))


def fv_from_nodes(node_of):
    """CSR FeatureVector from a per-feature node id (<0 = feature not in the vector, i.e. stopped word)."""
    node_of = np.asarray(node_of)
    ids = np.unique(node_of[node_of >= 0])
    off = [0]
    idx = []
    for nid in ids:
        members = np.nonzero(node_of == nid)[0]
        idx.append(members)
        off.append(off[-1] + len(members))
    idx = np.concatenate(idx) if idx else np.zeros(0, np.int64)
    return (ids.astype(np.uint32), np.array(off, np.int32), idx.astype(np.int32))


def make_matched_set(rng, desc_a, kps_a, m, width, height, match_frac=0.6, d_lo=0.05, d_hi=0.45, rot_offset=20.0,
                     rot_sigma=5.0, pos_sigma=None, rot_outlier_frac=0.08):
    """Second descriptor/keypoint set related to (desc_a, kps_a).  Returns desc_b, kps_b, gt (index into A or -1)."""
    n = len(desc_a)
    n_match = min(int(round(match_frac * min(n, m))), n, m)
    src = rng.permutation(n)[:n_match]
    slots = rng.permutation(m)
    desc_b = make_descriptors(rng, m)
    kps_b = make_keypoints(rng, m, width, height)
    gt = np.full(m, -1, np.int64)
    target_d = rng.uniform(d_lo, d_hi, n_match)
    sigma = np.sqrt(target_d / DESC_DIM).astype(np.float32)
    noisy = desc_a[src] + rng.standard_normal((n_match, DESC_DIM)).astype(np.float32) * sigma[:, None]
    noisy /= np.sqrt((noisy * noisy).sum(axis=1, keepdims=True, dtype=np.float32))
    dst = slots[:n_match]
    desc_b[dst] = noisy.astype(np.float32)
    gt[dst] = src
    ang = kps_a["angle"][src] - np.float32(rot_offset) + rng.normal(0, rot_sigma, n_match).astype(np.float32)
    wild = rng.random(n_match) < rot_outlier_frac  # matches whose rotation disagrees: pruned by the histogram
    ang = np.where(wild, rng.uniform(0, 360, n_match), ang)
    kps_b["angle"][dst] = np.mod(ang, 360.0).astype(np.float32)
    kps_b["octave"][dst] = kps_a["octave"][src]
    if pos_sigma is not None:
        kps_b["x"][dst] = np.clip(kps_a["x"][src] + rng.normal(0, pos_sigma, n_match), 0, width - 1e-3).astype(np.float32)
        kps_b["y"][dst] = np.clip(kps_a["y"][src] + rng.normal(0, pos_sigma, n_match), 0, height - 1e-3).astype(np.float32)
    return np.ascontiguousarray(desc_b, np.float32), kps_b, gt


def rng_for(config, extra=0):
    return np.random.default_rng(SEED_BASE + config * 1000 + extra)


# ----------------------------------------------------------------------------------------- configs
def config1(n=2000, m=2000, seed_extra=0, n_nodes=1):
    """C1: single frame pair, SearchByBoW(KF, F) semantics with a one-node FeatureVector."""
    rng = rng_for(1, seed_extra)
    W, H = KITTI["width"], KITTI["height"]
    desc_a = make_descriptors(rng, n)
    kps_a = make_keypoints(rng, n, W, H)
    desc_b, kps_b, gt = make_matched_set(rng, desc_a, kps_a, m, W, H)
    if n_nodes == 1:
        fv_a, fv_b = dense_fv(n), dense_fv(m)
    else:
        node_a = rng.integers(0, n_nodes, n)
        node_b = rng.integers(0, n_nodes, m)
        has = gt >= 0
        same = rng.random(m) < 0.9  # most true matches land in the same vocabulary node
        node_b[has & same] = node_a[gt[has & same]]
        node_a[rng.random(n) < 0.01] = -1  # stopped words
        node_b[rng.random(m) < 0.01] = -1
        fv_a, fv_b = fv_from_nodes(node_a * 7 + 3), fv_from_nodes(np.where(node_b >= 0, node_b * 7 + 3, -1))
        fv_a = fv_from_nodes(np.where(node_a >= 0, node_a * 7 + 3, -1))
    valid_a = np.ones(n, np.uint8)
    valid_b = np.ones(m, np.uint8)
    return dict(desc_a=desc_a, kps_a=kps_a, fv_a=fv_a, valid_a=valid_a, desc_b=desc_b, kps_b=kps_b, fv_b=fv_b,
                valid_b=valid_b, gt=gt, nnratio=0.7, check_ori=1, grid=image_grid(W, H))


def config2(n_kp=2000, n_mp=8000, n_in_view=2000, th=1.0, seed_extra=0):
    """C2: SearchByProjection(Frame, local map points, th) -- 8000 map points vs 2000 keypoints, KITTI geometry."""
    rng = rng_for(2, seed_extra)
    W, H = KITTI["width"], KITTI["height"]
    sf = scale_factors()
    desc_f = make_descriptors(rng, n_kp)
    kps_f = make_keypoints(rng, n_kp, W, H)
    desc_q = make_descriptors(rng, n_mp)
    q = np.zeros(n_mp, PROJ_QUERY_DTYPE)
    level = np.zeros(n_mp, np.int32)
    view_cos = np.where(rng.random(n_mp) < 0.5, 0.999, 0.9).astype(np.float32)
    in_view = rng.permutation(n_mp)[:n_in_view]
    src = rng.permutation(n_kp)[:n_in_view] if n_in_view <= n_kp else rng.integers(0, n_kp, n_in_view)
    q["valid"][in_view] = 1
    q["blocks"] = 1
    q["blocks"][rng.random(n_mp) < 0.02] = 0  # a few map points without observations
    q["u"] = rng.uniform(0, W, n_mp).astype(np.float32)
    q["v"] = rng.uniform(0, H, n_mp).astype(np.float32)
    q["u"][in_view] = (kps_f["x"][src] + rng.normal(0, 1.5, n_in_view)).astype(np.float32)
    q["v"][in_view] = (kps_f["y"][src] + rng.normal(0, 1.5, n_in_view)).astype(np.float32)
    level[in_view] = np.minimum(kps_f["octave"][src] + rng.integers(0, 2, n_in_view), N_LEVELS - 1)
    # descriptor of an in-view map point = noisy copy of the keypoint it projects onto (80 % of them)
    good = rng.random(n_in_view) < 0.8
    tgt_d = rng.uniform(0.05, 0.9, n_in_view)
    sigma = np.sqrt(tgt_d / DESC_DIM).astype(np.float32)
    noisy = desc_f[src] + rng.standard_normal((n_in_view, DESC_DIM)).astype(np.float32) * sigma[:, None]
    noisy /= np.sqrt((noisy * noisy).sum(axis=1, keepdims=True, dtype=np.float32))
    desc_q[in_view[good]] = noisy[good].astype(np.float32)
    # window: r = RadiusByViewingCos(cos) [* th] * scaleFactor[level]   (reference ORBmatcher.cc:62-68)
    base = np.where(view_cos.astype(np.float64) > 0.998, np.float32(2.5), np.float32(4.0)).astype(np.float32)
    if th != 1.0:
        base = (base * np.float32(th)).astype(np.float32)
    q["r"] = (base * sf[level]).astype(np.float32)
    q["min_level"] = level - 1
    q["max_level"] = level
    blocked0 = (rng.random(n_kp) < 0.05).astype(np.uint8)
    return dict(desc_t=desc_f, kps_t=kps_f, grid=image_grid(W, H), queries=q, desc_q=np.ascontiguousarray(desc_q),
                blocked0=blocked0, mode=0, th_accept=1.5, best_init=256.0, nnratio=0.8, check_ori=0, count_twice=1,
                level=level, view_cos=view_cos, th=th)


def config3(n=1500, seed_extra=0, th=15.0):
    """C3: EuRoC-shaped 752x480. (a) left/right row-band stereo descriptor stage; (b) frame-to-frame
    SearchByProjection(Cur, Last, th=15, mono)."""
    rng = rng_for(3, seed_extra)
    W, H = EUROC["width"], EUROC["height"]
    sf = scale_factors()
    desc_l = make_descriptors(rng, n)
    kps_l = make_keypoints(rng, n, W, H)
    # right image: same rows (+-1 px), disparity in [0, 60]
    desc_r, kps_r, gt_r = make_matched_set(rng, desc_l, kps_l, n, W, H, match_frac=0.7, d_lo=0.05, d_hi=1.2)
    has = gt_r >= 0
    disp = rng.uniform(0, 60, has.sum())
    kps_r["x"][has] = np.maximum(kps_l["x"][gt_r[has]] - disp, 0).astype(np.float32)
    kps_r["y"][has] = np.clip(kps_l["y"][gt_r[has]] + rng.normal(0, 0.7, has.sum()), 0, H - 1e-3).astype(np.float32)
    bf, fx = 47.9, EUROC["fx"]
    mb = np.float32(bf / fx)
    stereo = dict(desc_l=desc_l, kps_l=kps_l, desc_r=desc_r, kps_r=kps_r, scale_factors=sf, n_rows=H,
                  min_d=0.0, max_d=float(np.float32(bf) / mb))
    # frame-to-frame: "last" frame map points project near current keypoints
    desc_c, kps_c, gt_c = make_matched_set(rng, desc_l, kps_l, n, W, H, match_frac=0.7, d_lo=0.05, d_hi=1.4,
                                           pos_sigma=4.0)
    q = np.zeros(n, PROJ_QUERY_DTYPE)
    q["valid"] = (rng.random(n) < 0.6).astype(np.uint8)  # keypoints of the last frame that hold a map point
    q["blocks"] = 1
    q["blocks"][rng.random(n) < 0.03] = 0
    q["u"] = (kps_l["x"] + rng.normal(0, 3.0, n)).astype(np.float32)
    q["v"] = (kps_l["y"] + rng.normal(0, 3.0, n)).astype(np.float32)
    oct_l = kps_l["octave"]
    q["r"] = (np.float32(th) * sf[oct_l]).astype(np.float32)
    q["min_level"] = oct_l - 1
    q["max_level"] = oct_l + 1
    q["angle"] = kps_l["angle"]
    inb = (q["u"] >= 0) & (q["u"] <= W) & (q["v"] >= 0) & (q["v"] <= H)
    q["valid"] &= inb.astype(np.uint8)
    blocked0 = np.zeros(n, np.uint8)
    f2f = dict(desc_t=desc_c, kps_t=kps_c, grid=image_grid(W, H), queries=q, desc_q=desc_l, blocked0=blocked0,
               mode=1, th_accept=1.5, best_init=100.0, nnratio=0.8, check_ori=1, count_twice=0)
    return dict(stereo=stereo, f2f=f2f)


def _random_F12(rng):
    """Fundamental matrix of a synthetic relative pose with KITTI intrinsics, plus the epipole in image 2."""
    K = np.array([[KITTI["fx"], 0, KITTI["cx"]], [0, KITTI["fy"], KITTI["cy"]], [0, 0, 1.0]])
    ang = rng.normal(0, 0.03, 3)
    Rx = np.array([[1, 0, 0], [0, np.cos(ang[0]), -np.sin(ang[0])], [0, np.sin(ang[0]), np.cos(ang[0])]])
    Ry = np.array([[np.cos(ang[1]), 0, np.sin(ang[1])], [0, 1, 0], [-np.sin(ang[1]), 0, np.cos(ang[1])]])
    Rz = np.array([[np.cos(ang[2]), -np.sin(ang[2]), 0], [np.sin(ang[2]), np.cos(ang[2]), 0], [0, 0, 1]])
    R12 = Rz @ Ry @ Rx
    t12 = np.array([rng.normal(0, 0.1), rng.normal(0, 0.05), rng.uniform(0.5, 1.5)])
    tx = np.array([[0, -t12[2], t12[1]], [t12[2], 0, -t12[0]], [-t12[1], t12[0], 0]])
    Kinv = np.linalg.inv(K)
    F12 = Kinv.T @ tx @ R12 @ Kinv  # reference LocalMapping.cc ComputeF12
    # camera-1 centre seen from camera 2: C2 = R21 * 0 + t21 = -R12^T t12
    C2 = -R12.T @ t12
    ex = KITTI["fx"] * C2[0] / C2[2] + KITTI["cx"]
    ey = KITTI["fy"] * C2[1] / C2[2] + KITTI["cy"]
    return F12.astype(np.float32), np.float32(ex), np.float32(ey), R12, t12, K


def make_keyframes(rng, n_kf, n, width, height, base_frac=0.6, n_nodes=1, mask_prob=0.5):
    """n_kf keyframes observing a common pool of landmarks: each keyframe holds noisy copies of a random
    60 % of the pool plus outliers, so every pair of keyframes shares true matches."""
    pool = make_descriptors(rng, n)
    pool_angle = rng.uniform(0, 360, n).astype(np.float32)
    pool_oct = rng.choice(N_LEVELS, size=n, p=octave_probs()).astype(np.int32)
    pool_node = rng.integers(0, max(n_nodes, 1), n)
    kfs = []
    for k in range(n_kf):
        desc = make_descriptors(rng, n)
        kps = make_keypoints(rng, n, width, height)
        n_match = int(base_frac * n)
        src = rng.permutation(n)[:n_match]
        dst = rng.permutation(n)[:n_match]
        tgt = rng.uniform(0.02, 0.2, n_match)  # pairwise distance is about twice this
        sigma = np.sqrt(tgt / DESC_DIM).astype(np.float32)
        noisy = pool[src] + rng.standard_normal((n_match, DESC_DIM)).astype(np.float32) * sigma[:, None]
        noisy /= np.sqrt((noisy * noisy).sum(axis=1, keepdims=True, dtype=np.float32))
        desc[dst] = noisy.astype(np.float32)
        rot = np.float32(rng.uniform(0, 360))
        ang = pool_angle[src] + rot + rng.normal(0, 5.0, n_match)
        ang = np.where(rng.random(n_match) < 0.08, rng.uniform(0, 360, n_match), ang)
        kps["angle"][dst] = np.mod(ang, 360.0).astype(np.float32)
        kps["octave"][dst] = pool_oct[src]
        landmark = np.full(n, -1, np.int64)
        landmark[dst] = src
        if n_nodes <= 1:
            fv = dense_fv(n)
        else:
            node = rng.integers(0, n_nodes, n)
            node[dst] = pool_node[src]
            fv = fv_from_nodes(node)
        mask = (rng.random(n) < mask_prob).astype(np.uint8)
        kfs.append(dict(desc=np.ascontiguousarray(desc, np.float32), kps=kps, fv=fv, mask=mask, landmark=landmark))
    return kfs


def config4(n=2000, n_pairs=20, seed_extra=0, n_nodes=1):
    """C4: SearchForTriangulation between a new keyframe (index 0) and its 20 covisible neighbours."""
    rng = rng_for(4, seed_extra)
    W, H = KITTI["width"], KITTI["height"]
    kfs = make_keyframes(rng, n_pairs + 1, n, W, H, n_nodes=n_nodes, mask_prob=0.5)
    for kf in kfs:
        kf["has_mp"] = kf["mask"]
    ia = np.zeros(n_pairs, np.int32)
    ib = np.arange(1, n_pairs + 1, dtype=np.int32)
    F = np.zeros((n_pairs, 9), np.float32)
    ep = np.zeros((n_pairs, 2), np.float32)
    for p in range(n_pairs):
        F12, ex, ey, R12, t12, K = _random_F12(rng)
        F[p] = F12.reshape(9)
        ep[p] = (ex, ey)
        # move the true matches of KF2 onto the epipolar line of their KF1 partner so the gate passes
        A, B = kfs[0], kfs[ib[p]]
        lm_to_a = {int(l): i for i, l in enumerate(A["landmark"]) if l >= 0}
        for j, l in enumerate(B["landmark"]):
            i = lm_to_a.get(int(l), -1)
            if i < 0:
                continue
            x1 = np.array([A["kps"]["x"][i], A["kps"]["y"][i], 1.0])
            line = x1 @ F12.astype(np.float64)  # l = x1' F12
            a, b, c = line
            if abs(b) < 1e-9:
                continue
            x2 = rng.uniform(0, W)
            y2 = -(a * x2 + c) / b + rng.normal(0, 0.5)
            if 0 <= y2 < H:
                B["kps"]["x"][j] = np.float32(x2)
                B["kps"]["y"][j] = np.float32(y2)
    return dict(kfs=kfs, ia=ia, ib=ib, F12=F, epipoles=ep, scale_factors=scale_factors(),
                level_sigma2=level_sigma2(), check_ori=0, grid=image_grid(W, H))


def config5(n=4000, n_pairs=512, n_kf=64, seed_extra=0, n_nodes=1):
    """C5: loop-closure verification sweep, SearchByBoW(KF, KF) over candidate pairs; 64 distinct keyframes."""
    rng = rng_for(5, seed_extra)
    W, H = KITTI["width"], KITTI["height"]
    kfs = make_keyframes(rng, n_kf, n, W, H, n_nodes=n_nodes, mask_prob=0.9)
    for kf in kfs:
        kf["valid"] = kf["mask"]
    ia = rng.integers(0, n_kf, n_pairs).astype(np.int32)
    ib = ((ia + 1 + rng.integers(0, n_kf - 1, n_pairs)) % n_kf).astype(np.int32)
    return dict(kfs=kfs, ia=ia, ib=ib, nnratio=0.85, check_ori=1, grid=image_grid(W, H))


def config_init(n=4000, seed_extra=0, window=100):
    """SearchForInitialization (a10): two 4000-keypoint frames (Tracking.cc:85), window 100 px."""
    rng = rng_for(6, seed_extra)
    W, H = KITTI["width"], KITTI["height"]
    desc1 = make_descriptors(rng, n)
    kps1 = make_keypoints(rng, n, W, H)
    desc2, kps2, gt = make_matched_set(rng, desc1, kps1, n, W, H, pos_sigma=20.0)
    prev = np.stack([kps1["x"], kps1["y"]], axis=1).astype(np.float32)  # Tracking.cc:404-406
    return dict(desc1=desc1, kps1=kps1, desc2=desc2, kps2=kps2, grid2=image_grid(W, H), prev_matched=prev,
                window=window, nnratio=0.9, check_ori=1, gt=gt)


def config_distinctive(n_kf=12, n=600, n_points=900, seed_extra=0, max_obs=None, dup_frac=0.05):
    """MapPoint::ComputeDistinctiveDescriptors inputs: `n_points` map points, each observed in 1..max_obs
    keyframes (a few in none), observation descriptors = a per-point base descriptor + noise; some observations are
    exact duplicates so that medians tie.  Observations are listed per point in ascending keyframe order, one per
    keyframe (the reference iterates a std::map<KeyFrame*, size_t>)."""
    rng = rng_for(7, seed_extra)
    max_obs = n_kf if max_obs is None else min(max_obs, n_kf)
    descs = [make_descriptors(rng, n) for _ in range(n_kf)]
    next_row = np.zeros(n_kf, np.int64)
    off, okf, orow = [0], [], []
    for p in range(n_points):
        r = rng.random()
        k = 0 if r < 0.02 else (1 if r < 0.07 else int(rng.integers(2, max_obs + 1)))
        cand = [kf for kf in range(n_kf) if next_row[kf] < n]
        kfs = sorted(rng.choice(cand, size=min(k, len(cand)), replace=False).tolist()) if cand and k else []
        base = make_descriptors(rng, 1)[0]
        prev = None
        for kf in kfs:
            row = int(next_row[kf])
            next_row[kf] += 1
            if prev is not None and rng.random() < dup_frac:
                d = prev.copy()
            else:
                d = base + rng.normal(0, rng.uniform(0.01, 0.08), DESC_DIM).astype(np.float32)
                d = (d / np.sqrt((d * d).sum(dtype=np.float32))).astype(np.float32)
            descs[kf][row] = d
            prev = d
            okf.append(kf)
            orow.append(row)
        off.append(len(okf))
    return dict(descs=descs, obs_off=np.asarray(off, np.int32), obs_kf=np.asarray(okf, np.int32),
                obs_row=np.asarray(orow, np.int32))


def vocab_tree(k=10, L=5, seed_extra=0, ragged=0.0, stop_frac=0.05, weighting=0, scoring=0, dup_children=False):
    """Synthetic DBoW2 vocabulary (flat arrays of include/asdmatch.h asdm_vocab_desc) shaped like the output of
    TemplatedVocabulary::create (TemplatedVocabulary.h:557-600): a k-ary tree of depth L whose node descriptors are
    unit-norm perturbations of their parent's, node ids in creation (depth-first) order, word ids over the leaves in
    ascending node id (createWords, :813-835). ragged: probability that an inner node above level L stays a leaf or
    loses some children (k-means leaves clusters with one feature unsplit, :700-716); stop_frac: share of words with
    weight 0 (idf of a word seen in every image); dup_children: some siblings share one descriptor (exact ties)."""
    rng = rng_for(9, seed_extra)
    parent, level, desc = [0], [0], [np.zeros(DESC_DIM, np.float32)]
    children = [[]]
    stack = [0]
    while stack:
        p = stack.pop()
        if level[p] >= L:
            continue
        kp = k
        if p != 0 and ragged > 0 and rng.random() < ragged:
            kp = int(rng.integers(0, k))  # 0 = stays a leaf
        first = None
        for c in range(kp):
            nid = len(parent)
            base = desc[p] if p else np.zeros(DESC_DIM, np.float32)
            d = base + rng.standard_normal(DESC_DIM).astype(np.float32) * np.float32(0.6 ** level[p])
            d = (d / np.sqrt((d * d).sum(dtype=np.float32))).astype(np.float32)
            if dup_children and first is not None and rng.random() < 0.2:
                d = first.copy()
            if first is None:
                first = d
            parent.append(p); level.append(level[p] + 1); desc.append(d); children.append([])
            children[p].append(nid)
        stack.extend(reversed(children[p]))
    nn = len(parent)
    off = np.zeros(nn + 1, np.int32)
    for p in range(nn):
        off[p + 1] = off[p] + len(children[p])
    idx = np.array([c for ch in children for c in ch], np.int32)
    leaves = np.nonzero(off[1:] == off[:-1])[0]
    word = np.zeros(nn, np.int32)
    word[leaves] = np.arange(len(leaves), dtype=np.int32)
    w = np.zeros(nn, np.float64)
    w[leaves] = rng.uniform(0.1, 5.0, len(leaves))
    w[leaves[rng.random(len(leaves)) < stop_frac]] = 0.0
    return dict(k=k, L=L, weighting=weighting, scoring=scoring, child_off=off, child_idx=idx,
                desc=np.ascontiguousarray(np.stack(desc), np.float32), weight=w, word_id=word)


def vocab_queries(tree, n, seed_extra=0, sigma=0.05):
    """Descriptors near random leaves of `tree` (so the descent has near-ties at every level), unit norm."""
    rng = rng_for(10, seed_extra)
    off = tree["child_off"]
    leaves = np.nonzero(off[1:] == off[:-1])[0]
    d = tree["desc"][rng.choice(leaves, n)] + rng.standard_normal((n, DESC_DIM)).astype(np.float32) * np.float32(sigma)
    d /= np.sqrt((d * d).sum(axis=1, keepdims=True, dtype=np.float32))
    return np.ascontiguousarray(d, np.float32)


def config_kfdb(n_kf=300, n_words=100000, words_per_kf=900, n_places=60, seed_extra=0, erase_frac=0.05, global_frac=0.7):
    """A keyframe database for KeyFrameDatabase::DetectLoopCandidates / DetectRelocalizationCandidates
    (KeyFrameDatabase.cc:80-320): a trajectory that visits `n_places` places (consecutive keyframes share a place's
    word distribution, the end of the trajectory revisits the start), BowVectors = ascending word ids with
    L1-normalised tf-idf-like doubles, GetBestCovisibilityKeyFrames(10) = the nearest keyframes along the trajectory,
    a few erased keyframes, a global-map flag, and queries (BowVector + connected set) taken near the revisit."""
    rng = rng_for(9, seed_extra)
    per_place = max(1, n_kf // n_places)
    place_words = [rng.choice(n_words, size=3 * words_per_kf, replace=False) for _ in range(n_places)]
    common_pool = rng.choice(n_words, size=words_per_kf, replace=False)  # stop-word-like: seen everywhere

    def bow(place):
        k = int(words_per_kf * rng.uniform(0.7, 1.0))
        w = np.concatenate([rng.choice(place_words[place], size=int(0.7 * k), replace=False),
                            rng.choice(common_pool, size=int(0.2 * k), replace=False),
                            rng.integers(0, n_words, size=k - int(0.7 * k) - int(0.2 * k))])
        w = np.unique(w).astype(np.uint32)
        v = rng.gamma(2.0, 1.0, size=len(w))
        v = v / v.sum()  # L1 norm (BowVector::normalize); the exact values are arbitrary doubles
        return w, v.astype(np.float64)

    place_of = np.minimum(np.arange(n_kf) // per_place, n_places - 1)
    place_of[-2 * per_place:] = place_of[:2 * per_place]  # the loop: the trajectory ends where it began
    bows = [bow(int(p)) for p in place_of]
    off = np.zeros(n_kf + 1, np.int32)
    off[1:] = np.cumsum([len(w) for w, _ in bows])
    word = np.concatenate([w for w, _ in bows]).astype(np.uint32)
    val = np.concatenate([v for _, v in bows]).astype(np.float64)
    alive = (rng.uniform(size=n_kf) >= erase_frac).astype(np.uint8)
    glob = (rng.uniform(size=n_kf) < global_frac).astype(np.uint8)
    neigh_off = np.zeros(n_kf + 1, np.int32)
    neigh = []
    for s in range(n_kf):
        cand = [s + d for d in (1, -1, 2, -2, 3, -3, 4, -4, 5, -5, 6, -6) if 0 <= s + d < n_kf]
        k = int(rng.integers(0, 11))
        neigh.extend(cand[:k])
        neigh_off[s + 1] = len(neigh)
    neigh = np.asarray(neigh, np.int32)
    queries = []
    for qi in range(6):
        src = n_kf - 1 - qi * 3  # a keyframe at the end of the trajectory looks for the beginning
        qw, qv = bow(int(place_of[src]))
        conn = np.zeros(n_kf, np.uint8)
        conn[max(0, src - 15):] = 1  # its covisibility neighbourhood: never a loop candidate
        queries.append(dict(word=qw, val=qv, connected=conn, min_score=float(rng.uniform(0.005, 0.05))))
    return dict(n_kf=n_kf, n_words=n_words, bow_off=off, bow_word=word, bow_val=val, alive=alive, global_flag=glob,
                neigh_off=neigh_off, neigh_slot=neigh, queries=queries)
