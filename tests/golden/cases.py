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
    return out
