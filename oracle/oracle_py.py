"""ctypes wrapper over oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module (see oracle/asd_oracle.h).  The product never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("angle", "<f4"), ("octave", "<i4")])
PROJ_QUERY_DTYPE = np.dtype(
    [("valid", "u1"), ("blocks", "u1"), ("pad", "<i2"), ("u", "<f4"), ("v", "<f4"), ("r", "<f4"),
     ("min_level", "<i4"), ("max_level", "<i4"), ("angle", "<f4")])
assert KP_DTYPE.itemsize == 16 and PROJ_QUERY_DTYPE.itemsize == 28


class Grid(C.Structure):
    _fields_ = [("min_x", C.c_float), ("min_y", C.c_float), ("max_x", C.c_float), ("max_y", C.c_float),
                ("inv_w", C.c_float), ("inv_h", C.c_float), ("cols", C.c_int32), ("rows", C.c_int32)]


class FeatVec(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("node_ids", C.c_void_p), ("off", C.c_void_p), ("idx", C.c_void_p)]


class ProjParams(C.Structure):
    _fields_ = [("mode", C.c_int32), ("th_accept", C.c_float), ("best_init", C.c_float), ("nnratio", C.c_float),
                ("check_ori", C.c_int32), ("count_twice", C.c_int32), ("chi2_max", C.c_double), ("n_levels", C.c_int32),
                ("reserved", C.c_int32), ("inv_level_sigma2", C.c_float * 32)]


class KeyFrame(C.Structure):
    _fields_ = [("desc", C.c_void_p), ("kps", C.c_void_p), ("fv", C.POINTER(FeatVec)), ("mask", C.c_void_p)]


def build(force=False):
    """Compile liboracle.so (and nothing else) with the committed Makefile."""
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < max(
            os.path.getmtime(os.path.join(_HERE, f)) for f in ("asd_oracle.cc", "asd_oracle.h")):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
        _lib.ora_descriptor_distance.restype = C.c_float
        _lib.ora_radius_by_viewing_cos.restype = C.c_float
        _lib.ora_radius_by_viewing_cos.argtypes = [C.c_float]
        _lib.ora_local_map_window.restype = C.c_float
        _lib.ora_local_map_window.argtypes = [C.c_float, C.c_float, C.c_float]
        _lib.ora_predict_scale.argtypes = [C.c_float, C.c_float, C.c_float, C.c_int]
        _lib.ora_rot_bin.argtypes = [C.c_float, C.c_float]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


class FV:
    """Keeps the numpy arrays of a CSR feature vector alive next to the C struct."""

    def __init__(self, node_ids, off, idx):
        self.node_ids = np.ascontiguousarray(node_ids, dtype=np.uint32)
        self.off = np.ascontiguousarray(off, dtype=np.int32)
        self.idx = np.ascontiguousarray(idx, dtype=np.int32)
        self.c = FeatVec(len(self.node_ids), _p(self.node_ids), _p(self.off), _p(self.idx))

    @staticmethod
    def of(fv):
        return fv if isinstance(fv, FV) else FV(*fv)


def grid_struct(g):
    return g if isinstance(g, Grid) else Grid(*g)


def descriptor_distance(a, b):
    a, b = _f32(a), _f32(b)
    return float(lib().ora_descriptor_distance(_p(a), _p(b)))


def compute_three_maxima(sizes):
    s = np.ascontiguousarray(sizes, dtype=np.int32)
    i1, i2, i3 = C.c_int32(-1), C.c_int32(-1), C.c_int32(-1)
    lib().ora_compute_three_maxima(_p(s), len(s), C.byref(i1), C.byref(i2), C.byref(i3))
    return i1.value, i2.value, i3.value


def rot_bin(a1, a2):
    return int(lib().ora_rot_bin(float(a1), float(a2)))


def assign_features_to_grid(kps, grid):
    n = len(kps)
    g = grid_struct(grid)
    nc = g.cols * g.rows
    cell_start = np.zeros(nc + 1, np.int32)
    cell_items = np.zeros(max(n, 1), np.int32)
    cell_of = np.zeros(max(n, 1), np.int32)
    lib().ora_assign_features_to_grid(_p(kps), n, C.byref(g), _p(cell_start), _p(cell_items), _p(cell_of))
    return cell_start, cell_items[:cell_start[nc]], cell_of[:n]


def get_features_in_area(kps, grid, cell_start, cell_items, x, y, r, min_level=-1, max_level=-1):
    g = grid_struct(grid)
    out = np.zeros(max(len(kps), 1), np.int32)
    lib().ora_get_features_in_area.argtypes = None
    n = lib().ora_get_features_in_area(_p(kps), len(kps), C.byref(g), _p(cell_start), _p(cell_items),
                                       C.c_float(x), C.c_float(y), C.c_float(r), int(min_level), int(max_level),
                                       _p(out))
    return out[:n].copy()


def search_by_bow_kf_f(desc_kf, kps_kf, fv_kf, valid_kf, desc_f, kps_f, fv_f, nnratio, check_ori):
    desc_kf, desc_f, valid_kf = _f32(desc_kf), _f32(desc_f), _u8(valid_kf)
    fa, fb = FV.of(fv_kf), FV.of(fv_f)
    n_f = len(kps_f)
    match = np.empty(n_f, np.int32)
    dist = np.empty(n_f, np.float32)
    nm = lib().ora_search_by_bow_kf_f(_p(desc_kf), _p(kps_kf), len(kps_kf), C.byref(fa.c), _p(valid_kf), _p(desc_f),
                                      _p(kps_f), n_f, C.byref(fb.c), C.c_float(nnratio), int(check_ori), _p(match),
                                      _p(dist))
    return nm, match, dist


def search_by_bow_kf_kf(desc1, kps1, fv1, valid1, desc2, kps2, fv2, valid2, nnratio, check_ori):
    desc1, desc2, valid1, valid2 = _f32(desc1), _f32(desc2), _u8(valid1), _u8(valid2)
    fa, fb = FV.of(fv1), FV.of(fv2)
    n1 = len(kps1)
    match = np.empty(n1, np.int32)
    dist = np.empty(n1, np.float32)
    nm = lib().ora_search_by_bow_kf_kf(_p(desc1), _p(kps1), n1, C.byref(fa.c), _p(valid1), _p(desc2), _p(kps2),
                                       len(kps2), C.byref(fb.c), _p(valid2), C.c_float(nnratio), int(check_ori),
                                       _p(match), _p(dist))
    return nm, match, dist


def search_for_triangulation(desc1, kps1, fv1, has_mp1, desc2, kps2, fv2, has_mp2, F12, ex, ey, scale_factors2,
                             level_sigma2_2, check_ori):
    desc1, desc2, has_mp1, has_mp2 = _f32(desc1), _f32(desc2), _u8(has_mp1), _u8(has_mp2)
    F12, sf, ls = _f32(F12).reshape(9), _f32(scale_factors2), _f32(level_sigma2_2)
    fa, fb = FV.of(fv1), FV.of(fv2)
    n1 = len(kps1)
    match = np.empty(n1, np.int32)
    dist = np.empty(n1, np.float32)
    nm = lib().ora_search_for_triangulation(_p(desc1), _p(kps1), n1, C.byref(fa.c), _p(has_mp1), _p(desc2), _p(kps2),
                                            len(kps2), C.byref(fb.c), _p(has_mp2), _p(F12), C.c_float(ex),
                                            C.c_float(ey), _p(sf), _p(ls), int(check_ori), _p(match), _p(dist))
    return nm, match, dist


def search_by_projection(desc_t, kps_t, grid, queries, desc_q, blocked0, mode, th_accept, best_init, nnratio,
                         check_ori, count_twice, chi2_max=0.0, inv_level_sigma2=None):
    desc_t, desc_q = _f32(desc_t), _f32(desc_q)
    g = grid_struct(grid)
    n_t = len(kps_t)
    queries = np.ascontiguousarray(queries, dtype=PROJ_QUERY_DTYPE)
    b0 = None if blocked0 is None else _u8(blocked0)
    p = ProjParams(int(mode), th_accept, best_init, nnratio, int(check_ori), int(count_twice))
    if chi2_max > 0:
        inv = _f32(inv_level_sigma2)
        p.chi2_max, p.n_levels = float(chi2_max), len(inv)
        for l, v in enumerate(inv):
            p.inv_level_sigma2[l] = float(v)
    assign = np.empty(n_t, np.int32)
    dist = np.empty(n_t, np.float32)
    nm = lib().ora_search_by_projection(_p(desc_t), _p(kps_t), n_t, C.byref(g), _p(queries), _p(desc_q), len(queries),
                                        _p(b0), C.byref(p), _p(assign), _p(dist))
    return nm, assign, dist


def search_by_projection_q(desc_t, kps_t, grid, queries, desc_q, blocked0, mode, th_accept, best_init, nnratio,
                           chi2_max=0.0, inv_level_sigma2=None):
    """Per-query result of the same search: (best keypoint or -1, distance) for every query."""
    desc_t, desc_q = _f32(desc_t), _f32(desc_q)
    g = grid_struct(grid)
    queries = np.ascontiguousarray(queries, dtype=PROJ_QUERY_DTYPE)
    b0 = None if blocked0 is None else _u8(blocked0)
    p = ProjParams(int(mode), th_accept, best_init, nnratio, 0, 0)
    if chi2_max > 0:
        inv = _f32(inv_level_sigma2)
        p.chi2_max, p.n_levels = float(chi2_max), len(inv)
        for l, v in enumerate(inv):
            p.inv_level_sigma2[l] = float(v)
    best = np.empty(len(queries), np.int32)
    dist = np.empty(len(queries), np.float32)
    lib().ora_search_by_projection_q(_p(desc_t), _p(kps_t), len(kps_t), C.byref(g), _p(queries), _p(desc_q), len(queries),
                                     _p(b0), C.byref(p), _p(best), _p(dist))
    return best, dist


def search_for_initialization(desc1, kps1, desc2, kps2, grid2, prev_matched, window_size, nnratio, check_ori):
    desc1, desc2 = _f32(desc1), _f32(desc2)
    g = grid_struct(grid2)
    prev = np.array(prev_matched, dtype=np.float32, copy=True).reshape(-1)
    n1 = len(kps1)
    match = np.empty(n1, np.int32)
    dist = np.empty(n1, np.float32)
    nm = lib().ora_search_for_initialization(_p(desc1), _p(kps1), n1, _p(desc2), _p(kps2), len(kps2), C.byref(g),
                                             _p(prev), int(window_size), C.c_float(nnratio), int(check_ori), _p(match),
                                             _p(dist))
    return nm, match, dist, prev.reshape(-1, 2)


def stereo_row_match(desc_l, kps_l, desc_r, kps_r, scale_factors, n_rows, min_d, max_d):
    desc_l, desc_r, sf = _f32(desc_l), _f32(desc_r), _f32(scale_factors)
    n_l = len(kps_l)
    best = np.empty(n_l, np.int32)
    dist = np.empty(n_l, np.float32)
    lib().ora_stereo_row_match(_p(desc_l), _p(kps_l), n_l, _p(desc_r), _p(kps_r), len(kps_r), _p(sf), int(n_rows),
                               C.c_float(min_d), C.c_float(max_d), _p(best), _p(dist))
    return best, dist


def _kf_array(kfs, mask_key):
    """kfs: list of dicts with desc, kps, fv (FV), and mask under `mask_key`."""
    keep = []
    arr = (KeyFrame * len(kfs))()
    for k, kf in enumerate(kfs):
        d, m, fv = _f32(kf["desc"]), _u8(kf[mask_key]), FV.of(kf["fv"])
        keep += [d, m, fv, kf["kps"]]
        arr[k] = KeyFrame(_p(d), _p(kf["kps"]), C.pointer(fv.c), _p(m))
    return arr, keep


def batch_bow_kf_kf(kfs, ia, ib, nnratio, check_ori, threads=1, mask_key="valid"):
    arr, keep = _kf_array(kfs, mask_key)
    n = len(kfs[0]["kps"])
    ia, ib = np.ascontiguousarray(ia, np.int32), np.ascontiguousarray(ib, np.int32)
    P = len(ia)
    match = np.empty((P, n), np.int32)
    dist = np.empty((P, n), np.float32)
    nm = np.empty(P, np.int32)
    lib().ora_batch_bow_kf_kf(arr, n, _p(ia), _p(ib), P, C.c_float(nnratio), int(check_ori), int(threads), _p(match),
                              _p(dist), _p(nm))
    return nm, match, dist


def batch_triangulation(kfs, ia, ib, F12, epipoles, scale_factors, level_sigma2, check_ori, threads=1,
                        mask_key="has_mp"):
    arr, keep = _kf_array(kfs, mask_key)
    n = len(kfs[0]["kps"])
    ia, ib = np.ascontiguousarray(ia, np.int32), np.ascontiguousarray(ib, np.int32)
    F12, ep = _f32(F12).reshape(-1), _f32(epipoles).reshape(-1)
    sf, ls = _f32(scale_factors), _f32(level_sigma2)
    P = len(ia)
    match = np.empty((P, n), np.int32)
    dist = np.empty((P, n), np.float32)
    nm = np.empty(P, np.int32)
    lib().ora_batch_triangulation(arr, n, _p(ia), _p(ib), P, _p(F12), _p(ep), _p(sf), _p(ls), int(check_ori),
                                  int(threads), _p(match), _p(dist), _p(nm))
    return nm, match, dist


def max_threads():
    return int(lib().ora_max_threads())


def _obs_arrays(kf_descs, obs_off, obs_kf, obs_row):
    descs = [_f32(d) for d in kf_descs]
    ptrs = (C.c_void_p * len(descs))(*[d.ctypes.data for d in descs])
    off = np.ascontiguousarray(obs_off, np.int32)
    kf = np.ascontiguousarray(obs_kf, np.int32)
    row = np.ascontiguousarray(obs_row, np.int32)
    return descs, ptrs, off, kf, row


def distinctive(kf_descs, obs_off, obs_kf, obs_row):
    """MapPoint::ComputeDistinctiveDescriptors over a batch of map points (MapPoint.cc:271-338)."""
    descs, ptrs, off, kf, row = _obs_arrays(kf_descs, obs_off, obs_kf, obs_row)
    n = len(off) - 1
    best = np.empty(n, np.int32)
    med = np.empty(n, np.float32)
    lib().ora_distinctive(ptrs, n, _p(off), _p(kf), _p(row), _p(best), _p(med))
    return best, med


def voc_transform(tree, desc, levelsup=4):
    """TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (TemplatedVocabulary.h:1125-1192,
    1216-1257; FSift.cpp:86-101) on the flat tree `tree` (keys as asdm_vocab_desc)."""
    desc = _f32(desc)
    n = len(desc)
    off = np.ascontiguousarray(tree["child_off"], np.int32)
    idx = np.ascontiguousarray(tree["child_idx"], np.int32)
    nd = _f32(tree["desc"])
    w = np.ascontiguousarray(tree["weight"], np.float64)
    word = np.ascontiguousarray(tree["word_id"], np.int32)
    word_id = np.empty(n, np.int32)
    node_id = np.empty(n, np.int32)
    inc = np.empty(n, np.uint8)
    bid = np.empty(max(n, 1), np.uint32)
    bval = np.empty(max(n, 1), np.float64)
    f = lib().ora_voc_transform
    f.restype = C.c_int
    nb = f(C.c_int(int(tree["L"])), C.c_int(int(tree["weighting"])), C.c_int(int(tree["scoring"])), C.c_int(len(w)),
           _p(off), _p(idx), _p(nd), _p(w), _p(word), _p(desc), C.c_int(n), C.c_int(int(levelsup)), _p(word_id),
           _p(node_id), _p(inc), _p(bid), _p(bval))
    return dict(word_id=word_id, node_id=node_id, in_fv=inc, bow_id=bid[:nb].copy(), bow_val=bval[:nb].copy())


def map_write_block(desc):
    """Descriptor block of one frame in chamo.map (System.cc:333-343 + visual_map_seri.cc:134-146) as bytes."""
    desc = _f32(desc)
    n, d = desc.shape
    out = np.empty(8 + n * d * 4, np.uint8)
    f = lib().ora_map_write_block
    f.restype = C.c_long
    nb = f(_p(desc), C.c_int(n), C.c_int(d), _p(out))
    assert nb == len(out)
    return out


def map_read_block(block):
    """vm::loader_visual_map's descriptor block reader (visual_map_seri.cc:272-285): returns the N x D float rows."""
    block = np.ascontiguousarray(block, np.uint8)
    w = C.c_int()
    f = lib().ora_map_read_block
    n = f(_p(block), C.c_long(len(block)), None, C.c_int(0), C.byref(w))
    if n < 0:
        return None
    rows = np.empty((n, w.value), np.float32)
    f(_p(block), C.c_long(len(block)), _p(rows), C.c_int(n), C.byref(w))
    return rows


def _db_args(db):
    return (C.c_int(int(db["n_kf"])), _p(np.ascontiguousarray(db["bow_off"], np.int32)),
            _p(np.ascontiguousarray(db["bow_word"], np.uint32)), _p(np.ascontiguousarray(db["bow_val"], np.float64)),
            _p(np.ascontiguousarray(db["alive"], np.uint8)))


def bow_score_l1(w1, v1, w2, v2, _lib=None):
    """DBoW2::L1Scoring::score (ScoringObject.cpp:24-61) of two BowVectors given as ascending (word, value) arrays."""
    w1, w2 = np.ascontiguousarray(w1, np.uint32), np.ascontiguousarray(w2, np.uint32)
    v1, v2 = np.ascontiguousarray(v1, np.float64), np.ascontiguousarray(v2, np.float64)
    f = (_lib or lib()).ora_bow_score_l1 if _lib is None else _lib.ref_bow_score_l1
    f.restype = C.c_double
    return float(f(_p(w1), _p(v1), C.c_int(len(w1)), _p(w2), _p(v2), C.c_int(len(w2))))


def detect_loop_candidates(db, q_word, q_val, connected, min_score, only_global_map=False, _ref=None):
    """KeyFrameDatabase::DetectLoopCandidates (KeyFrameDatabase.cc:80-216) over the flat database `db`
    (synth.config_kfdb keys).  Returns (candidate slots in the reference's order, mnLoopWords, mLoopScore)."""
    n = int(db["n_kf"])
    qw, qv = np.ascontiguousarray(q_word, np.uint32), np.ascontiguousarray(q_val, np.float64)
    conn, glob = np.ascontiguousarray(connected, np.uint8), np.ascontiguousarray(db["global_flag"], np.uint8)
    no, ns = np.ascontiguousarray(db["neigh_off"], np.int32), np.ascontiguousarray(db["neigh_slot"], np.int32)
    out, common, score = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.float32)
    tail = (_p(out), _p(common), _p(score))
    if _ref is None:
        k = lib().ora_detect_loop_candidates(*_db_args(db), _p(qw), _p(qv), C.c_int(len(qw)), _p(conn), _p(glob), _p(no), _p(ns),
                                             C.c_float(min_score), C.c_int(int(only_global_map)), *tail)
    else:
        k = _ref.ref_detect_loop_candidates(*_db_args(db), _p(qw), _p(qv), C.c_int(len(qw)), _p(conn), _p(glob), _p(no), _p(ns),
                                            C.c_float(min_score), C.c_int(int(only_global_map)), C.c_uint(int(db["n_words"])),
                                            *tail)
    return out[:k].copy(), common[:n], score[:n]


def detect_reloc_candidates(db, q_word, q_val, reloc_score, only_global_map=False, _ref=None):
    """KeyFrameDatabase::DetectRelocalizationCandidates (KeyFrameDatabase.cc:219-320). reloc_score: the keyframes'
    persistent mRelocScore (float32 [n_kf], updated in place).  Returns (candidate slots, mnRelocWords)."""
    n = int(db["n_kf"])
    assert reloc_score.dtype == np.float32 and reloc_score.flags["C_CONTIGUOUS"] and len(reloc_score) == n
    qw, qv = np.ascontiguousarray(q_word, np.uint32), np.ascontiguousarray(q_val, np.float64)
    glob = np.ascontiguousarray(db["global_flag"], np.uint8)
    no, ns = np.ascontiguousarray(db["neigh_off"], np.int32), np.ascontiguousarray(db["neigh_slot"], np.int32)
    out, common = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.int32)
    if _ref is None:
        k = lib().ora_detect_reloc_candidates(*_db_args(db), _p(qw), _p(qv), C.c_int(len(qw)), _p(glob), _p(no), _p(ns),
                                              C.c_int(int(only_global_map)), _p(reloc_score), _p(out), _p(common))
    else:
        k = _ref.ref_detect_reloc_candidates(*_db_args(db), _p(qw), _p(qv), C.c_int(len(qw)), _p(glob), _p(no), _p(ns),
                                             C.c_int(int(only_global_map)), C.c_uint(int(db["n_words"])), _p(reloc_score),
                                             _p(out), _p(common))
    return out[:k].copy(), common[:n]
