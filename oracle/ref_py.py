"""ctypes wrapper over oracle/_ref/libref_orbmatcher.so -- the reference's OWN ORBmatcher.cc compiled in place
against header shims (oracle/Makefile `ref`, oracle/ref_driver.cc).  TEST INFRASTRUCTURE ONLY.

Used by tests/test_oracle_vs_ref.py to pin the CPU oracle and by tests/golden/make_golden.py to generate the
committed fixtures.  The library can only be BUILT where /root/reference exists (this container); the prebuilt .so
travels to the GPU box.  `available()` says whether it is there.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from oracle_py import FV, Grid, PROJ_QUERY_DTYPE, _f32, _p, _u8, grid_struct

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_ref", "libref_orbmatcher.so")
REF_ROOT = "/root/reference/src/vslam"


def build():
    if os.path.isdir(REF_ROOT):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)
    return os.path.exists(_LIB)


def available():
    return os.path.exists(_LIB) or build()


_lib = None
_path = None  # library the ref_* calls go to; None = the reference build


def use(path=None):
    """Route the ref_* calls to another build of ref_driver.cc: the C++ adaptor under test
    (asd-slam_b200/host/libadaptor_shim.so = our ORBmatcher_cuda.cc behind the same harness). None = the reference."""
    global _lib, _path
    _lib, _path = None, path


def lib():
    global _lib
    if _lib is None:
        if _path is None and not available():
            raise RuntimeError("oracle/_ref/libref_orbmatcher.so is missing and /root/reference is not mounted")
        _lib = C.CDLL(_path or _LIB)
        _lib.ref_descriptor_distance.restype = C.c_float
        _lib.ref_radius_by_viewing_cos.restype = C.c_float
        _lib.ref_radius_by_viewing_cos.argtypes = [C.c_float]
        _lib.ref_predict_scale.argtypes = [C.c_float, C.c_float, C.c_float, C.c_int]
        _lib.ref_check_dist_epipolar_line.argtypes = [C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_void_p,
                                                      C.c_void_p, C.c_int]
    return _lib


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def descriptor_distance(a, b):
    a, b = _f32(a), _f32(b)
    return float(lib().ref_descriptor_distance(_p(a), _p(b)))


def compute_three_maxima(sizes):
    s = _i32(sizes)
    i1, i2, i3 = C.c_int32(-1), C.c_int32(-1), C.c_int32(-1)
    lib().ref_compute_three_maxima(_p(s), len(s), C.byref(i1), C.byref(i2), C.byref(i3))
    return i1.value, i2.value, i3.value


def radius_by_viewing_cos(c):
    return float(lib().ref_radius_by_viewing_cos(float(c)))


def predict_scale(max_distance, current_dist, log_scale_factor, n_levels):
    return int(lib().ref_predict_scale(float(max_distance), float(current_dist), float(log_scale_factor), int(n_levels)))


def check_dist_epipolar_line(x1, y1, x2, y2, octave2, F12, level_sigma2):
    F12, ls = _f32(F12).reshape(9), _f32(level_sigma2)
    return int(lib().ref_check_dist_epipolar_line(float(x1), float(y1), float(x2), float(y2), int(octave2), _p(F12),
                                                  _p(ls), len(ls)))


def frame_features_in_area(kps, grid, x, y, r, min_level=-1, max_level=-1):
    g = grid_struct(grid)
    out = np.zeros(max(len(kps), 1), np.int32)
    n = lib().ref_frame_features_in_area(_p(kps), len(kps), C.byref(g), C.c_float(x), C.c_float(y), C.c_float(r),
                                         int(min_level), int(max_level), _p(out))
    return out[:n].copy()


def keyframe_features_in_area(kps, grid, x, y, r):
    g = grid_struct(grid)
    out = np.zeros(max(len(kps), 1), np.int32)
    n = lib().ref_keyframe_features_in_area(_p(kps), len(kps), C.byref(g), C.c_float(x), C.c_float(y), C.c_float(r),
                                            _p(out))
    return out[:n].copy()


def search_by_bow_kf_f(desc_kf, kps_kf, fv_kf, valid_kf, desc_f, kps_f, fv_f, nnratio, check_ori):
    desc_kf, desc_f, valid_kf = _f32(desc_kf), _f32(desc_f), _u8(valid_kf)
    fa, fb = FV.of(fv_kf), FV.of(fv_f)
    match = np.empty(len(kps_f), np.int32)
    nm = lib().ref_search_by_bow_kf_f(_p(desc_kf), _p(kps_kf), len(kps_kf), C.byref(fa.c), _p(valid_kf), _p(desc_f),
                                      _p(kps_f), len(kps_f), C.byref(fb.c), C.c_float(nnratio), int(check_ori),
                                      _p(match))
    return nm, match


def search_by_bow_kf_kf(desc1, kps1, fv1, valid1, desc2, kps2, fv2, valid2, nnratio, check_ori):
    desc1, desc2, valid1, valid2 = _f32(desc1), _f32(desc2), _u8(valid1), _u8(valid2)
    fa, fb = FV.of(fv1), FV.of(fv2)
    match = np.empty(len(kps1), np.int32)
    nm = lib().ref_search_by_bow_kf_kf(_p(desc1), _p(kps1), len(kps1), C.byref(fa.c), _p(valid1), _p(desc2), _p(kps2),
                                       len(kps2), C.byref(fb.c), _p(valid2), C.c_float(nnratio), int(check_ori),
                                       _p(match))
    return nm, match


def search_for_triangulation(desc1, kps1, fv1, has_mp1, desc2, kps2, fv2, has_mp2, F12, ex, ey, scale_factors2,
                             level_sigma2_2, check_ori):
    desc1, desc2, has_mp1, has_mp2 = _f32(desc1), _f32(desc2), _u8(has_mp1), _u8(has_mp2)
    F12, sf, ls = _f32(F12).reshape(9), _f32(scale_factors2), _f32(level_sigma2_2)
    fa, fb = FV.of(fv1), FV.of(fv2)
    match = np.empty(len(kps1), np.int32)
    nm = lib().ref_search_for_triangulation(_p(desc1), _p(kps1), len(kps1), C.byref(fa.c), _p(has_mp1), _p(desc2),
                                            _p(kps2), len(kps2), C.byref(fb.c), _p(has_mp2), _p(F12), C.c_float(ex),
                                            C.c_float(ey), _p(sf), _p(ls), len(sf), int(check_ori), _p(match))
    return nm, match


def search_for_initialization(desc1, kps1, desc2, kps2, grid2, prev_matched, window_size, nnratio, check_ori):
    desc1, desc2 = _f32(desc1), _f32(desc2)
    g = grid_struct(grid2)
    prev = np.array(prev_matched, dtype=np.float32, copy=True).reshape(-1)
    match = np.empty(len(kps1), np.int32)
    nm = lib().ref_search_for_initialization(_p(desc1), _p(kps1), len(kps1), _p(desc2), _p(kps2), len(kps2),
                                             C.byref(g), _p(prev), int(window_size), C.c_float(nnratio),
                                             int(check_ori), _p(match))
    return nm, match, prev.reshape(-1, 2)


def _proj_common(desc_t, kps_t, grid, queries, level, desc_q, blocked0, scale_factors):
    desc_t, desc_q, sf = _f32(desc_t), _f32(desc_q), _f32(scale_factors)
    q = np.ascontiguousarray(queries, dtype=PROJ_QUERY_DTYPE)
    b0 = None if blocked0 is None else _u8(blocked0)
    return desc_t, desc_q, sf, q, _i32(level), b0, grid_struct(grid), np.empty(len(kps_t), np.int32)


def sbp_local_map(desc_t, kps_t, grid, queries, level, view_cos, desc_q, blocked0, scale_factors, th, nnratio):
    desc_t, desc_q, sf, q, lvl, b0, g, assign = _proj_common(desc_t, kps_t, grid, queries, level, desc_q, blocked0,
                                                              scale_factors)
    vc = _f32(view_cos)
    nm = lib().ref_sbp_local_map(_p(desc_t), _p(kps_t), len(kps_t), C.byref(g), _p(q), _p(lvl), _p(vc), _p(desc_q),
                                 len(q), _p(b0), _p(sf), len(sf), C.c_float(th), C.c_float(nnratio), _p(assign))
    return nm, assign


def sbp_frame_frame(desc_t, kps_t, grid, queries, level, desc_q, blocked0, scale_factors, th, nnratio, check_ori):
    desc_t, desc_q, sf, q, lvl, b0, g, assign = _proj_common(desc_t, kps_t, grid, queries, level, desc_q, blocked0,
                                                              scale_factors)
    nm = lib().ref_sbp_frame_frame(_p(desc_t), _p(kps_t), len(kps_t), C.byref(g), _p(q), _p(lvl), _p(desc_q), len(q),
                                   _p(b0), _p(sf), len(sf), C.c_float(th), C.c_float(nnratio), int(check_ori),
                                   _p(assign))
    return nm, assign


def sbp_frame_keyframe(desc_t, kps_t, grid, queries, level, desc_q, blocked0, scale_factors, th, orb_dist, nnratio,
                       check_ori):
    desc_t, desc_q, sf, q, lvl, b0, g, assign = _proj_common(desc_t, kps_t, grid, queries, level, desc_q, blocked0,
                                                              scale_factors)
    nm = lib().ref_sbp_frame_keyframe(_p(desc_t), _p(kps_t), len(kps_t), C.byref(g), _p(q), _p(lvl), _p(desc_q),
                                      len(q), _p(b0), _p(sf), len(sf), C.c_float(th), C.c_float(orb_dist),
                                      C.c_float(nnratio), int(check_ori), _p(assign))
    return nm, assign


def sbp_keyframe_sim3(desc_t, kps_t, grid, queries, level, desc_q, blocked0, scale_factors, th, nnratio):
    desc_t, desc_q, sf, q, lvl, b0, g, assign = _proj_common(desc_t, kps_t, grid, queries, level, desc_q, blocked0,
                                                              scale_factors)
    nm = lib().ref_sbp_keyframe_sim3(_p(desc_t), _p(kps_t), len(kps_t), C.byref(g), _p(q), _p(lvl), _p(desc_q), len(q),
                                     _p(b0), _p(sf), len(sf), int(th), C.c_float(nnratio), _p(assign))
    return nm, assign


def fuse(desc_t, kps_t, grid, queries, level, desc_q, occ_obs, mp_obs, scale_factors, th):
    desc_t, desc_q, sf, q, lvl, _, g, _ = _proj_common(desc_t, kps_t, grid, queries, level, desc_q, None, scale_factors)
    occ, mob = _i32(occ_obs), _i32(mp_obs)
    action = np.empty(len(q), np.int32)
    nf = lib().ref_fuse(_p(desc_t), _p(kps_t), len(kps_t), C.byref(g), _p(q), _p(lvl), _p(desc_q), len(q), _p(occ), _p(mob),
                        _p(sf), len(sf), C.c_float(th), _p(action))
    return nf, action


def fuse_sim3(desc_t, kps_t, grid, queries, level, desc_q, occ_obs, mp_obs, scale_factors, th):
    desc_t, desc_q, sf, q, lvl, _, g, _ = _proj_common(desc_t, kps_t, grid, queries, level, desc_q, None, scale_factors)
    occ, mob = _i32(occ_obs), _i32(mp_obs)
    replace, added = np.empty(len(q), np.int32), np.empty(len(q), np.int32)
    nf = lib().ref_fuse_sim3(_p(desc_t), _p(kps_t), len(kps_t), C.byref(g), _p(q), _p(lvl), _p(desc_q), len(q), _p(occ),
                             _p(mob), _p(sf), len(sf), C.c_float(th), _p(replace), _p(added))
    return nf, replace, added


def search_by_sim3(desc1, kps1, desc2, kps2, grid, valid1, valid2, uv1, uv2, level1, level2, mpdesc1, mpdesc2, already1,
                   scale_factors, th):
    desc1, desc2, mp1, mp2, sf = _f32(desc1), _f32(desc2), _f32(mpdesc1), _f32(mpdesc2), _f32(scale_factors)
    v1, v2, a1 = _u8(valid1), _u8(valid2), _u8(already1)
    uv1, uv2, l1, l2 = _f32(uv1), _f32(uv2), _i32(level1), _i32(level2)
    g = grid_struct(grid)
    match = np.empty(len(kps1), np.int32)
    nf = lib().ref_search_by_sim3(_p(desc1), _p(kps1), len(kps1), _p(desc2), _p(kps2), len(kps2), C.byref(g), _p(v1), _p(v2),
                                  _p(uv1), _p(uv2), _p(l1), _p(l2), _p(mp1), _p(mp2), _p(a1), _p(sf), len(sf),
                                  C.c_float(th), _p(match))
    return nf, match


def distinctive(kf_descs, obs_off, obs_kf, obs_row):
    """The reference's own MapPoint::ComputeDistinctiveDescriptors; returns the chosen descriptor of every point
    (NaN rows for points without observations, which the reference leaves untouched)."""
    descs = [_f32(d) for d in kf_descs]
    ptrs = (C.c_void_p * len(descs))(*[d.ctypes.data for d in descs])
    ns = _i32([len(d) for d in descs])
    off, kf, row = _i32(obs_off), _i32(obs_kf), _i32(obs_row)
    n = len(off) - 1
    out = np.full((n, 128), np.nan, np.float32)
    lib().ref_distinctive(ptrs, _p(ns), len(descs), n, _p(off), _p(kf), _p(row), _p(out))
    return out


# ------------------------------------------------------------------ vocabulary tree (oracle/ref_voc_driver.cc)
_VLIB = os.path.join(_HERE, "_ref", "libref_vocab.so")
_vlib = None
_vpath = None


def use_voc(path=None):
    """Route the ref_voc_* calls to another build of ref_voc_driver.cc: the drop-in DeviceVocabulary under test
    (asd-slam_b200/host/libadaptor_vocab_shim.so). None = the reference."""
    global _vlib, _vpath
    _vlib, _vpath = None, path


def voc_available():
    return os.path.exists(_VLIB) or (build() and os.path.exists(_VLIB))


def vlib():
    global _vlib
    if _vlib is None:
        if _vpath is None and not voc_available():
            raise RuntimeError("oracle/_ref/libref_vocab.so is missing and /root/reference is not mounted")
        _vlib = C.CDLL(_vpath or _VLIB)
        _vlib.ref_voc_train.restype = C.c_void_p
        _vlib.ref_voc_load.restype = C.c_void_p
        _vlib.ref_fsift_distance.restype = C.c_double
    return _vlib


class RefVocabulary:
    """The reference's own TemplatedVocabulary<FSift::TDescriptor, FSift> behind a handle."""

    def __init__(self, h):
        self.h = C.c_void_p(h)

    @classmethod
    def train(cls, desc, n_images, k, L, weighting=0, scoring=0, seed=1):
        desc = _f32(desc)
        return cls(vlib().ref_voc_train(_p(desc), len(desc), n_images, k, L, weighting, scoring, C.c_uint(seed)))

    @classmethod
    def load(cls, tree):
        """tree: flat arrays as asdm_vocab_desc; goes through the reference's load() (text descriptors)."""
        off, idx = _i32(tree["child_off"]), _i32(tree["child_idx"])
        nn = len(off) - 1
        parent = np.zeros(nn, np.int32)
        for p in range(nn):
            parent[idx[off[p]:off[p + 1]]] = p
        desc, w, word = _f32(tree["desc"]), np.ascontiguousarray(tree["weight"], np.float64), _i32(tree["word_id"])
        leaves = np.nonzero(off[1:] == off[:-1])[0]
        word_node = np.zeros(len(leaves), np.int32)
        word_node[word[leaves]] = leaves
        return cls(vlib().ref_voc_load(int(tree["k"]), int(tree["L"]), int(tree["weighting"]), int(tree["scoring"]), nn,
                                       _p(parent), _p(desc), _p(w), len(leaves), _p(word_node)))

    def export(self):
        sizes = np.zeros(6, np.int32)
        vlib().ref_voc_sizes(self.h, _p(sizes))
        nn = int(sizes[0])
        off, idx = np.zeros(nn + 1, np.int32), np.zeros(max(nn - 1, 1), np.int32)
        desc, w, word = np.zeros((nn, 128), np.float32), np.zeros(nn, np.float64), np.zeros(nn, np.int32)
        vlib().ref_voc_export(self.h, _p(off), _p(idx), _p(desc), _p(w), _p(word))
        return dict(k=int(sizes[1]), L=int(sizes[2]), weighting=int(sizes[3]), scoring=int(sizes[4]), child_off=off,
                    child_idx=idx[:nn - 1], desc=desc, weight=w, word_id=word)

    def transform(self, desc, levelsup=4):
        desc = _f32(desc)
        n = len(desc)
        word, node, order = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(max(n, 1), np.int32)
        bid, bval = np.empty(max(n, 1), np.uint32), np.empty(max(n, 1), np.float64)
        nfv = C.c_int32()
        nb = vlib().ref_voc_transform(self.h, _p(desc), n, levelsup, _p(word), _p(node), _p(order), C.byref(nfv), _p(bid),
                                      _p(bval))
        return dict(word_id=word, fv_node=node, fv_order=order[:nfv.value].copy(), bow_id=bid[:nb].copy(),
                    bow_val=bval[:nb].copy())

    def free(self):
        if self.h:
            vlib().ref_voc_free(self.h)
            self.h = None


def fsift_distance(a, b):
    a, b = _f32(a), _f32(b)
    return vlib().ref_fsift_distance(_p(a), _p(b))


# ------------------------------------------------------------------ chamo.map descriptor blocks (oracle/ref_map_driver.cc)
_MLIB = os.path.join(_HERE, "_ref", "libref_map.so")


def map_available():
    return os.path.exists(_MLIB) or (build() and os.path.exists(_MLIB))


def map_write_block(desc, path):
    """System::saveToVisualMap + vm::save_visual_map, the reference's own loops; returns the file's bytes."""
    desc = _f32(desc)
    m = C.CDLL(_MLIB)
    m.ref_map_write_block.restype = C.c_long
    nb = m.ref_map_write_block(_p(desc), len(desc), path.encode())
    data = np.fromfile(path, np.uint8)
    assert nb == len(data)
    return data


def map_read_block(path, cap_rows):
    m = C.CDLL(_MLIB)
    rows = np.zeros((max(cap_rows, 1), 128), np.float32)
    dims = np.zeros(2, np.int32)
    n = m.ref_map_read_block(path.encode(), _p(rows), cap_rows, _p(dims))
    return rows[:n], dims


# ---------------------------------------------------------------- KeyFrameDatabase.cc (oracle/_ref/libref_kfdb.so)
_KLIB = os.path.join(_HERE, "_ref", "libref_kfdb.so")
_klib = None


def kfdb_available():
    return os.path.exists(_KLIB) or (build() and os.path.exists(_KLIB))


def klib():
    global _klib
    if _klib is None:
        if not kfdb_available():
            raise RuntimeError("oracle/_ref/libref_kfdb.so is missing and /root/reference is not mounted")
        _klib = C.CDLL(_KLIB)
    return _klib


def bow_score_l1(w1, v1, w2, v2):
    import oracle_py
    return oracle_py.bow_score_l1(w1, v1, w2, v2, _lib=klib())


def detect_loop_candidates(db, q_word, q_val, connected, min_score, only_global_map=False):
    import oracle_py
    return oracle_py.detect_loop_candidates(db, q_word, q_val, connected, min_score, only_global_map, _ref=klib())


def detect_reloc_candidates(db, q_word, q_val, reloc_score, only_global_map=False):
    import oracle_py
    return oracle_py.detect_reloc_candidates(db, q_word, q_val, reloc_score, only_global_map, _ref=klib())


def stereo_row_match(desc_l, kps_l, desc_r, kps_r, scale_factors, n_rows, max_d):
    """Descriptor stage of the reference's Frame::ComputeStereoMatches (Frame.cc:366-447, its own text): (best_r, dist)."""
    desc_l, desc_r, sf = _f32(desc_l), _f32(desc_r), _f32(scale_factors)
    best = np.empty(len(kps_l), np.int32)
    dist = np.empty(len(kps_l), np.float32)
    lib().ref_stereo_row_match(_p(desc_l), _p(kps_l), len(kps_l), _p(desc_r), _p(kps_r), len(kps_r), _p(sf), len(sf),
                               int(n_rows), C.c_float(max_d), _p(best), _p(dist))
    return best, dist


def set_scene(mode=0, seed=0):
    """mode 0: identity poses (the flat goldens); mode 1: real intrinsics, random SE3 / Sim3 poses, points behind the
    camera, outside the distance band and beyond the viewing gate (oracle/ref_driver.cc `Scene`). Global: reset to 0."""
    lib().ref_set_scene(int(mode), C.c_uint(int(seed)))
