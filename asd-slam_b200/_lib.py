"""ctypes binding of libasdmatch_cuda.so (include/asdmatch.h).

The product path has no CPU fallback: if the library is missing, cannot be loaded, or finds no B200, the
calls below raise.  Nothing in this package imports oracle/.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libasdmatch_cuda.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("angle", "<f4"), ("octave", "<i4")])
PROJ_QUERY_DTYPE = np.dtype(
    [("valid", "u1"), ("blocks", "u1"), ("pad", "<i2"), ("u", "<f4"), ("v", "<f4"), ("r", "<f4"),
     ("min_level", "<i4"), ("max_level", "<i4"), ("angle", "<f4")])
assert KP_DTYPE.itemsize == 16 and PROJ_QUERY_DTYPE.itemsize == 28

ASDM_JOB_BOW_KF_KF = 1
ASDM_JOB_TRIANGULATION = 2
ASDM_PROJ_RATIO_LEVEL = 0
ASDM_PROJ_BEST_ONLY = 1
TH_LOW = 0.5
TH_HIGH = 1.5
HISTO_LENGTH = 30

# every symbol include/asdmatch.h declares (tests/test_abi.py checks the .so exports exactly these)
EXPORTS = [
    "asdm_create", "asdm_fork", "asdm_destroy", "asdm_last_error", "asdm_stream", "asdm_synchronize", "asdm_set_stream",
    "asdm_host_alloc", "asdm_host_free", "asdm_set_option", "asdm_get_stat",
    "asdm_set_create", "asdm_set_create_async", "asdm_set_create_batch", "asdm_set_retain", "asdm_set_release", "asdm_set_size", "asdm_set_featvec",
    "asdm_set_read_desc", "asdm_desc_block_bytes", "asdm_set_save_blocks", "asdm_set_load_block",
    "asdm_search_by_bow_kf_f", "asdm_search_by_bow_kf_kf", "asdm_search_for_triangulation",
    "asdm_search_by_projection", "asdm_search_by_projection_q", "asdm_search_for_initialization", "asdm_stereo_row_match",
    "asdm_stereo_row_match_static",
    "asdm_job_prepare", "asdm_job_run", "asdm_job_fetch", "asdm_job_fetch_async", "asdm_job_settle", "asdm_job_out_stride",
    "asdm_job_device_outputs", "asdm_job_destroy",
    "asdm_shared_alloc", "asdm_shared_open", "asdm_shared_close", "asdm_shared_free", "asdm_job_set_publish",
    "asdm_group_create", "asdm_group_join", "asdm_group_destroy", "asdm_group_last_error", "asdm_group_world",
    "asdm_group_local_count", "asdm_group_local_rank", "asdm_group_ctx", "asdm_group_host_alloc", "asdm_group_host_free",
    "asdm_group_host_barrier", "asdm_group_sets_create", "asdm_group_job_prepare", "asdm_group_plan_units", "asdm_group_job_units",
    "asdm_group_job_stride", "asdm_group_job_run", "asdm_group_job_wait", "asdm_group_job_counts", "asdm_group_job_fetch",
    "asdm_group_job_gathered", "asdm_group_job_flush", "asdm_group_job_destroy",
    "asdm_distinctive_descriptors",
    "asdm_vocab_create", "asdm_vocab_release", "asdm_vocab_transform",
    "asdm_bowdb_create", "asdm_bowdb_release", "asdm_bowdb_add", "asdm_bowdb_erase", "asdm_bowdb_clear", "asdm_bowdb_size",
    "asdm_bowdb_query", "asdm_bowdb_detect_loop", "asdm_bowdb_detect_reloc", "asdm_bowdb_reloc_scores",
    "asdm_bowdb_detect_loop_full", "asdm_bowdb_detect_reloc_full", "asdm_bowdb_set_reloc_scores",
]


class BowdbGates(C.Structure):
    _fields_ = [("connected", C.c_void_p), ("global_flag", C.c_void_p), ("neigh_off", C.c_void_p), ("neigh_slot", C.c_void_p),
                ("min_score", C.c_float), ("only_global_map", C.c_int32)]


class Grid(C.Structure):
    _fields_ = [("min_x", C.c_float), ("min_y", C.c_float), ("max_x", C.c_float), ("max_y", C.c_float),
                ("inv_w", C.c_float), ("inv_h", C.c_float), ("cols", C.c_int32), ("rows", C.c_int32)]


class FeatVec(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("node_ids", C.c_void_p), ("off", C.c_void_p), ("idx", C.c_void_p)]


class SetDesc(C.Structure):
    _fields_ = [("desc", C.c_void_p), ("n", C.c_int32), ("row_stride_bytes", C.c_int32), ("kps", C.c_void_p),
                ("grid", C.POINTER(Grid)), ("featvec", C.POINTER(FeatVec))]


class ProjParams(C.Structure):
    _fields_ = [("mode", C.c_int32), ("th_accept", C.c_float), ("best_init", C.c_float), ("nnratio", C.c_float),
                ("check_ori", C.c_int32), ("count_twice", C.c_int32), ("chi2_max", C.c_double), ("n_levels", C.c_int32),
                ("reserved", C.c_int32), ("inv_level_sigma2", C.c_float * 32)]


class JobDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("n_kf", C.c_int32), ("sets", C.POINTER(C.c_void_p)),
                ("masks", C.POINTER(C.c_void_p)), ("n_pairs", C.c_int32), ("ia", C.c_void_p), ("ib", C.c_void_p),
                ("nnratio", C.c_float), ("check_ori", C.c_int32), ("F12", C.c_void_p), ("epipoles", C.c_void_p),
                ("scale_factors", C.c_void_p), ("level_sigma2", C.c_void_p), ("n_levels", C.c_int32),
                ("row_begin", C.c_void_p), ("row_end", C.c_void_p)]


class GJobDesc(C.Structure):
    _fields_ = [("job", JobDesc), ("blocks_per_pair", C.c_int32), ("gather", C.c_int32)]


ASDM_GATHER_DEVICE = 1
ASDM_GATHER_NO_DIST = 2
ALLGATHER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t)


class VocabDesc(C.Structure):
    _fields_ = [("k", C.c_int32), ("L", C.c_int32), ("weighting", C.c_int32), ("scoring", C.c_int32),
                ("n_nodes", C.c_int32), ("child_off", C.c_void_p), ("child_idx", C.c_void_p), ("desc", C.c_void_p),
                ("weight", C.c_void_p), ("word_id", C.c_void_p)]


class AsdmError(RuntimeError):
    pass


_lib = None


def load():
    """Loads libasdmatch_cuda.so; raises if it was not built (python -c 'import __graft_entry__ as g; g.build()')."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise AsdmError(f"{LIB_PATH} is missing: build it with `make -C {os.path.join(_HERE, 'csrc')}` "
                        "(there is no CPU fallback for the matcher)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, f32 = C.c_void_p, C.c_int32, C.c_float
    lib.asdm_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.asdm_fork.argtypes = [vp, C.POINTER(vp)]
    lib.asdm_destroy.argtypes = [vp]
    lib.asdm_last_error.argtypes = [vp]
    lib.asdm_last_error.restype = C.c_char_p
    lib.asdm_stream.argtypes = [vp]
    lib.asdm_stream.restype = vp
    lib.asdm_synchronize.argtypes = [vp]
    lib.asdm_set_stream.argtypes = [vp, vp]
    lib.asdm_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    lib.asdm_host_free.argtypes = [vp]
    lib.asdm_set_option.argtypes = [vp, C.c_char_p, C.c_int]
    lib.asdm_get_stat.argtypes = [vp, C.c_char_p, C.POINTER(C.c_int64)]
    lib.asdm_set_create.argtypes = [vp, vp, C.c_int, C.c_int, vp, C.POINTER(Grid), C.POINTER(vp)]
    lib.asdm_set_create_async.argtypes = lib.asdm_set_create.argtypes
    lib.asdm_set_create_batch.argtypes = [vp, C.POINTER(SetDesc), C.c_int, C.POINTER(vp)]
    lib.asdm_set_retain.argtypes = [vp]
    lib.asdm_set_release.argtypes = [vp]
    lib.asdm_set_size.argtypes = [vp]
    lib.asdm_set_featvec.argtypes = [vp, C.POINTER(FeatVec)]
    lib.asdm_set_read_desc.argtypes = [vp, C.c_int, C.c_int, vp]
    lib.asdm_search_by_bow_kf_f.argtypes = [vp, vp, vp, vp, f32, C.c_int, vp, vp, C.POINTER(i32)]
    lib.asdm_search_by_bow_kf_kf.argtypes = [vp, vp, vp, vp, vp, f32, C.c_int, vp, vp, C.POINTER(i32)]
    lib.asdm_search_for_triangulation.argtypes = [vp, vp, vp, vp, vp, vp, f32, f32, vp, vp, C.c_int, C.c_int, vp, vp,
                                                  C.POINTER(i32)]
    lib.asdm_search_by_projection.argtypes = [vp, vp, vp, C.c_int, vp, vp, vp, vp, C.POINTER(ProjParams), vp, vp,
                                              C.POINTER(i32)]
    lib.asdm_search_by_projection_q.argtypes = [vp, vp, vp, C.c_int, vp, vp, vp, vp, C.POINTER(ProjParams), vp, vp,
                                                C.POINTER(i32)]
    lib.asdm_search_for_initialization.argtypes = [vp, vp, vp, vp, C.c_int, f32, C.c_int, vp, vp, C.POINTER(i32)]
    lib.asdm_stereo_row_match.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, f32, f32, vp, vp]
    lib.asdm_stereo_row_match_static.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, f32, f32, vp, vp]
    lib.asdm_job_prepare.argtypes = [vp, C.POINTER(JobDesc), C.POINTER(vp)]
    lib.asdm_job_run.argtypes = [vp]
    lib.asdm_job_fetch.argtypes = [vp, vp, vp, vp]
    lib.asdm_job_fetch_async.argtypes = [vp, vp, vp, vp]
    lib.asdm_job_settle.argtypes = [vp]
    lib.asdm_job_out_stride.argtypes = [vp]
    lib.asdm_job_device_outputs.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    lib.asdm_shared_alloc.argtypes = [vp, C.c_size_t, C.POINTER(vp), vp]
    lib.asdm_shared_open.argtypes = [vp, vp, C.POINTER(vp)]
    lib.asdm_shared_close.argtypes = [vp, vp]
    lib.asdm_shared_free.argtypes = [vp, vp]
    lib.asdm_job_set_publish.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.c_int, C.c_int, C.c_int]
    lib.asdm_job_destroy.argtypes = [vp]
    lib.asdm_group_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_size_t, C.POINTER(vp)]
    lib.asdm_group_join.argtypes = [vp, C.c_int, C.c_int, C.c_size_t, ALLGATHER_FN, vp, C.POINTER(vp)]
    lib.asdm_group_destroy.argtypes = [vp]
    lib.asdm_group_last_error.argtypes = [vp]
    lib.asdm_group_last_error.restype = C.c_char_p
    lib.asdm_group_world.argtypes = [vp]
    lib.asdm_group_local_count.argtypes = [vp]
    lib.asdm_group_local_rank.argtypes = [vp, C.c_int]
    lib.asdm_group_ctx.argtypes = [vp, C.c_int]
    lib.asdm_group_ctx.restype = vp
    lib.asdm_group_host_alloc.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
    lib.asdm_group_host_free.argtypes = [vp, vp]
    lib.asdm_group_host_barrier.argtypes = [vp]
    lib.asdm_group_sets_create.argtypes = [vp, C.POINTER(SetDesc), C.c_int, C.POINTER(vp)]
    lib.asdm_group_job_prepare.argtypes = [vp, C.POINTER(GJobDesc), C.POINTER(vp)]
    lib.asdm_group_plan_units.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, C.c_int, C.POINTER(C.c_int)]
    lib.asdm_group_job_units.argtypes = [vp]
    lib.asdm_group_job_stride.argtypes = [vp]
    lib.asdm_group_job_run.argtypes = [vp, vp, vp, vp]
    lib.asdm_group_job_wait.argtypes = [vp]
    lib.asdm_group_job_counts.argtypes = [vp, vp, vp]
    lib.asdm_group_job_fetch.argtypes = [vp, vp, vp, vp]
    lib.asdm_group_job_gathered.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    lib.asdm_group_job_flush.argtypes = [vp]
    lib.asdm_group_job_destroy.argtypes = [vp]
    lib.asdm_distinctive_descriptors.argtypes = [vp, C.POINTER(vp), C.c_int, C.c_int, vp, vp, vp, vp, vp]
    lib.asdm_desc_block_bytes.argtypes = [C.c_int]
    lib.asdm_desc_block_bytes.restype = C.c_size_t
    lib.asdm_set_save_blocks.argtypes = [vp, C.POINTER(vp), C.c_int, vp, C.c_size_t, vp]
    lib.asdm_set_load_block.argtypes = [vp, vp, C.c_size_t, vp, C.POINTER(Grid), C.POINTER(vp), C.POINTER(C.c_size_t)]
    lib.asdm_vocab_create.argtypes = [vp, C.POINTER(VocabDesc), C.POINTER(vp)]
    lib.asdm_vocab_release.argtypes = [vp]
    lib.asdm_vocab_transform.argtypes = [vp, vp, vp, C.c_int, vp, vp, vp, C.POINTER(i32), vp, vp]
    lib.asdm_bowdb_create.argtypes = [vp, C.POINTER(vp)]
    lib.asdm_bowdb_release.argtypes = [vp]
    lib.asdm_bowdb_add.argtypes = [vp, vp, vp, C.c_int, C.POINTER(i32)]
    lib.asdm_bowdb_erase.argtypes = [vp, C.c_int32]
    lib.asdm_bowdb_clear.argtypes = [vp]
    lib.asdm_bowdb_size.argtypes = [vp]
    lib.asdm_bowdb_query.argtypes = [vp, vp, vp, C.c_int, vp, vp, vp]
    lib.asdm_bowdb_detect_loop.argtypes = [vp, vp, vp, C.c_int, C.POINTER(BowdbGates), vp, C.c_int, C.POINTER(i32)]
    lib.asdm_bowdb_detect_reloc.argtypes = [vp, vp, vp, C.c_int, C.POINTER(BowdbGates), vp, C.c_int, C.POINTER(i32)]
    lib.asdm_bowdb_reloc_scores.argtypes = [vp, vp]
    lib.asdm_bowdb_set_reloc_scores.argtypes = [vp, vp]
    lib.asdm_bowdb_detect_loop_full.argtypes = [vp, vp, vp, C.c_int, C.POINTER(BowdbGates), vp, C.c_int, C.POINTER(i32), vp, vp, vp]
    lib.asdm_bowdb_detect_reloc_full.argtypes = [vp, vp, vp, C.c_int, C.POINTER(BowdbGates), vp, C.c_int, C.POINTER(i32), vp, vp, vp]
    _lib = lib
    return lib


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)
