"""asd-slam_b200 -- B200-native (sm_100a) ASD descriptor matching: the data-parallel hot path of
mataiyuan/ASD-SLAM (reference src/vslam/src/ORBmatcher.cc) behind a C ABI (include/asdmatch.h).

The directory name carries a hyphen, so import it with
    importlib.import_module("asd-slam_b200")
(see tests/conftest.py).  `matcher` mirrors the reference's ORBmatcher interface; `shard` deals batched
keyframe-pair searches over the ranks of a torch.distributed group; `synth` builds the seeded inputs of the five
BASELINE configs.  Importing the package does not load the CUDA library; creating a
`Context` does, and raises when the library or a B200 is missing (no CPU fallback).
"""
from . import _lib, synth  # noqa: F401
from ._lib import (ASDM_JOB_BOW_KF_KF, ASDM_JOB_TRIANGULATION, ASDM_PROJ_BEST_ONLY, ASDM_PROJ_RATIO_LEVEL,  # noqa: F401
                   KP_DTYPE, PROJ_QUERY_DTYPE, AsdmError)
from .matcher import (ComputeDistinctiveDescriptors, Context, DescriptorSet, KeyFrameDatabase, ORBmatcher, ORBVocabulary,  # noqa: F401
                      PairJob, StreamedPairSweep, save_map_blocks)
from .group import Group, GroupJob, StreamedGroupSweep, torch_allgather  # noqa: F401
from .shard import GatheredResults, PairShards, ReplicatedDescriptors, ShardedPairJob, all_gather_results  # noqa: F401
