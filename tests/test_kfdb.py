"""KeyFrameDatabase: BoW candidate detection, the caller that hands the loop-closure sweep (C5) its candidate pairs
(reference src/vslam/src/KeyFrameDatabase.cc:34-320 + DBoW2::L1Scoring::score, src/dbow2/DBoW2/ScoringObject.cpp:24-61).

  * CPU: the oracle restatement against the committed outputs of the reference's own KeyFrameDatabase.cc
    (tests/golden/kfdb_golden.npz, made by tests/golden/make_kfdb_golden.py) and, where oracle/_ref is present, against
    that library directly (L1 score on random vectors, extra seeds).
  * GPU: asdm_bowdb_* (k_bowdb_query + the list logic) against the same goldens and against the oracle on a larger
    database; add / erase / clear bookkeeping; error behaviour.
Tolerance: 0 -- candidate slots and their order, common-word counts, and the bits of every float score."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(HERE, "golden"))
import kfdb_cases  # noqa: E402
import make_kfdb_golden  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "kfdb_golden.npz"))


def _ref():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_py
    return ref_py if ref_py.kfdb_available() else None


def _check(name, got):
    keys = [k for k in GOLD.files if k.startswith(name + "/")]
    assert keys and len(keys) == len(got)
    for k in keys:
        g = got[k.split("/", 1)[1]]
        assert g.dtype == GOLD[k].dtype and np.array_equal(g, GOLD[k]), k


# ------------------------------------------------------------------------------------------------ CPU: oracle
@pytest.mark.parametrize("name", sorted(kfdb_cases.CASES))
def test_oracle_vs_reference_golden(pkg, oracle, name):
    db = kfdb_cases.build(pkg, name)
    _check(name, make_kfdb_golden.run(oracle, db))


def test_oracle_l1_score_vs_reference(oracle):
    ref = _ref()
    if ref is None:
        pytest.skip("oracle/_ref/libref_kfdb.so not built (needs /root/reference)")
    rng = np.random.default_rng(77)
    for _ in range(200):
        n1, n2 = rng.integers(0, 60, 2)
        w1 = np.unique(rng.integers(0, 80, n1)).astype(np.uint32)
        w2 = np.unique(rng.integers(0, 80, n2)).astype(np.uint32)
        v1, v2 = rng.standard_normal(len(w1)), rng.standard_normal(len(w2))  # signs too: fabs() matters
        a, b = oracle.bow_score_l1(w1, v1, w2, v2), ref.bow_score_l1(w1, v1, w2, v2)
        assert np.float64(a).view(np.uint64) == np.float64(b).view(np.uint64)


@pytest.mark.parametrize("seed", [11, 12, 13])
def test_oracle_vs_reference_live(pkg, oracle, seed):
    ref = _ref()
    if ref is None:
        pytest.skip("oracle/_ref/libref_kfdb.so not built (needs /root/reference)")
    db = pkg.synth.config_kfdb(n_kf=90, n_words=3000, words_per_kf=150, n_places=9, seed_extra=seed, erase_frac=0.15)
    a, b = make_kfdb_golden.run(oracle, db), make_kfdb_golden.run(ref, db)
    assert a.keys() == b.keys()
    for k in a:
        assert np.array_equal(a[k], b[k]), k


# ------------------------------------------------------------------------------------------------ GPU
class _Device:
    """make_kfdb_golden.run's `impl` interface over the CUDA library: the database is rebuilt per call sequence the same
    way the reference driver does (add every keyframe in order, erase the dead ones)."""

    def __init__(self, pkg, ctx, db):
        self.db, self.kdb = db, pkg.KeyFrameDatabase(ctx)
        o = db["bow_off"]
        for s in range(db["n_kf"]):
            assert self.kdb.add(db["bow_word"][o[s]:o[s + 1]], db["bow_val"][o[s]:o[s + 1]]) == s
        for s in np.nonzero(db["alive"] == 0)[0]:
            self.kdb.erase(int(s))

    def detect_loop_candidates(self, db, w, v, connected, min_score, only_global):
        cand = self.kdb.DetectLoopCandidates(w, v, connected, min_score, only_global, db["global_flag"], db["neigh_off"],
                                             db["neigh_slot"])
        common, first, score = self.kdb.query(w, v)
        # the reference reports mnLoopWords / mLoopScore of LISTED keyframes only, scores only above the common-word gate
        listed = (common > 0) & (np.asarray(connected) == 0) & (db["alive"] != 0)
        common = np.where(listed, common, 0).astype(np.int32)
        mx = int(common.max()) if len(common) else 0
        score = np.where(listed & (common > int(np.float32(mx) * np.float32(0.6))), score, np.float32(-1)).astype(np.float32)
        return cand, common, score

    def detect_reloc_candidates(self, db, w, v, rs, only_global):
        cand = self.kdb.DetectRelocalizationCandidates(w, v, only_global, db["global_flag"], db["neigh_off"], db["neigh_slot"])
        common, _, _ = self.kdb.query(w, v)
        rs[...] = self.kdb.reloc_scores()
        return cand, np.where(db["alive"] != 0, common, 0).astype(np.int32)


@pytest.mark.gpu
@pytest.mark.parametrize("search", [0, 1], ids=["hash", "bsearch"])
@pytest.mark.parametrize("name", sorted(kfdb_cases.CASES))
def test_gpu_vs_reference_golden(pkg, ctx, name, search):
    """both device passes: the shared-memory hash table (k_bowdb_query_hash) and the binary search (k_bowdb_query)"""
    db = kfdb_cases.build(pkg, name)
    ctx.set_option("bowdb_search", search)
    try:
        dev = _Device(pkg, ctx, db)
        _check(name, make_kfdb_golden.run(dev, db))
        assert ctx.stat("gpu_launches") == 1
        dev.kdb.release()
    finally:
        ctx.set_option("bowdb_search", 0)


@pytest.mark.gpu
def test_gpu_hash_equals_bsearch_and_fallbacks(pkg, ctx):
    """The two kernels agree bit for bit on every slot; queries the table cannot serve (word id 0xffffffff = the empty
    mark, or too many words for shared memory) take the binary search and still answer correctly."""
    rng = np.random.default_rng(77)
    kdb = pkg.KeyFrameDatabase(ctx)
    kfs = []
    for s in range(700):
        w = np.unique(rng.integers(0, 60000, int(rng.integers(1, 900)))).astype(np.uint32)
        if s % 50 == 0:
            w = np.unique(np.concatenate([w, [0xffffffff, 0xfffffffe]])).astype(np.uint32)
        v = rng.random(len(w))
        kdb.add(w, v)
        kfs.append((w, v))
    kdb.erase(13)
    queries = [kfs[5], kfs[50], kfs[699], (np.zeros(0, np.uint32), np.zeros(0)),
               (np.array([7], np.uint32), np.array([1.0])),
               (np.unique(rng.integers(0, 60000, 20000)).astype(np.uint32),) * 1]
    queries[-1] = (queries[-1][0], rng.random(len(queries[-1][0])))  # > 12800 words: no table
    for w, v in queries:
        ctx.set_option("bowdb_search", 1)
        a = kdb.query(w, v)
        ctx.set_option("bowdb_search", 0)
        b = kdb.query(w, v)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2].view(np.uint32), b[2].view(np.uint32))
    # a keyframe queried against the database finds itself with all its words
    common, first, score = kdb.query(*kfs[50])
    assert common[50] == len(kfs[50][0]) and first[50] == kfs[50][0][0] and common[13] == 0
    kdb.release()


@pytest.mark.gpu
def test_gpu_vs_oracle_large(pkg, ctx, oracle):
    db = pkg.synth.config_kfdb(n_kf=3000, n_words=200000, words_per_kf=1200, n_places=300, seed_extra=21)
    dev = _Device(pkg, ctx, db)
    a, b = make_kfdb_golden.run(oracle, db), make_kfdb_golden.run(dev, db)
    for k in a:
        assert np.array_equal(a[k], b[k]), k
    # the device pass alone: L1 score of EVERY keyframe == the oracle's merge, bit for bit
    q = db["queries"][0]
    common, first, score = dev.kdb.query(q["word"], q["val"])
    o = db["bow_off"]
    for s in range(0, db["n_kf"], 37):
        w, v = db["bow_word"][o[s]:o[s + 1]], db["bow_val"][o[s]:o[s + 1]]
        if not db["alive"][s]:
            assert common[s] == 0
            continue
        shared = np.intersect1d(w, q["word"])
        assert common[s] == len(shared) and (first[s] == shared[0] if len(shared) else first[s] == 0xFFFFFFFF)
        assert np.float32(oracle.bow_score_l1(q["word"], q["val"], w, v)).view(np.uint32) == score[s].view(np.uint32)
    dev.kdb.release()


@pytest.mark.gpu
def test_gpu_bookkeeping_and_errors(pkg, ctx):
    kdb = pkg.KeyFrameDatabase(ctx)
    assert len(kdb) == 0
    assert len(kdb.DetectLoopCandidates(np.array([1, 2], np.uint32), np.array([.5, .5]), np.zeros(0, np.uint8), 0.0)) == 0
    w, v = np.array([3, 9, 40], np.uint32), np.array([.2, .3, .5])
    assert [kdb.add(w, v), kdb.add(w[:2], v[:2]), kdb.add(w[:0], v[:0])] == [0, 1, 2]
    conn = np.zeros(3, np.uint8)
    assert list(kdb.DetectLoopCandidates(w, v, conn, 0.0)) == [0]     # 1 shares 2 of 3 words: 2 > int(3 * 0.6f) = 1 too
    assert list(kdb.query(w, v)[0]) == [3, 2, 0]
    kdb.erase(0)
    kdb.erase(0)  # erasing twice is a no-op, as in the reference
    assert list(kdb.DetectLoopCandidates(w, v, conn, 0.0)) == [1]
    assert kdb.add(w, v) == 3  # slots are never reused: the re-added keyframe stands at the end of every list
    assert list(kdb.DetectLoopCandidates(w, v, np.zeros(4, np.uint8), 0.0)) == [3]
    kdb.clear()
    assert len(kdb) == 4 and len(kdb.DetectLoopCandidates(w, v, np.zeros(4, np.uint8), 0.0)) == 0
    with pytest.raises(pkg.AsdmError):
        kdb.add(np.array([5, 5], np.uint32), np.array([.5, .5]))  # a std::map has no duplicate keys
    with pytest.raises(pkg.AsdmError):
        kdb.erase(99)
    with pytest.raises(pkg.AsdmError):  # only_global_map without flags
        kdb.DetectRelocalizationCandidates(w, v, only_global_map=True)
    kdb.release()


# ------------------------------------------------------------------------------------------------ GPU: the C++ drop-in
ADAPTOR = os.path.join(ROOT, "asd-slam_b200", "host", "libadaptor_kfdb_shim.so")


class _Adaptor:
    """host/KeyFrameDatabase_cuda.cc (our bodies of ORB_SLAM2::KeyFrameDatabase, compiled against the reference's own
    KeyFrameDatabase.h) behind the flat harness that drives the reference's KeyFrameDatabase.cc (oracle/ref_kfdb_driver.cc):
    add / erase per keyframe, the detections, and the bookkeeping members they leave on the KeyFrames."""

    def __init__(self, oracle):
        import ctypes
        self.oracle, self.lib = oracle, ctypes.CDLL(ADAPTOR)

    def detect_loop_candidates(self, db, w, v, connected, min_score, only_global):
        return self.oracle.detect_loop_candidates(db, w, v, connected, min_score, only_global, _ref=self.lib)

    def detect_reloc_candidates(self, db, w, v, rs, only_global):
        return self.oracle.detect_reloc_candidates(db, w, v, rs, only_global, _ref=self.lib)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(kfdb_cases.CASES))
def test_adaptor_vs_reference_golden(pkg, ctx, oracle, name):
    if not os.path.exists(ADAPTOR):
        pytest.skip("libadaptor_kfdb_shim.so not built (needs /root/reference at build time)")
    db = kfdb_cases.build(pkg, name)
    _check(name, make_kfdb_golden.run(_Adaptor(oracle), db))


@pytest.mark.gpu
def test_adaptor_vs_oracle_large(pkg, ctx, oracle):
    if not os.path.exists(ADAPTOR):
        pytest.skip("libadaptor_kfdb_shim.so not built (needs /root/reference at build time)")
    db = pkg.synth.config_kfdb(n_kf=1500, n_words=100000, words_per_kf=800, n_places=150, seed_extra=33, erase_frac=0.1)
    a, b = make_kfdb_golden.run(oracle, db), make_kfdb_golden.run(_Adaptor(oracle), db)
    for k in a:
        assert np.array_equal(a[k], b[k]), k
