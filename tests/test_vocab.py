"""DBoW2 vocabulary transform: TemplatedVocabulary::transform (reference
src/dbow2/include/TemplatedVocabulary.h:1125-1192, 1216-1257) with FSift::distance (src/dbow2/DBoW2/FSift.cpp:86-101).

  * CPU: the oracle restatement (ora_voc_transform) against the reference's own code (oracle/_ref/libref_vocab.so, when
    built) and against the committed outputs of that code (tests/golden/vocab_golden.npz).
  * GPU: asdm_vocab_transform through the C ABI against the same golden outputs and the oracle, plus the FeatureVector it
    attaches on the device (checked through the searches it gates).
Tolerance: 0 -- word / node ids identical, BowVector doubles bit-identical (uint64 view).

FeatureVector node of a feature whose branch ends above level L - levelsup: the reference reads an uninitialised
NodeId there (:1148), so goldens are compared only where it is defined (vocab_cases.leaf_depth_reaches)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import vocab_cases as V  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "vocab_golden.npz"))
ALL = V.CASES + [V.TRAINED]
IDS = [c[0] for c in ALL]


def bits64(a):
    return np.asarray(a, np.float64).view(np.uint64)


def case_inputs(i):
    name, kw, n, levels = ALL[i]
    if name == V.TRAINED[0]:
        tree = {k: GOLD[f"{name}/tree/{k}"] for k in V.TREE_KEYS}
        tree.update({k: int(GOLD[f"{name}/tree/{k}"]) for k in ("k", "L", "weighting", "scoring")})
        q = V.queries(tree, n, 100)
    else:
        tree = V.synth.vocab_tree(**kw)
        q = V.queries(tree, n, i)
    assert bytes(GOLD[name + "/sha"]).hex() == V.digest(tree, q), "synthetic generator drifted: regenerate the golden file"
    return name, tree, q, levels


def check_against_gold(name, tree, lu, got):
    g = {k: GOLD[f"{name}/{lu}/{k}"] for k in ("word_id", "fv_node", "fv_order", "bow_id", "bow_val")}
    assert np.array_equal(got["word_id"], g["word_id"])
    assert np.array_equal(got["in_fv"] > 0, g["fv_node"] >= 0)
    defined = (g["fv_node"] >= 0) & V.leaf_depth_reaches(tree, g["word_id"], int(tree["L"]) - lu)
    assert np.array_equal(got["node_id"][defined], g["fv_node"][defined])
    assert defined.sum() > 0
    assert np.array_equal(got["bow_id"], g["bow_id"])
    assert np.array_equal(bits64(got["bow_val"]), bits64(g["bow_val"]))
    if defined.sum() == (g["fv_node"] >= 0).sum():  # whole FeatureVector defined: its iteration order too
        inc = np.nonzero(got["in_fv"])[0]
        order = inc[np.lexsort((inc, got["node_id"][inc]))]
        assert np.array_equal(order, g["fv_order"])


# ------------------------------------------------------------------------------------------------ CPU: oracle
@pytest.mark.parametrize("i", range(len(ALL)), ids=IDS)
def test_oracle_vs_golden(oracle, i):
    name, tree, q, levels = case_inputs(i)
    for lu in levels:
        check_against_gold(name, tree, lu, oracle.voc_transform(tree, q, lu))


@pytest.fixture(scope="module")
def ref():
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
    import ref_py
    if not ref_py.voc_available():
        pytest.skip("oracle/_ref/libref_vocab.so unavailable")
    return ref_py


def test_oracle_vs_ref_trained(oracle, ref):
    """A second tree built by the reference's create(), transformed by the reference and by the oracle."""
    rng = np.random.default_rng(123)
    train = V.synth.make_descriptors(rng, 2500)
    voc = ref.RefVocabulary.train(train, 7, 6, 3, weighting=0, scoring=0, seed=99)
    tree = voc.export()
    q = np.concatenate([V.synth.make_descriptors(rng, 300), train[:200]])
    for lu in (0, 1, 2, 3, 4):
        r, o = voc.transform(q, lu), oracle.voc_transform(tree, q, lu)
        assert np.array_equal(r["word_id"], o["word_id"])
        assert np.array_equal(r["bow_id"], o["bow_id"]) and np.array_equal(bits64(r["bow_val"]), bits64(o["bow_val"]))
        assert np.array_equal(r["fv_node"] >= 0, o["in_fv"] > 0)
        ok = (r["fv_node"] >= 0) & V.leaf_depth_reaches(tree, o["word_id"], tree["L"] - lu)
        assert np.array_equal(r["fv_node"][ok], o["node_id"][ok])
    voc.free()


def test_fsift_distance_vs_ref(oracle, ref):
    """FSift::distance accumulates fp32 products in a double: check the oracle's arithmetic through a 1-level tree whose
    best distance decides the word, on inputs with wide dynamic range (where double sums are not exact)."""
    rng = np.random.default_rng(5)
    a = (rng.standard_normal((64, 128)) * np.exp(rng.uniform(-12, 4, (64, 128)))).astype(np.float32)
    b = (rng.standard_normal((64, 128)) * np.exp(rng.uniform(-12, 4, (64, 128)))).astype(np.float32)
    for x, y in zip(a, b):
        d = ref.fsift_distance(x, y)
        acc = 0.0
        for i in range(128):
            t = np.float32(x[i] - y[i])
            acc += float(np.float32(t * t))
        assert d == acc


# ------------------------------------------------------------------------------------------------ GPU: product
@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["fast", "exact", "direct"])
@pytest.mark.parametrize("i", range(len(ALL)), ids=IDS)
def test_cuda_vs_golden(pkg, ctx, oracle, i, mode):
    """fast: child blocks staged through shared memory with cp.async (fan-out <= 32), fp32 pass + exact chain where the
    margin is too small to call; exact: the same kernel, exact chain at every level; direct: the kernel that reads the
    children straight from global memory (also the one that handles fan-outs above 32, case k40L2)."""
    name, tree, q, levels = case_inputs(i)
    voc = pkg.ORBVocabulary(ctx, tree)
    s = pkg.DescriptorSet(ctx, q)
    for lu in levels:
        ctx.set_option("voc_direct", mode == "direct")
        ctx.set_option("voc_exact", mode == "exact")
        got = voc.transform(s, lu)
        ctx.set_option("voc_direct", 0)
        ctx.set_option("voc_exact", 0)
        assert ctx.stat("gpu_launches") == 2
        check_against_gold(name, tree, lu, got)
        o = oracle.voc_transform(tree, q, lu)
        for k in ("word_id", "node_id", "in_fv", "bow_id"):
            assert np.array_equal(got[k], o[k]), k
        assert np.array_equal(bits64(got["bow_val"]), bits64(o["bow_val"]))
    s.release()
    voc.release()


@pytest.mark.gpu
@pytest.mark.parametrize("rel", [0.0, 1e-9, 1e-7, 3e-6, 1e-4])
def test_cuda_near_ties_take_the_exact_chain(pkg, ctx, oracle, rel):
    """Siblings that are copies of one centre up to a relative perturbation `rel` (0 = exact duplicates, first child
    wins): the fp32 pass cannot separate them below its margin (2^-16), so the decision must come from the reference's
    fp32-product / double-sum chain -- and the answer must still be the oracle's, word for word."""
    tree = dict(pkg.synth.vocab_tree(k=8, L=3, seed_extra=77))
    rng = np.random.default_rng(771)
    desc = tree["desc"].copy()
    off, idx = tree["child_off"], tree["child_idx"]
    for p in range(len(off) - 1):
        ch = idx[off[p]:off[p + 1]]
        if len(ch) > 1:
            jit = 1.0 + rel * rng.standard_normal((len(ch) - 1, 128))
            desc[ch[1:]] = (desc[ch[0]][None, :] * jit).astype(np.float32)
    tree["desc"] = desc
    q = pkg.synth.vocab_queries(tree, 600, seed_extra=78, sigma=0.05)
    voc = pkg.ORBVocabulary(ctx, tree)
    s = pkg.DescriptorSet(ctx, q)
    got, o = voc.transform(s, 2), oracle.voc_transform(tree, q, 2)
    n_exact = ctx.stat("voc_exact_levels")
    for k in ("word_id", "node_id", "in_fv", "bow_id"):
        assert np.array_equal(got[k], o[k]), k
    assert np.array_equal(bits64(got["bow_val"]), bits64(o["bow_val"]))
    if rel <= 1e-7:
        assert n_exact >= 600  # every feature met at least one level it could not call in fp32
    if rel == 0.0:
        assert len(np.unique(got["word_id"])) == 1  # strict <: the first of the identical siblings wins at every level
    s.release()
    voc.release()


@pytest.mark.gpu
@pytest.mark.parametrize("levelsup", [2, 4])
def test_device_featvec_gates_like_host_featvec(pkg, ctx, oracle, levelsup):
    """SearchByBoW(KF,KF) on sets whose FeatureVector was attached on the device by asdm_vocab_transform == the same
    search with the FeatureVector built on the host from the oracle's node ids == the oracle's search."""
    tree = pkg.synth.vocab_tree(k=6, L=4, seed_extra=20, stop_frac=0.1)
    c = pkg.synth.config5(n=1200, n_pairs=1, n_kf=2)
    kfs = c["kfs"]
    voc = pkg.ORBVocabulary(ctx, tree)
    res = []
    fvs = []
    for dev in (True, False):
        sets = []
        for k in range(2):
            s = pkg.DescriptorSet(ctx, kfs[k]["desc"], kfs[k]["kps"], c["grid"])
            o = oracle.voc_transform(tree, kfs[k]["desc"], levelsup)
            if dev:
                voc.transform(s, levelsup, host=False)
            else:
                fv = pkg.synth.fv_from_nodes(np.where(o["in_fv"] > 0, o["node_id"], -1))
                s.set_featvec(fv)
                fvs.append(fv)
            sets.append(s)
        mt = pkg.ORBmatcher(ctx, 0.85, True)
        res.append(mt.SearchByBoW_KF_KF(sets[0], kfs[0]["valid"], sets[1], kfs[1]["valid"]))
        for s in sets:
            s.release()
    voc.release()
    assert res[0][0] == res[1][0] and np.array_equal(res[0][1], res[1][1]) and np.array_equal(res[0][2], res[1][2])
    onm, om, od = oracle.search_by_bow_kf_kf(kfs[0]["desc"], kfs[0]["kps"], fvs[0], kfs[0]["valid"], kfs[1]["desc"],
                                             kfs[1]["kps"], fvs[1], kfs[1]["valid"], 0.85, 1)
    assert res[0][0] == onm and np.array_equal(res[0][1], om)
    assert onm > 50


@pytest.mark.gpu
def test_cuda_full_size_tree(pkg, ctx, oracle):
    """k=10, L=5 (the DBoW2 default shape: 111 111 nodes, 57 MB of centres), 4000 features, levelsup 4 as
    Frame::ComputeBoW calls it (Frame.cc:294)."""
    tree = pkg.synth.vocab_tree(k=10, L=5, seed_extra=30)
    q = pkg.synth.vocab_queries(tree, 4000, seed_extra=31, sigma=0.08)
    voc = pkg.ORBVocabulary(ctx, tree)
    s = pkg.DescriptorSet(ctx, q)
    got, o = voc.transform(s, 4), oracle.voc_transform(tree, q, 4)
    for k in ("word_id", "node_id", "in_fv", "bow_id"):
        assert np.array_equal(got[k], o[k]), k
    assert np.array_equal(bits64(got["bow_val"]), bits64(o["bow_val"]))
    assert len(np.unique(got["node_id"])) == 10 and len(got["bow_id"]) > 2000
    assert ctx.stat("voc_exact_levels") < 200  # of 20 000 decisions: the fp32 pass settles the rest
    s.release()
    voc.release()


@pytest.mark.gpu
def test_vocab_rejects_bad_trees(pkg, ctx):
    tree = pkg.synth.vocab_tree(k=3, L=2)
    bad = dict(tree)
    bad["child_idx"] = tree["child_idx"].copy()
    bad["child_idx"][0] = bad["child_idx"][1]  # a node listed twice
    with pytest.raises(pkg.AsdmError):
        pkg.ORBVocabulary(ctx, bad)
    empty = dict(tree)
    for k in ("desc", "weight", "word_id"):
        empty[k] = tree[k][:1]
    empty["child_off"], empty["child_idx"] = np.zeros(2, np.int32), np.zeros(0, np.int32)
    with pytest.raises(pkg.AsdmError):
        pkg.ORBVocabulary(ctx, empty)
    # empty frame: nothing to do, FeatureVector empty
    voc = pkg.ORBVocabulary(ctx, tree)
    s = pkg.DescriptorSet(ctx, np.zeros((0, 128), np.float32))
    got = voc.transform(s, 4)
    assert len(got["word_id"]) == 0 and len(got["bow_id"]) == 0
    s.release()
    voc.release()


ADAPTOR = os.path.join(os.path.dirname(HERE), "asd-slam_b200", "host", "libadaptor_vocab_shim.so")


@pytest.mark.gpu
@pytest.mark.parametrize("i", range(len(ALL)), ids=IDS)
def test_device_vocabulary_adaptor_vs_golden(ctx, i):
    """asd-slam_b200/host/asdm_vocabulary.h (DeviceVocabulary: the drop-in for `new ORBVocabulary()`, System.cc:114)
    behind the harness the goldens were made with: the reference's load() fills the tree, the virtual
    transform(features, BowVector&, FeatureVector&, levelsup) runs on the GPU and fills the caller's std::maps."""
    if not os.path.exists(ADAPTOR):
        pytest.skip("libadaptor_vocab_shim.so not built (needs /root/reference at build time)")
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
    import ref_py
    ref_py.use_voc(ADAPTOR)
    try:
        name, tree, q, levels = case_inputs(i)
        voc = ref_py.RefVocabulary.load(tree)
        for lu in levels:
            r = voc.transform(q, lu)
            g = {k: GOLD[f"{name}/{lu}/{k}"] for k in ("word_id", "fv_node", "fv_order", "bow_id", "bow_val")}
            assert np.array_equal(r["bow_id"], g["bow_id"]) and np.array_equal(bits64(r["bow_val"]), bits64(g["bow_val"]))
            assert np.array_equal(r["fv_node"] >= 0, g["fv_node"] >= 0)
            ok = (g["fv_node"] >= 0) & V.leaf_depth_reaches(tree, g["word_id"], int(tree["L"]) - lu)
            assert np.array_equal(r["fv_node"][ok], g["fv_node"][ok])
            if ok.sum() == (g["fv_node"] >= 0).sum():
                assert np.array_equal(r["fv_order"], g["fv_order"])
        voc.free()
    finally:
        ref_py.use_voc(None)
