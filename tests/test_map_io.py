"""Map save / load of descriptors: the per-frame descriptor block of the reference's binary map file
`chamo.map` -- System::saveToVisualMap (reference src/vslam/src/System.cc:333-343) + vm::save_visual_map
(src/visual_map/src/visual_map_seri.cc:134-146) and vm::loader_visual_map (:272-285).

  * CPU: the oracle's writer / reader against the reference's own loops (oracle/_ref/libref_map.so, extracted by line
    range at build time) and against a committed digest of the reference's bytes.
  * GPU: asdm_set_save_blocks == those bytes; asdm_set_load_block -> the same resident rows; a search on loaded sets ==
    the search on uploaded sets; encode -> decode round trip at C5 keyframe size.
Tolerance: 0 (bytes)."""
import hashlib
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

# sha256 of the bytes the REFERENCE writes (ref_py.map_write_block) for seeded_desc(n); regenerate with
#   python tests/test_map_io.py
GOLD = {
    0: "b1b0bee5378188f5250138bcce25855f2617f9c55b20b9628e13d367c47404a9",
    1: "efe63a94005d0414caa5a95f8aeb302cf13f9af4742e583c5738051620101131",
    37: "c9f3f851425d5b7d17eb8ef1ecb8277852a37677c1d2f3743e4e193719ec0f82",
    2000: "ef74676ffaa93e548a344beb9ba474ed31176d25c36bc0824a46d3f79d03a829",
}


def seeded_desc(n):
    rng = np.random.default_rng(0xC4A30 + n)
    d = rng.standard_normal((n, 128)).astype(np.float32)
    if n > 3:
        d[1, 5] = -0.0
        d[2, 7] = np.float32(1e-42)  # denormal: the file stores raw bits
        d[3, 9] = np.inf
    return d


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _ref():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_py
    return ref_py if ref_py.map_available() else None


@pytest.mark.parametrize("n", sorted(GOLD))
def test_oracle_block_vs_golden_digest(oracle, n):
    block = oracle.map_write_block(seeded_desc(n))
    assert len(block) == 8 + n * 512
    assert np.array_equal(block[:8].view(np.int32), [128, n])
    assert sha(block) == GOLD[n]
    rows = oracle.map_read_block(block)
    assert rows.shape == (n, 128) and np.array_equal(rows.view(np.uint32), seeded_desc(n).view(np.uint32))
    if n:
        assert oracle.map_read_block(block[:-4]) is None  # truncated


@pytest.mark.parametrize("n", sorted(GOLD))
def test_oracle_block_vs_reference(oracle, n, tmp_path):
    ref = _ref()
    if ref is None:
        pytest.skip("oracle/_ref/libref_map.so unavailable")
    d = seeded_desc(n)
    path = str(tmp_path / "blk.bin")
    theirs = ref.map_write_block(d, path)
    assert np.array_equal(theirs, oracle.map_write_block(d)) and sha(theirs) == GOLD[n]
    rows, dims = ref.map_read_block(path, n)
    assert list(dims) == [128, n] and np.array_equal(rows.view(np.uint32), d.view(np.uint32))


# ------------------------------------------------------------------------------------------------ GPU: product
@pytest.mark.gpu
def test_save_blocks_are_the_reference_bytes(pkg, ctx, oracle):
    ns = sorted(GOLD)
    sets = [pkg.DescriptorSet(ctx, seeded_desc(n)) for n in ns]
    out, offs = pkg.save_map_blocks(ctx, sets)
    for k, n in enumerate(ns):
        blk = out[offs[k]:offs[k + 1]]
        assert sha(blk) == GOLD[n]
        assert np.array_equal(blk, oracle.map_write_block(seeded_desc(n)))
    # too small a buffer is refused, nothing is written past it
    small = np.zeros(16, np.uint8)
    import ctypes as C
    hs = (C.c_void_p * len(sets))(*[s.h for s in sets])
    assert ctx.lib.asdm_set_save_blocks(ctx.h, hs, len(sets), small.ctypes.data_as(C.c_void_p), 16, None) == -1
    for s in sets:
        s.release()


@pytest.mark.gpu
def test_load_block_round_trip_and_errors(pkg, ctx, oracle):
    blocks = [oracle.map_write_block(seeded_desc(n)) for n in sorted(GOLD)]
    stream = np.concatenate(blocks)  # frames follow each other in the file
    pos = 0
    for n in sorted(GOLD):
        s, used = pkg.DescriptorSet.from_map_block(ctx, stream[pos:])
        assert used == 8 + n * 512 and s.n == n
        if n:
            assert np.array_equal(s.read_desc().view(np.uint32), seeded_desc(n).view(np.uint32))
        pos += used
        s.release()
    assert pos == len(stream)
    with pytest.raises(pkg.AsdmError):  # truncated payload
        pkg.DescriptorSet.from_map_block(ctx, blocks[-1][:-4])
    bad = blocks[-1].copy()
    bad[:4].view(np.int32)[0] = 32  # a binary-ORB map (32-byte descriptors) is not an ASD map
    with pytest.raises(pkg.AsdmError):
        pkg.DescriptorSet.from_map_block(ctx, bad)


@pytest.mark.gpu
def test_search_on_loaded_sets_equals_search_on_uploaded_sets(pkg, ctx, oracle):
    c = pkg.synth.config5(n=4000, n_pairs=1, n_kf=2, n_nodes=8)
    kfs = c["kfs"]
    up = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    saved, offs = pkg.save_map_blocks(ctx, up)
    ld = []
    for k, kf in enumerate(kfs):
        s, _ = pkg.DescriptorSet.from_map_block(ctx, saved[offs[k]:offs[k + 1]], kf["kps"], c["grid"])
        s.set_featvec(kf["fv"])
        ld.append(s)
    mt = pkg.ORBmatcher(ctx, 0.85, True)
    a = mt.SearchByBoW_KF_KF(up[0], kfs[0]["valid"], up[1], kfs[1]["valid"])
    b = mt.SearchByBoW_KF_KF(ld[0], kfs[0]["valid"], ld[1], kfs[1]["valid"])
    assert a[0] == b[0] > 100 and np.array_equal(a[1], b[1]) and np.array_equal(a[2].view(np.uint32), b[2].view(np.uint32))
    for s in up + ld:
        s.release()


if __name__ == "__main__":
    ref = _ref()
    for n in sorted(GOLD):
        print(f'    {n}: "{sha(ref.map_write_block(seeded_desc(n), "/tmp/_blk.bin"))}",')
