"""asdm_stereo_row_match_static (include/asdmatch.h): the stereo descriptor stage with its device scratch in a file-scope
device array of the library -- no cudaMalloc / cudaFree / pool on the call path. The entry point has no counterpart in the
reference; the check is against the existing entry point
asdm_stereo_row_match through the C ABI, bit for bit (tolerance 0, as everywhere in tests/), on two calls in a row.

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _static_call(pkg, ctx, L, R, s, out_ptr, stream_ptr=None):
    sf = np.ascontiguousarray(s["scale_factors"], np.float32)
    ctx.check(ctx.lib.asdm_stereo_row_match_static(ctx.h, L.h, R.h, pkg._lib.ptr(sf), len(sf), int(s["n_rows"]), float(s["min_d"]),
                                                   float(s["max_d"]), C.c_void_p(out_ptr), C.c_void_p(stream_ptr)))


def _split(out_words, n):
    pad = (n + 3) & ~3
    return out_words[:n].copy(), out_words[pad:pad + n].copy().view(np.float32)


def test_static_scratch_stereo_equals_pooled_entry_point_twice(pkg, ctx):
    import torch
    s1 = pkg.synth.config3()["stereo"]              # 1500 x 1500, EuRoC-shaped
    s2 = dict(s1)                                   # second call: fewer, different left keypoints (ragged length)
    keep = np.arange(len(s1["kps_l"]))[::2][:601]
    s2["desc_l"], s2["kps_l"] = np.ascontiguousarray(s1["desc_l"][keep]), np.ascontiguousarray(s1["kps_l"][keep])
    mt = pkg.ORBmatcher(ctx, 0.7, True)
    grid = pkg.synth.config3()["f2f"]["grid"]
    pad_max = (len(s1["kps_l"]) + 3) & ~3
    h_out = ctx.pinned((2 * pad_max,), np.int32)    # page-locked host memory the kernel stores into
    d_out = torch.empty(2 * pad_max, dtype=torch.int32, device="cuda:0")  # ... and device memory, on a caller stream
    side = torch.cuda.Stream(device=0)
    sets = []
    for s in (s1, s2):
        L = pkg.DescriptorSet(ctx, s["desc_l"], s["kps_l"], grid)
        R = pkg.DescriptorSet(ctx, s["desc_r"], s["kps_r"], grid)
        sets.append((L, R))
    # first call on this device: resolves the array's address (and, with lazy module loading, loads the kernel)
    _static_call(pkg, ctx, sets[0][0], sets[0][1], s1, h_out.ctypes.data)
    ctx.synchronize()
    for (L, R), s in zip(sets, (s1, s2)):           # two calls in a row share the one static array
        n = L.n
        h_out[...] = 0x5a5a5a5a
        d_out.fill_(0x5a5a5a5a)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(0)[0]
        _static_call(pkg, ctx, L, R, s, h_out.ctypes.data)          # the context's own stream
        ctx.synchronize()
        _static_call(pkg, ctx, L, R, s, d_out.data_ptr(), side.cuda_stream)   # a stream of the caller
        side.synchronize()
        assert torch.cuda.mem_get_info(0)[0] >= free0   # the calls took nothing from the device
        best_a, dist_a = _split(h_out, n)
        best_b, dist_b = _split(d_out.cpu().numpy(), n)
        best, dist = mt.StereoRowMatch(L, R, s["scale_factors"], s["n_rows"], s["min_d"], s["max_d"])
        assert (best >= 0).sum() > n // 4           # the synthetic stereo pair really matches
        for got_b, got_d in ((best_a, dist_a), (best_b, dist_b)):
            assert np.array_equal(got_b, best)
            assert np.array_equal(got_d.view(np.uint32), dist.view(np.uint32))  # tolerance 0
    for L, R in sets:
        L.release(); R.release()


def test_static_scratch_rejects_sizes_beyond_its_maximum(pkg, ctx):
    s = pkg.synth.config3()["stereo"]
    grid = pkg.synth.config3()["f2f"]["grid"]
    n_big = 16384 + 1                               # ASDM_STATIC_MAX_LEFT + 1
    rng = np.random.default_rng(5)
    desc = rng.standard_normal((n_big, 128)).astype(np.float32)
    kps = np.zeros(n_big, pkg._lib.KP_DTYPE)
    kps["x"], kps["y"] = rng.uniform(20, 700, n_big), rng.uniform(20, 460, n_big)
    L = pkg.DescriptorSet(ctx, desc, kps, grid)
    R = pkg.DescriptorSet(ctx, s["desc_r"], s["kps_r"], grid)
    out = ctx.pinned((2 * ((n_big + 3) & ~3),), np.int32)
    with pytest.raises(pkg._lib.AsdmError, match="ASDM_STATIC_MAX_LEFT"):
        _static_call(pkg, ctx, L, R, s, out.ctypes.data)
    # the context stays usable and the pooled entry point takes the same input
    mt = pkg.ORBmatcher(ctx, 0.7, True)
    best, _ = mt.StereoRowMatch(L, R, s["scale_factors"], s["n_rows"], s["min_d"], s["max_d"])
    assert best.shape == (n_big,)
    L.release(); R.release()
