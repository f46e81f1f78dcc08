"""GPU parity: libasdmatch_cuda.so (through the C ABI) against the CPU oracle on the same seeded inputs.

Bar (BASELINE.json north_star): match index lists bit-exact; distances are the reference's own fp32 values, so they
are compared bit-exactly too (tolerance stated: 0, tighter than the 1e-5 relative the north star allows).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _sets_pair(pkg, ctx, c):
    A = pkg.DescriptorSet(ctx, c["desc_a"], c["kps_a"], c["grid"], c["fv_a"])
    B = pkg.DescriptorSet(ctx, c["desc_b"], c["kps_b"], c["grid"], c["fv_b"])
    return A, B


def _eq(a, b, what):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, what
    bad = np.nonzero(a != b)[0] if a.ndim == 1 else np.argwhere(a != b)
    assert len(bad) == 0, f"{what}: {len(bad)} mismatches, first at {bad[:5].tolist()}: {a[tuple(bad[0].T)] if a.ndim > 1 else a[bad[:5]]} vs {b[tuple(bad[0].T)] if b.ndim > 1 else b[bad[:5]]}"


def _eq_dist(a, b, what):
    # bit-exact comparison of fp32 distances (tolerance 0)
    _eq(np.asarray(a, np.float32).view(np.uint32), np.asarray(b, np.float32).view(np.uint32), what)


# ------------------------------------------------------------------------------------------------ sets
def test_set_roundtrip_and_grid(pkg, ctx, oracle):
    c = pkg.synth.config1(777, 500)
    A = pkg.DescriptorSet(ctx, c["desc_a"], c["kps_a"], c["grid"], c["fv_a"])
    assert np.array_equal(A.read_desc(), c["desc_a"])
    assert np.array_equal(A.read_desc(100, 17), c["desc_a"][100:117])
    A.release()


@pytest.mark.parametrize("force_exact", [1, 0])
@pytest.mark.parametrize("n,m,nodes,ratio,ori", [(2000, 2000, 1, 0.7, 1), (1500, 2000, 1, 0.75, 0), (2000, 1777, 40, 0.7, 1),
                                                   (300, 250, 5, 0.9, 1), (64, 64, 1, 0.6, 1), (1, 1, 1, 0.7, 1)])
def test_c1_search_by_bow_kf_f(pkg, ctx, oracle, n, m, nodes, ratio, ori, force_exact):
    c = pkg.synth.config1(n, m, n_nodes=nodes)
    rng = np.random.default_rng(n * 31 + m)
    valid = (rng.random(n) < 0.9).astype(np.uint8)
    ctx.set_option("force_exact", force_exact)
    A, B = _sets_pair(pkg, ctx, c)
    mt = pkg.ORBmatcher(ctx, ratio, bool(ori))
    nm, match, dist = mt.SearchByBoW_KF_F(A, valid, B)
    onm, omatch, odist = oracle.search_by_bow_kf_f(c["desc_a"], c["kps_a"], c["fv_a"], valid, c["desc_b"], c["kps_b"],
                                                   c["fv_b"], ratio, ori)
    ctx.set_option("force_exact", 0)
    _eq(match, omatch, "match_f")
    _eq_dist(dist, odist, "dist_f")
    assert nm == onm
    if n >= 1000:
        assert nm > n // 4  # the synthetic pair really has matches
    A.release(); B.release()


@pytest.mark.parametrize("force_exact", [1, 0])
@pytest.mark.parametrize("n,nodes,ratio", [(2000, 1, 0.85), (1200, 25, 0.85), (130, 3, 0.6)])
def test_search_by_bow_kf_kf(pkg, ctx, oracle, n, nodes, ratio, force_exact):
    c = pkg.synth.config5(n=n, n_pairs=3, n_kf=4, n_nodes=nodes)
    kfs = c["kfs"]
    ctx.set_option("force_exact", force_exact)
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    mt = pkg.ORBmatcher(ctx, ratio, True)
    for a, b in [(0, 1), (2, 3), (1, 0)]:
        nm, match, dist = mt.SearchByBoW_KF_KF(sets[a], kfs[a]["valid"], sets[b], kfs[b]["valid"])
        onm, omatch, odist = oracle.search_by_bow_kf_kf(kfs[a]["desc"], kfs[a]["kps"], kfs[a]["fv"], kfs[a]["valid"],
                                                        kfs[b]["desc"], kfs[b]["kps"], kfs[b]["fv"], kfs[b]["valid"],
                                                        ratio, 1)
        _eq(match, omatch, "match12")
        _eq_dist(dist, odist, "dist12")
        assert nm == onm
    ctx.set_option("force_exact", 0)
    for s in sets:
        s.release()


@pytest.mark.parametrize("force_exact", [1, 0])
def test_exact_ties_and_duplicates(pkg, ctx, oracle, force_exact):
    """Adversarial: duplicated descriptors give exactly equal distances; scan-order tie-breaking must match
    (first wins in SearchByBoW, last wins in SearchForTriangulation)."""
    c = pkg.synth.config1(400, 400, n_nodes=3)
    rng = np.random.default_rng(5)
    db = c["desc_b"].copy()
    src = rng.integers(0, 400, 120)
    dst = rng.integers(0, 400, 120)
    db[dst] = db[src]  # exact duplicates inside B
    da = c["desc_a"].copy()
    da[rng.integers(0, 400, 60)] = db[rng.integers(0, 400, 60)]  # zero-distance pairs
    c["desc_a"], c["desc_b"] = da, db
    ctx.set_option("force_exact", force_exact)
    A, B = _sets_pair(pkg, ctx, c)
    valid = np.ones(400, np.uint8)
    for ratio in (0.7, 1.5):
        mt = pkg.ORBmatcher(ctx, ratio, True)
        nm, match, dist = mt.SearchByBoW_KF_F(A, valid, B)
        onm, omatch, odist = oracle.search_by_bow_kf_f(da, c["kps_a"], c["fv_a"], valid, db, c["kps_b"], c["fv_b"], ratio, 1)
        _eq(match, omatch, "match_f (ties)")
        _eq_dist(dist, odist, "dist_f (ties)")
        assert nm == onm
        nm, match, dist = mt.SearchByBoW_KF_KF(A, valid, B, valid)
        onm, omatch, odist = oracle.search_by_bow_kf_kf(da, c["kps_a"], c["fv_a"], valid, db, c["kps_b"], c["fv_b"], valid,
                                                        ratio, 1)
        _eq(match, omatch, "match12 (ties)")
        assert nm == onm
    # triangulation with wide-open geometry gates so that the tie rule decides
    F12 = np.array([0, 0, 0, 0, 0, 1, 0, -1, 0.0], np.float32)  # epipolar lines y2 = y1... any; gate opened below
    sf = pkg.synth.scale_factors()
    ls = np.full(8, 1e12, np.float32)
    nomp = np.zeros(400, np.uint8)
    mt = pkg.ORBmatcher(ctx, 0.6, False)
    nm, match, dist = mt.SearchForTriangulation(A, nomp, B, nomp, F12, -1e4, -1e4, sf, ls)
    onm, omatch, odist = oracle.search_for_triangulation(da, c["kps_a"], c["fv_a"], nomp, db, c["kps_b"], c["fv_b"], nomp,
                                                         F12, -1e4, -1e4, sf, ls, 0)
    _eq(match, omatch, "triangulation (ties)")
    _eq_dist(dist, odist, "triangulation dist (ties)")
    assert nm == onm and nm > 50
    ctx.set_option("force_exact", 0)
    A.release(); B.release()


@pytest.mark.parametrize("force_exact", [1, 0])
@pytest.mark.parametrize("n,pairs,nodes,ori", [(2000, 20, 1, 0), (900, 6, 30, 1)])
def test_c4_triangulation_job(pkg, ctx, oracle, n, pairs, nodes, ori, force_exact):
    c = pkg.synth.config4(n=n, n_pairs=pairs, n_nodes=nodes)
    kfs = c["kfs"]
    ctx.set_option("force_exact", force_exact)
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    job = pkg.PairJob(ctx, pkg.ASDM_JOB_TRIANGULATION, sets, [k["has_mp"] for k in kfs], c["ia"], c["ib"], 0.6, ori,
                      c["F12"], c["epipoles"], c["scale_factors"], c["level_sigma2"])
    job.run()
    nm, match, dist = job.fetch()
    assert ctx.stat("gpu_launches") >= 2
    onm, omatch, odist = oracle.batch_triangulation(kfs, c["ia"], c["ib"], c["F12"], c["epipoles"], c["scale_factors"],
                                                    c["level_sigma2"], ori, threads=oracle.max_threads())
    ctx.set_option("force_exact", 0)
    _eq(match[:, :n], omatch, "triangulation match12")
    _eq_dist(dist[:, :n], odist, "triangulation dist12")
    _eq(nm, onm, "nmatches")
    assert onm.sum() > pairs * n // 50
    # single-pair entry point agrees with the job
    mt = pkg.ORBmatcher(ctx, 0.6, bool(ori))
    p = pairs - 1
    a, b = c["ia"][p], c["ib"][p]
    nm1, m1, d1 = mt.SearchForTriangulation(sets[a], kfs[a]["has_mp"], sets[b], kfs[b]["has_mp"], c["F12"][p],
                                            c["epipoles"][p][0], c["epipoles"][p][1], c["scale_factors"],
                                            c["level_sigma2"])
    _eq(m1, omatch[p], "single-pair triangulation")
    assert nm1 == onm[p]
    job.destroy()
    for s in sets:
        s.release()


@pytest.mark.parametrize("force_exact", [1, 0])
def test_c5_loop_sweep_job(pkg, ctx, oracle, force_exact):
    n, pairs = 4000, 6
    c = pkg.synth.config5(n=n, n_pairs=pairs, n_kf=5)
    kfs = c["kfs"]
    ctx.set_option("force_exact", force_exact)
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    job = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, [k["valid"] for k in kfs], c["ia"], c["ib"], c["nnratio"],
                      c["check_ori"])
    job.run()
    nm, match, dist = job.fetch()
    ctx.set_option("force_exact", 0)
    onm, omatch, odist = oracle.batch_bow_kf_kf(kfs, c["ia"], c["ib"], c["nnratio"], c["check_ori"],
                                                threads=oracle.max_threads())
    _eq(match[:, :n], omatch, "bow match12")
    _eq_dist(dist[:, :n], odist, "bow dist12")
    _eq(nm, onm, "nmatches")
    assert onm.min() > n // 20
    # properties that hold at any size: one-to-one (vbMatched2), matched targets valid
    for p in range(pairs):
        m = match[p, :n]
        got = m[m >= 0]
        assert len(np.unique(got)) == len(got)
        assert kfs[c["ib"][p]]["valid"][got].all() and kfs[c["ia"][p]]["valid"][np.nonzero(m >= 0)[0]].all()
    job.destroy()
    for s in sets:
        s.release()


@pytest.mark.parametrize("level", [0, 1, 2])
def test_resolve_kernel_levels_agree_with_oracle(pkg, ctx, oracle, level):
    """k_resolve_bow keeps its records in shared memory under two budgets (levels 0, 1) and falls back to the global-memory
    kernel (level 2); many pairs in one launch so that CTAs of different speeds overlap (a race in the sweep flag once
    slipped through the small cases). All three must equal the oracle."""
    c = pkg.synth.config5(n=1500, n_pairs=96, n_kf=12, n_nodes=3)
    kfs = c["kfs"]
    ctx.set_option("resolve_level", level)
    try:
        sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
        job = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, [k["valid"] for k in kfs], c["ia"], c["ib"], c["nnratio"], c["check_ori"])
        for rep in range(3):
            job.run()
            nm, match, dist = job.fetch()
            if rep == 0:
                onm, omatch, odist = oracle.batch_bow_kf_kf(kfs, c["ia"], c["ib"], c["nnratio"], c["check_ori"], threads=oracle.max_threads())
            _eq(nm, onm, "nmatches")
            _eq(match[:, :1500], omatch, "match12")
            _eq_dist(dist[:, :1500], odist, "dist12")
        job.destroy()
        for s in sets:
            s.release()
    finally:
        ctx.set_option("resolve_level", 0)


def test_resolve_record_budget_escalates(pkg, ctx, oracle):
    """More queries with candidates than the 100 KB record budget holds (4000-feature keyframes, most features matched):
    the job must notice, rerun with the larger budget and still equal the oracle."""
    c = pkg.synth.config5(n=4000, n_pairs=3, n_kf=3, n_nodes=1)
    kfs = c["kfs"]
    ones = [np.ones(4000, np.uint8) for _ in kfs]
    for k, v in zip(kfs, ones):
        k["valid"] = v
    ctx.set_option("resolve_level", 0)
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    job = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, ones, c["ia"], c["ib"], 0.95, True)
    job.run()
    nm, match, dist = job.fetch()
    onm, omatch, odist = oracle.batch_bow_kf_kf(kfs, c["ia"], c["ib"], 0.95, 1, threads=oracle.max_threads())
    _eq(nm, onm, "nmatches")
    _eq(match[:, :4000], omatch, "match12")
    _eq_dist(dist[:, :4000], odist, "dist12")
    job.destroy()
    for s in sets:
        s.release()
    ctx.set_option("resolve_level", 0)


def test_small_cand_cap_retries(pkg, ctx, oracle):
    """Candidate pools start tiny and must grow transparently (no silent truncation)."""
    c = pkg.synth.config1(500, 500)
    db = c["desc_b"].copy()
    db[:300] = db[0]  # 300 identical targets: hundreds of candidates for some queries
    c["desc_a"][:50] = db[0]
    c["desc_b"] = db
    ctx.set_option("cand_cap", 1)
    A, B = _sets_pair(pkg, ctx, c)
    valid = np.ones(500, np.uint8)
    mt = pkg.ORBmatcher(ctx, 0.7, True)
    nm, match, dist = mt.SearchByBoW_KF_F(A, valid, B)
    ctx.set_option("cand_cap", 8)
    onm, omatch, odist = oracle.search_by_bow_kf_f(c["desc_a"], c["kps_a"], c["fv_a"], valid, db, c["kps_b"], c["fv_b"],
                                                   0.7, 1)
    _eq(match, omatch, "match_f after growth")
    assert nm == onm
    A.release(); B.release()


# ------------------------------------------------------------------------------- projection-gated searches
@pytest.fixture(params=[0, 1], ids=["fused", "legacy"])
def proj_path(request, ctx):
    """Both projection-search implementations must equal the oracle: k_proj_fused (default) and the three-kernel path."""
    ctx.set_option("proj_legacy", request.param)
    yield request.param
    ctx.set_option("proj_legacy", 0)


@pytest.mark.parametrize("th,nkp,nmp,inview", [(1.0, 2000, 8000, 2000), (5.0, 2000, 8000, 3000), (1.0, 50, 200, 100),
                                               (3.0, 1, 5, 5)])
def test_c2_search_by_projection_local_map(pkg, ctx, oracle, th, nkp, nmp, inview, proj_path):
    c = pkg.synth.config2(n_kp=nkp, n_mp=nmp, n_in_view=inview, th=th)
    F = pkg.DescriptorSet(ctx, c["desc_t"], c["kps_t"], c["grid"])
    mt = pkg.ORBmatcher(ctx, c["nnratio"], True)
    q = c["queries"]
    nm, assign, dist = mt.SearchByProjection_LocalMap(F, q["valid"], q["u"], q["v"], c["level"], c["view_cos"], q["blocks"],
                                                      c["desc_q"], pkg.synth.scale_factors(), th=th,
                                                      blocked0=c["blocked0"])
    if proj_path == 0:  # fused: one launch per call (+ the grid bucket on the first search of a set)
        assert ctx.stat("gpu_launches") <= 2
    onm, oassign, odist = oracle.search_by_projection(c["desc_t"], c["kps_t"], c["grid"], q, c["desc_q"], c["blocked0"],
                                                      c["mode"], c["th_accept"], c["best_init"], c["nnratio"],
                                                      c["check_ori"], c["count_twice"])
    _eq(assign, oassign, "assign")
    _eq_dist(dist, odist, "dist")
    assert nm == onm
    if nkp >= 2000:
        assert (oassign >= 0).sum() > 500
    F.release()


def test_c3_stereo_and_frame_to_frame(pkg, ctx, oracle, proj_path):
    c = pkg.synth.config3()
    s, f = c["stereo"], c["f2f"]
    L = pkg.DescriptorSet(ctx, s["desc_l"], s["kps_l"], f["grid"])
    R = pkg.DescriptorSet(ctx, s["desc_r"], s["kps_r"], f["grid"])
    mt = pkg.ORBmatcher(ctx, 0.8, True)
    best, dist = mt.StereoRowMatch(L, R, s["scale_factors"], s["n_rows"], s["min_d"], s["max_d"])
    obest, odist = oracle.stereo_row_match(s["desc_l"], s["kps_l"], s["desc_r"], s["kps_r"], s["scale_factors"],
                                           s["n_rows"], s["min_d"], s["max_d"])
    _eq(best, obest, "stereo best_r")
    _eq_dist(dist, odist, "stereo dist")
    assert (obest >= 0).sum() > 200
    # frame-to-frame SearchByProjection(Cur, Last, th=15, mono): queries are rows of the last frame's set
    Cur = pkg.DescriptorSet(ctx, f["desc_t"], f["kps_t"], f["grid"])
    nm, assign, dist = mt.SearchByProjection(Cur, f["queries"], q_set=L, blocked0=f["blocked0"],
                                             mode=pkg.ASDM_PROJ_BEST_ONLY, th_accept=1.5, best_init=100.0)
    onm, oassign, odist = oracle.search_by_projection(f["desc_t"], f["kps_t"], f["grid"], f["queries"], f["desc_q"],
                                                      f["blocked0"], f["mode"], f["th_accept"], f["best_init"],
                                                      f["nnratio"], f["check_ori"], f["count_twice"])
    _eq(assign, oassign, "f2f assign")
    _eq_dist(dist, odist, "f2f dist")
    assert nm == onm and (oassign >= 0).sum() > 200
    # same through host descriptors and an explicit row map
    rows = np.arange(len(f["queries"]), dtype=np.int32)
    nm2, assign2, _ = mt.SearchByProjection(Cur, f["queries"], q_set=L, q_rows=rows, blocked0=f["blocked0"],
                                            mode=pkg.ASDM_PROJ_BEST_ONLY, th_accept=1.5, best_init=100.0)
    _eq(assign2, oassign, "f2f assign (row map)")
    for x in (L, R, Cur):
        x.release()


def test_projection_entry_pool_overflow_is_retried(pkg, ctx, oracle, proj_path):
    """The candidate pool starts far too small: the search must notice, grow it and repeat -- never walk the hole an
    overflowing query leaves in the pool (kernels_proj.cuh:112) and never truncate silently."""
    f = pkg.synth.config3()["f2f"]
    s = pkg.synth.config3()["stereo"]
    L = pkg.DescriptorSet(ctx, s["desc_l"], s["kps_l"], f["grid"])
    Cur = pkg.DescriptorSet(ctx, f["desc_t"], f["kps_t"], f["grid"])
    onm, oassign, odist = oracle.search_by_projection(f["desc_t"], f["kps_t"], f["grid"], f["queries"], f["desc_q"],
                                                      f["blocked0"], f["mode"], f["th_accept"], f["best_init"],
                                                      f["nnratio"], f["check_ori"], f["count_twice"])
    mt = pkg.ORBmatcher(ctx, 0.8, True)
    ctx.set_option("proj_ent_cap", 64)
    try:
        for rep in range(2):  # the second call starts from the grown pool
            nm, assign, dist = mt.SearchByProjection(Cur, f["queries"], q_set=L, blocked0=f["blocked0"],
                                                     mode=pkg.ASDM_PROJ_BEST_ONLY, th_accept=1.5, best_init=100.0)
            if rep == 0:
                assert ctx.stat("retries") >= 1
            _eq(assign, oassign, "f2f assign after pool growth")
            _eq_dist(dist, odist, "f2f dist after pool growth")
            assert nm == onm
    finally:
        ctx.set_option("proj_ent_cap", 1 << 16)
    L.release(); Cur.release()


def test_projection_query_descriptor_sources(pkg, ctx, oracle):
    """Host descriptors are read in place when they are page-locked (asdm_host_alloc) and staged when pageable; both
    equal the oracle, and so does the per-query view of the same search."""
    c = pkg.synth.config2(n_kp=700, n_mp=2500, n_in_view=900, th=1.0)
    F = pkg.DescriptorSet(ctx, c["desc_t"], c["kps_t"], c["grid"])
    mt = pkg.ORBmatcher(ctx, c["nnratio"], True)
    q = c["queries"]
    onm, oassign, odist = oracle.search_by_projection(c["desc_t"], c["kps_t"], c["grid"], q, c["desc_q"], c["blocked0"],
                                                      c["mode"], c["th_accept"], c["best_init"], c["nnratio"],
                                                      c["check_ori"], c["count_twice"])
    pinned = ctx.pinned(c["desc_q"].shape, np.float32)
    pinned[...] = c["desc_q"]
    for dq in (c["desc_q"], pinned):
        nm, assign, dist = mt.SearchByProjection_LocalMap(F, q["valid"], q["u"], q["v"], c["level"], c["view_cos"], q["blocks"],
                                                          dq, pkg.synth.scale_factors(), th=1.0, blocked0=c["blocked0"])
        _eq(assign, oassign, "assign")
        _eq_dist(dist, odist, "dist")
        assert nm == onm
    F.release()


def test_empty_jobs_settle_and_fetch(pkg, ctx, oracle):
    """A job without pairs (a rank that was dealt no unit) must settle and fetch without touching the device flags
    (asdmatch.cu:1163)."""
    c = pkg.synth.config5(n=300, n_pairs=2, n_kf=3, n_nodes=4)
    kfs = c["kfs"]
    sets = [pkg.DescriptorSet(ctx, k["desc"], k["kps"], c["grid"], k["fv"]) for k in kfs]
    masks = [k["valid"] for k in kfs]
    # poison the block cache so that recycled scratch is not accidentally zero
    big = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, masks, c["ia"], c["ib"], 0.85, True)
    big.run(); big.fetch(); big.destroy()
    job = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, masks, np.zeros(0, np.int32), np.zeros(0, np.int32), 0.85, True)
    job.run()
    job.settle()
    nm, match, dist = job.fetch()
    assert nm.shape == (0,) and match.shape[0] == 0
    job.destroy()
    # world larger than the number of pairs: ranks 2 and 3 own nothing
    for rank in (0, 3):
        sj = pkg.ShardedPairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, masks, c["ia"], c["ib"], 0.85, True, rank=rank, world=4)
        sj.run()
        sj.settle()
        assert sj.n_local == (1 if rank == 0 else 0)
        sj.destroy()
    for s in sets:
        s.release()


@pytest.mark.parametrize("serial", [0, 1])
@pytest.mark.parametrize("n,window", [(4000, 100), (600, 40), (30, 500), (900, 12)])
def test_search_for_initialization(pkg, ctx, oracle, n, window, serial):
    """serial=0: Jacobi sweeps to the fixed point of the reference's serial replay (k_init_sweep); serial=1: the serial
    replay kernel itself (the fallback). Both must give the reference's lists."""
    c = pkg.synth.config_init(n=n, window=window)
    F1 = pkg.DescriptorSet(ctx, c["desc1"], c["kps1"], c["grid2"])
    F2 = pkg.DescriptorSet(ctx, c["desc2"], c["kps2"], c["grid2"])
    mt = pkg.ORBmatcher(ctx, c["nnratio"], True)
    ctx.set_option("init_serial", serial)
    nm, match, dist, prev = mt.SearchForInitialization(F1, F2, c["prev_matched"], window)
    ctx.set_option("init_serial", 0)
    if n >= 600:
        assert ctx.stat("init_sweeps") == 1 - serial
    onm, omatch, odist, oprev = oracle.search_for_initialization(c["desc1"], c["kps1"], c["desc2"], c["kps2"], c["grid2"],
                                                                 c["prev_matched"], window, c["nnratio"], 1)
    _eq(match, omatch, "matches12")
    _eq_dist(dist, odist, "dist12")
    assert nm == onm
    assert np.array_equal(prev, oprev)
    if n >= 4000:
        assert onm > 100
    F1.release(); F2.release()


def test_init_stealing(pkg, ctx, oracle):
    """Later queries steal a keypoint from an earlier owner (ORBmatcher.cc:474-478): make many F1 keypoints
    compete for the same F2 keypoints with decreasing distances."""
    rng = np.random.default_rng(11)
    S = pkg.synth
    n = 300
    W, H = S.KITTI["width"], S.KITTI["height"]
    d2 = S.make_descriptors(rng, n)
    k2 = S.make_keypoints(rng, n, W, H)
    k2["octave"] = 0
    d1 = np.empty_like(d2)
    k1 = S.make_keypoints(rng, n, W, H)
    k1["octave"] = (rng.random(n) < 0.2).astype(np.int32)
    tgt = rng.integers(0, 40, n)  # 300 queries fight over 40 targets
    for i in range(n):
        noise = rng.standard_normal(128).astype(np.float32) * np.float32(0.05 * (1.0 - i / (2 * n)))
        v = d2[tgt[i]] + noise
        d1[i] = v / np.sqrt((v * v).sum(dtype=np.float32))
        k1["x"][i], k1["y"][i] = k2["x"][tgt[i]] + rng.normal(0, 5), k2["y"][tgt[i]] + rng.normal(0, 5)
    g = S.image_grid(W, H)
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    F1 = pkg.DescriptorSet(ctx, d1, k1, g)
    F2 = pkg.DescriptorSet(ctx, d2, k2, g)
    mt = pkg.ORBmatcher(ctx, 0.9, True)
    nm, match, dist, pv = mt.SearchForInitialization(F1, F2, prev, 100)
    onm, omatch, odist, opv = oracle.search_for_initialization(d1, k1, d2, k2, g, prev, 100, 0.9, 1)
    _eq(match, omatch, "matches12 (steal)")
    assert nm == onm and np.array_equal(pv, opv)
    F1.release(); F2.release()


def test_errors_do_not_throw(pkg, ctx):
    """C ABI error behaviour: negative codes + message, never an exception from native code."""
    import ctypes as C
    lib = ctx.lib
    h = C.c_void_p()
    assert lib.asdm_set_create(ctx.h, None, 5, 0, None, None, C.byref(h)) == -1
    c = pkg.synth.config1(10, 10)
    A = pkg.DescriptorSet(ctx, c["desc_a"], c["kps_a"], c["grid"])  # no FeatureVector
    mt = pkg.ORBmatcher(ctx)
    with pytest.raises(pkg.AsdmError):
        mt.SearchByBoW_KF_KF(A, np.ones(10, np.uint8), A, np.ones(10, np.uint8))
    assert b"FeatureVector" in lib.asdm_last_error(ctx.h)
    A.release()


# ------------------------------------------------------------------ MapPoint::ComputeDistinctiveDescriptors
@pytest.mark.parametrize("n_kf,n,n_points,max_obs", [(20, 2000, 6000, 12), (48, 300, 200, 48), (2, 10, 5, 2)])
def test_compute_distinctive_descriptors(pkg, ctx, oracle, n_kf, n, n_points, max_obs):
    c = pkg.synth.config_distinctive(n_kf=n_kf, n=n, n_points=n_points, max_obs=max_obs, dup_frac=0.1)
    sets = [pkg.DescriptorSet(ctx, d) for d in c["descs"]]
    best, med = pkg.ComputeDistinctiveDescriptors(ctx, sets, c["obs_off"], c["obs_kf"], c["obs_row"])
    assert ctx.stat("gpu_launches") >= 1
    obest, omed = oracle.distinctive(c["descs"], c["obs_off"], c["obs_kf"], c["obs_row"])
    _eq(best, obest, "best observation")
    has = obest >= 0
    _eq_dist(med[has], omed[has], "best median")
    if max_obs > 32:
        assert (np.diff(c["obs_off"]) > 32).sum() > 10  # the large-point kernel ran
    # empty batch and points without observations
    b0, _ = pkg.ComputeDistinctiveDescriptors(ctx, sets, [0, 0, 0], [], [])
    assert list(b0) == [-1, -1]
    with pytest.raises(pkg.AsdmError):
        pkg.ComputeDistinctiveDescriptors(ctx, sets, [0, 1], [len(sets)], [0])
    for s in sets:
        s.release()


@pytest.mark.parametrize("bad", [False, True])
def test_streamed_sweep_and_speculative_tensor_route(pkg, ctx, oracle, bad):
    """StreamedPairSweep: keyframes uploaded asynchronously, pairs searched in readiness waves while later keyframes are
    still crossing PCIe. The jobs take the tensor-core route before the range flags of their sets are back; with
    `bad` one keyframe holds rows outside the shortlist's proven norm range (x 3000), so the speculation fails and the
    settle step must rerun that wave on the exact route. Either way: the oracle's lists, bit for bit."""
    n, pairs, n_kf = 1500, 40, 9
    c = pkg.synth.config5(n=n, n_pairs=pairs, n_kf=n_kf, n_nodes=6)
    kfs = c["kfs"]
    if bad:
        kfs[3]["desc"][::7] *= np.float32(3000.0)
    host = [ctx.pinned((n, 128), np.float32) for _ in kfs]
    for h, k in zip(host, kfs):
        h[...] = k["desc"]
    sweep = pkg.StreamedPairSweep(ctx, pkg.ASDM_JOB_BOW_KF_KF, n_kf, c["ia"], c["ib"], c["nnratio"], c["check_ori"],
                                  bounds=[3, 5, 7, 9])
    match, dist, nm = ctx.pinned((pairs, n), np.int32), ctx.pinned((pairs, n), np.float32), ctx.pinned((pairs,), np.int32)
    retries = []

    def make(lo, hi):
        if lo == 0:  # first wave one set at a time, the others as batches: both creation paths feed the same searches
            return [pkg.DescriptorSet(ctx, host[i], kfs[i]["kps"], c["grid"], kfs[i]["fv"], async_upload=True) for i in range(lo, hi)]
        return pkg.DescriptorSet.create_batch(ctx, host[lo:hi], [k["kps"] for k in kfs[lo:hi]], c["grid"],
                                              [k["fv"] for k in kfs[lo:hi]])
    sets = sweep.run(make, [k["valid"] for k in kfs], match, dist, nm)
    retries.append(ctx.stat("retries"))
    onm, omatch, odist = oracle.batch_bow_kf_kf(kfs, c["ia"], c["ib"], c["nnratio"], c["check_ori"], threads=oracle.max_threads())
    o = sweep.order
    assert sorted(o.tolist()) == list(range(pairs))
    _eq(match, omatch[o], "streamed match12")
    _eq_dist(dist, odist[o], "streamed dist12")
    _eq(nm, onm[o], "streamed nmatches")
    assert onm.max() > n // 20
    # the same pairs as one job on the (now flag-resolved) sets: the exact route is chosen up front when `bad`
    job = pkg.PairJob(ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, [k["valid"] for k in kfs], c["ia"], c["ib"], c["nnratio"], c["check_ori"])
    job.run()
    nm2, match2, _ = job.fetch()
    assert ctx.stat("tc_used") == (0 if bad else 1)
    _eq(match2[:, :n], omatch, "one-job match12")
    job.destroy()
    for s in sets:
        s.release()


def test_batch_created_sets_in_projection_search(pkg, ctx, oracle):
    """asdm_set_create_batch (keypoints + FeatureVector in one copy, one pack launch, grid built by the first projection
    search): the sets behave exactly like individually created ones -- rows, padding, grid scan order."""
    cs = [pkg.synth.config2(n_kp=n, n_mp=4 * n, n_in_view=n, th=2.0, seed_extra=e) for e, n in enumerate((1999, 300, 1))]
    descs = [np.ascontiguousarray(c["desc_t"], np.float32) for c in cs]
    sets = pkg.DescriptorSet.create_batch(ctx, descs, [c["kps_t"] for c in cs], cs[0]["grid"])
    for c, F, d in zip(cs, sets, descs):
        assert np.array_equal(F.read_desc().view(np.uint32), d.view(np.uint32))
        q = c["queries"]
        mt = pkg.ORBmatcher(ctx, c["nnratio"], True)
        for _ in range(2):  # second call: grid already built
            nm, assign, dist = mt.SearchByProjection_LocalMap(F, q["valid"], q["u"], q["v"], c["level"], c["view_cos"],
                                                              q["blocks"], c["desc_q"], pkg.synth.scale_factors(), th=2.0,
                                                              blocked0=c["blocked0"])
            onm, oassign, odist = oracle.search_by_projection(c["desc_t"], c["kps_t"], c["grid"], q, c["desc_q"], c["blocked0"],
                                                              c["mode"], c["th_accept"], c["best_init"], c["nnratio"],
                                                              c["check_ori"], c["count_twice"])
            _eq(assign, oassign, "assign")
            _eq_dist(dist, odist, "dist")
            assert nm == onm
    for F in sets:
        F.release()


def test_single_pair_inputs_survive_flag_resolution(pkg, ctx, oracle):
    """A single-pair search whose sets still have their range flags in flight resolves them INSIDE job_prepare, with
    the job's input block already staged in the page-locked arena at offset 0 (the previous call rewound it). Nothing
    staged or written by the host afterwards may land on that block before the device has pulled it (seen once: the
    result block of the call overlapped the masks and the tensor maps)."""
    c = pkg.synth.config1(4000, 4000, n_nodes=6)
    valid = np.ones(4000, np.uint8)
    mt = pkg.ORBmatcher(ctx, 0.8, True)
    onm, omatch, odist = oracle.search_by_bow_kf_f(c["desc_a"], c["kps_a"], c["fv_a"], valid, c["desc_b"], c["kps_b"],
                                                   c["fv_b"], 0.8, 1)
    old = pkg.DescriptorSet(ctx, c["desc_a"][:64])
    for _ in range(6):
        A, B = _sets_pair(pkg, ctx, c)           # flags in flight
        old.read_desc(0, 8)                      # waits for the stream and rewinds the arena
        nm, match, dist = mt.SearchByBoW_KF_F(A, valid, B)
        _eq(match, omatch, "match after flag resolution")
        _eq_dist(dist, odist, "dist after flag resolution")
        assert nm == onm
        A.release(); B.release()
    old.release()
