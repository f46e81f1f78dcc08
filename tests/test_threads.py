"""Three host threads, one context each on the same GPU (asdm_fork), sharing the descriptor sets: what upstream ORB-SLAM2
does with its tracking, local-mapping and loop-closing threads (a single context serialised them on its
mutex, blocking waits included). Every result of every thread must equal the oracle's, whatever the interleaving."""
import threading
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_forked_contexts_search_shared_sets_concurrently(pkg, ctx, oracle):
    root = ctx
    loop_ctx, track_ctx = root.fork(), root.fork()
    try:
        # keyframes created by the "mapping" thread's context (root) and searched by the others
        c5 = pkg.synth.config5(n=2000, n_pairs=48, n_kf=8, n_nodes=3)
        kfs = c5["kfs"]
        sets = [pkg.DescriptorSet(root, k["desc"], k["kps"], c5["grid"], k["fv"]) for k in kfs]
        masks = [k["valid"] for k in kfs]
        ref5 = oracle.batch_bow_kf_kf(kfs, c5["ia"], c5["ib"], c5["nnratio"], c5["check_ori"], threads=oracle.max_threads())
        c2 = pkg.synth.config2(n_kp=1500, n_mp=5000, n_in_view=1500, th=1.0)
        F = pkg.DescriptorSet(root, c2["desc_t"], c2["kps_t"], c2["grid"])  # grid bucket built by whoever searches first
        pool = pkg.DescriptorSet(root, c2["desc_q"])
        q = c2["queries"]
        ref2 = oracle.search_by_projection(c2["desc_t"], c2["kps_t"], c2["grid"], q, c2["desc_q"], c2["blocked0"], c2["mode"],
                                           c2["th_accept"], c2["best_init"], c2["nnratio"], c2["check_ori"], c2["count_twice"])
        c1 = pkg.synth.config1(900, 1000, n_nodes=4)
        ref1 = oracle.search_by_bow_kf_f(c1["desc_a"], c1["kps_a"], c1["fv_a"], c1["valid_a"], c1["desc_b"], c1["kps_b"], c1["fv_b"], 0.7, 1)
        errors, lat = [], []
        stop = threading.Event()

        def loop_closing():  # long batched jobs on its own context
            try:
                for _ in range(12):
                    job = pkg.PairJob(loop_ctx, pkg.ASDM_JOB_BOW_KF_KF, sets, masks, c5["ia"], c5["ib"], c5["nnratio"], c5["check_ori"])
                    job.run()
                    nm, match, dist = job.fetch()
                    job.destroy()
                    if not (np.array_equal(nm, ref5[0]) and np.array_equal(match[:, :2000], ref5[1]) and
                            np.array_equal(dist[:, :2000].view(np.uint32), ref5[2].view(np.uint32))):
                        errors.append("loop closing: sweep differs from the oracle")
            except Exception as e:  # pragma: no cover
                errors.append(f"loop closing: {e!r}")
            finally:
                stop.set()

        def tracking():  # short projection searches on its own context, against sets the root created
            try:
                mt = pkg.ORBmatcher(track_ctx, c2["nnratio"], True)
                while not stop.is_set():
                    t = time.perf_counter()
                    nm, assign, dist = mt.SearchByProjection(F, q, q_set=pool, blocked0=c2["blocked0"], mode=pkg.ASDM_PROJ_RATIO_LEVEL,
                                                             th_accept=1.5, best_init=256.0, count_twice=True, check_ori=False)
                    lat.append(time.perf_counter() - t)
                    if not (nm == ref2[0] and np.array_equal(assign, ref2[1]) and np.array_equal(dist.view(np.uint32), ref2[2].view(np.uint32))):
                        errors.append("tracking: projection search differs from the oracle")
                        break
            except Exception as e:  # pragma: no cover
                errors.append(f"tracking: {e!r}")

        def mapping():  # creates and releases sets on the root context while the others search
            try:
                mt = pkg.ORBmatcher(root, 0.7, True)
                while not stop.is_set():
                    A = pkg.DescriptorSet(root, c1["desc_a"], c1["kps_a"], c1["grid"], c1["fv_a"])
                    B = pkg.DescriptorSet(root, c1["desc_b"], c1["kps_b"], c1["grid"], c1["fv_b"])
                    nm, match, dist = mt.SearchByBoW_KF_F(A, c1["valid_a"], B)
                    A.release(); B.release()
                    if not (nm == ref1[0] and np.array_equal(match, ref1[1]) and np.array_equal(dist.view(np.uint32), ref1[2].view(np.uint32))):
                        errors.append("mapping: SearchByBoW differs from the oracle")
                        break
            except Exception as e:  # pragma: no cover
                errors.append(f"mapping: {e!r}")

        threads = [threading.Thread(target=f) for f in (loop_closing, tracking, mapping)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=300)
        assert not errors, errors
        assert len(lat) > 20, "the tracking thread hardly ran: it was serialised behind the other threads"
        print(f"tracking searches while the other threads ran: {len(lat)}, median {1e6 * np.median(lat):.0f} us")
        for s in sets + [F, pool]:
            s.release()
    finally:
        loop_ctx.close()
        track_ctx.close()
