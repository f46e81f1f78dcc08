#!/usr/bin/env python
"""Generates tests/golden/ref_golden.npz by running the REFERENCE's own ORBmatcher.cc (oracle/_ref, compiled from
/root/reference by oracle/Makefile `ref`) on the cases of cases.py.  Run in the container that mounts /root/reference:

    python tests/golden/make_golden.py

The stored outputs are index lists / counts (and the distances DescriptorDistance returns for the matched pairs)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle"))
import cases as C  # noqa: E402
import ref_py as ref  # noqa: E402


def pair_dists(da, db, ia, ib):
    return np.array([ref.descriptor_distance(da[i], db[j]) for i, j in zip(ia, ib)], np.float32)


def main():
    S = C.synth()
    out = {}
    for name, kind, a in C.cases():
        out[name + "/sha"] = np.frombuffer(bytes.fromhex(C.digest(a)), np.uint8)
        if kind == "bow":
            c = a["c"]
            nm, m = ref.search_by_bow_kf_f(c["desc_a"], c["kps_a"], c["fv_a"], c["valid_a"], c["desc_b"], c["kps_b"], c["fv_b"], a["ratio"], a["ori"])
            j = np.nonzero(m >= 0)[0]
            out[name + "/kf_f_nm"], out[name + "/kf_f_match"] = np.int32(nm), m
            out[name + "/kf_f_dist"] = pair_dists(c["desc_a"], c["desc_b"], m[j], j)
            nm, m = ref.search_by_bow_kf_kf(c["desc_a"], c["kps_a"], c["fv_a"], c["valid_a"], c["desc_b"], c["kps_b"], c["fv_b"], c["valid_b"], a["ratio"], a["ori"])
            i = np.nonzero(m >= 0)[0]
            out[name + "/kf_kf_nm"], out[name + "/kf_kf_match"] = np.int32(nm), m
            out[name + "/kf_kf_dist"] = pair_dists(c["desc_a"], c["desc_b"], i, m[i])
        elif kind == "tri":
            c = a["c"]
            kfs = c["kfs"]
            ms, nms = [], []
            for p in range(len(c["ia"])):
                x, y = kfs[c["ia"][p]], kfs[c["ib"][p]]
                nm, m = ref.search_for_triangulation(x["desc"], x["kps"], x["fv"], x["has_mp"], y["desc"], y["kps"], y["fv"], y["has_mp"],
                                                     c["F12"][p], c["epipoles"][p][0], c["epipoles"][p][1], c["scale_factors"], c["level_sigma2"], a["ori"])
                ms.append(m); nms.append(nm)
            out[name + "/match"], out[name + "/nm"] = np.stack(ms), np.array(nms, np.int32)
        elif kind == "init":
            c = a["c"]
            nm, m, prev = ref.search_for_initialization(c["desc1"], c["kps1"], c["desc2"], c["kps2"], c["grid2"], c["prev_matched"], a["window"], c["nnratio"], 1)
            out[name + "/nm"], out[name + "/match"], out[name + "/prev"] = np.int32(nm), m, prev
        elif kind == "proj_localmap":
            c = a["c"]
            nm, asg = ref.sbp_local_map(c["desc_t"], c["kps_t"], c["grid"], c["queries"], c["level"], c["view_cos"], c["desc_q"], c["blocked0"], S.scale_factors(), a["th"], 0.8)
            out[name + "/nm"], out[name + "/assign"] = np.int32(nm), asg
        elif kind == "proj_f2f":
            f = a["c"]
            q = f["queries"]
            nm, asg = ref.sbp_frame_frame(f["desc_t"], f["kps_t"], f["grid"], q, (q["min_level"] + 1).astype(np.int32), f["desc_q"], f["blocked0"], S.scale_factors(), a["th"], 0.8, 1)
            out[name + "/nm"], out[name + "/assign"] = np.int32(nm), asg
        elif kind == "proj_reloc":
            f = a["f"]
            nm, asg = ref.sbp_frame_keyframe(f["desc_t"], f["kps_t"], f["grid"], a["q"], a["level"], f["desc_q"], a["blocked0"], S.scale_factors(), a["th"], a["acc"], 0.9, 1)
            out[name + "/nm"], out[name + "/assign"] = np.int32(nm), asg
        elif kind == "proj_sim3":
            f = a["f"]
            nm, asg = ref.sbp_keyframe_sim3(f["desc_t"], f["kps_t"], f["grid"], a["q"], a["level"], f["desc_q"], a["blocked0"], S.scale_factors(), int(a["th"]), 0.85)
            out[name + "/nm"], out[name + "/assign"] = np.int32(nm), asg
        print(name, kind, "ok")
    path = os.path.join(HERE, "ref_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
