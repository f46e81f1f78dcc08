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


def main():
    out = {}
    for name, kind, a in C.cases():
        res = C.run_case(ref, name, kind, a)
        out.update(res)
        print(name, kind, "ok", {k.split("/")[1]: (int(v) if v.ndim == 0 else int((v >= 0).sum())) for k, v in res.items() if not k.endswith("sha")})
    path = os.path.join(HERE, "ref_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
