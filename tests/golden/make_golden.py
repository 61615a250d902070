#!/usr/bin/env python
"""Generates tests/golden/golden_v1.npz from the REFERENCE itself
(oracle/_ref/libmgardx_ref.so = /root/reference's SERIAL MGARD-X path compiled
by oracle/Makefile).  Run in the build container (needs /root/reference):

    make -C oracle ref && python tests/golden/make_golden.py

Every case runs in a fresh process (see oracle_lib.fresh for why) and is only
kept if the reference's output for it is fully defined, i.e. the out-of-bounds
frequency read of GenerateCL is either not hit or reproduces the restatement's
convention -- the script asserts it.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib  # noqa: E402
from oracle_lib import smooth_field  # noqa: E402

INF = float("inf")
# (shape, dtype, s, rel, tol)
CASES = [
    ((17, 17, 17), "float32", INF, 1, 1e-3),
    ((33, 33, 33), "float32", INF, 1, 1e-4),
    ((20, 11, 6), "float32", INF, 0, 1e-3),
    ((20, 11, 6), "float64", 0.0, 1, 1e-4),
    ((65, 33), "float64", 0.0, 0, 1e-5),
    ((129,), "float32", INF, 1, 1e-3),
    ((5, 6, 9, 9, 17), "float64", 1.0, 1, 1e-3),
    ((30, 41), "float32", INF, 1, 1e-2),
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    assert oracle_lib.have_ref(), "build oracle/_ref first (make -C oracle ref)"
    out = {"ncases": np.int64(len(CASES))}
    for i, (shape, dt, s, rel, tol) in enumerate(CASES):
        u = smooth_field(shape, np.dtype(dt), seed=100 + i)
        orth = not (np.isinf(s) and len(shape) <= 3)
        coeff = oracle_lib.fresh("ref", "decompose", u, orth)
        blob, norm = oracle_lib.fresh("ref", "compress", u, rel, tol, s)
        codes = oracle_lib.fresh("ref", "quantize", coeff, rel, tol, s, norm)
        dec = oracle_lib.fresh("ref", "decompress", blob, shape, np.dtype(dt), rel,
                               tol, s, norm)
        allowed = oracle_lib.ref_allowed_outputs("compress", u, rel, tol, s)
        assert oracle_lib.same_container(blob, allowed[0][0]), \
            "case %d: reference output is not the defined one -- pick another" % i
        p = "c%d_" % i
        out[p + "shape"] = np.array(shape, np.int64)
        out[p + "dtype"] = np.array(dt)
        out[p + "params"] = np.array([s, rel, tol, norm], np.float64)
        out[p + "seed"] = np.int64(100 + i)
        out[p + "coeff_sha"] = np.array(sha(coeff))
        out[p + "codes_sha"] = np.array(sha(codes))
        out[p + "dec_sha"] = np.array(sha(dec))
        out[p + "blob"] = np.delete(blob, 7)  # byte 7 is uninitialised padding
        if u.size <= 8192:
            out[p + "coeff"] = coeff
            out[p + "codes"] = codes.astype(np.int32)
        print("case", i, shape, dt, "CR %.2f" % (u.nbytes / blob.size))
    np.savez_compressed(os.path.join(HERE, "golden_v1.npz"), **out)


if __name__ == "__main__":
    main()
