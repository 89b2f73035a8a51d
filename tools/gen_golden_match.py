#!/usr/bin/env python
"""TEST INFRASTRUCTURE — writes tests/golden/match_*.npz: matcher outputs of the REFERENCE's own sources (oracle/_ref/libmatch_ref.so =
match/robust.cc, projection.cc, area.cc, fuse.cc, bow_tree.cc, data/common.cc, data/landmark.cc, camera/perspective.cc compiled
unmodified by oracle/Makefile) on the case generators of tests/test_match_gpu.py, after asserting that the port gives the same.
The fixtures hold outputs only; tests/test_match_golden_gpu.py rebuilds the inputs from the same seeded generators.
Run where /root/reference exists: python tools/gen_golden_match.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import match_cases  # noqa: E402
from oracle import pymatchref, pyoracle  # noqa: E402

pyoracle.build()
port, ref = pyoracle.port(), pymatchref.match_ref()
assert ref is not None, "oracle/_ref/libmatch_ref.so missing: run `make -f oracle/Makefile ref` where /root/reference exists"
out_dir = os.path.join(ROOT, "tests", "golden")
for name, run in match_cases.CASES.items():
    got_ref = run(ref, port)
    got_port = run(port, port)
    for k in got_ref:
        assert np.array_equal(np.asarray(got_ref[k]), np.asarray(got_port[k])), (name, k)
    np.savez_compressed(os.path.join(out_dir, f"match_{name}.npz"), **{k: np.asarray(v) for k, v in got_ref.items()})
    print(name, {k: (np.asarray(v).shape, int((np.asarray(v) >= 0).sum()) if np.asarray(v).dtype.kind == "i" else "") for k, v in got_ref.items()})
