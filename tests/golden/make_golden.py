#!/usr/bin/env python
"""Regenerates tests/golden/*.npy from the reference tree (run in the build container only).
peano_keys_4x4x4.npy = the literal result_keys[] of libgadget/tests/test_peano.cpp:102."""
import re, os, sys
import numpy as np
REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
src = open(os.path.join(REF, "libgadget/tests/test_peano.cpp")).read()
m = re.search(r"result_keys\[\]\s*=\s*\{([^}]*)\}", src)
keys = np.array([int(x) for x in m.group(1).split(",")], dtype=np.uint64)
assert len(keys) == 64
np.save(os.path.join(HERE, "peano_keys_4x4x4.npy"), keys)
print("wrote", len(keys), "golden keys")
