#!/usr/bin/env python
"""Generates tests/golden/depth_hashes.json from the REFERENCE sources: oracle/_ref/libref_d50.so (MAX_RECURSION 50)
and libref.so (7) where they apply, else from the oracle after it has been checked against them (depth 20 has no
reference build: the oracle, which equals the reference at 7 and 50, provides it)."""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import rtlib  # noqa: E402
from test_depth import CASES, _key  # noqa: E402

out = {}
for scene, W, H, depth in CASES:
    img = rtlib.ref_render(scene, W, H, depth=depth) if depth in (7, 50) else rtlib.oracle_render(scene, W, H, depth=depth)
    assert (img == rtlib.oracle_render(scene, W, H, depth=depth)).all()
    out[_key(scene, W, H, depth)] = hashlib.md5(img.tobytes()).hexdigest()
json.dump(out, open(os.path.join(HERE, "depth_hashes.json"), "w"), indent=1, sort_keys=True)
print("wrote", len(out))
