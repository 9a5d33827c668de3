#!/usr/bin/env python
"""Generates tests/golden/bin_hashes.json: md5 + length of the files the REFERENCE's own
ppm_WriteOutP6 / ppm_WriteOutP7 (C/ppm.c:193-227, from oracle/_ref/libref.so) write for the seeded
RGBA pixmaps of tests/test_host.py.  Run in the build container after `make -C oracle`."""
import hashlib
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import rtlib  # noqa: E402
from test_host import BIN_CASES, _rgba, _write_with  # noqa: E402

out = {}
r = rtlib.ref()
with tempfile.TemporaryDirectory() as d:
    for W, H in BIN_CASES:
        rgba, _ = _rgba(W, H)
        for fn in ("ppm_WriteOutP6", "ppm_WriteOutP7"):
            data = _write_with(r, fn, W, H, 4, rgba, os.path.join(d, "g.ppm"))
            out[f"{fn}_{W}x{H}"] = {"md5": hashlib.md5(data).hexdigest(), "bytes": len(data)}
json.dump(out, open(os.path.join(HERE, "bin_hashes.json"), "w"), indent=1, sort_keys=True)
print("wrote", len(out), "entries")
