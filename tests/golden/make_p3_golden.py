#!/usr/bin/env python
"""Generates tests/golden/p3_hashes.json: md5 + length of the P3 files the REFERENCE's own
ppm_WriteOutP3 (C/ppm.c:168-184, via oracle/_ref/libref.so: ref_write_p3) writes for the seeded
pixmaps of tests/rtlib.py:p3_pixmap.  Run in the build container (needs /root/reference built
into oracle/_ref by `make -C oracle`)."""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import rtlib  # noqa: E402
from test_p3 import CASES, _key  # noqa: E402

out = {}
for W, H, kind in CASES:
    txt = rtlib.ref_p3(rtlib.p3_pixmap(W, H, kind), W, H)
    out[_key(W, H, kind)] = {"md5": hashlib.md5(txt).hexdigest(), "bytes": len(txt)}
json.dump(out, open(os.path.join(HERE, "p3_hashes.json"), "w"), indent=1, sort_keys=True)
print("wrote", len(out), "entries")
