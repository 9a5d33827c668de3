#!/usr/bin/env python
"""SASS excerpt of the constant-bank sphere scan (scan_chunk_c) inside k_shade<16,false,false>: the window of
consecutive instructions with the most uniform-register FFMAs, i.e. one loop trip of the widest scan variant
(17 ray slots).  Evidence for profiles/r02_scan_shapes.md ("FFMA R, -R, UR, R").

  cuobjdump -sass raytracing-programs_b200/lib/librt_b200.so > /tmp/sass.txt
  python tools/sass_excerpt.py /tmp/sass.txt > profiles/r03_scan_chunk_c_sass.txt
"""
import re
import sys

txt = open(sys.argv[1]).read()
win = int(sys.argv[2]) if len(sys.argv) > 2 else 90
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = f.split("\n", 1)[0]
    if "k_shade" not in name or "Li16ELb0ELb0" not in name:
        continue
    ins = [l for l in f.splitlines() if re.search(r"/\*[0-9a-f]{4,6}\*/", l)]
    is_ur = [1 if re.search(r"FFMA[^;]*UR\d+", l) else 0 for l in ins]
    best, at, run = -1, 0, sum(is_ur[:win])
    for i in range(len(ins) - win):
        if run > best:
            best, at = run, i
        run += is_ur[i + win] - is_ur[i]
    seg = ins[at:at + win]
    ops = {}
    for l in seg:
        m = re.search(r"/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
        if m:
            ops[m.group(1).split(".")[0]] = ops.get(m.group(1).split(".")[0], 0) + 1
    print(f"# {name}")
    print(f"# function: {len(ins)} instructions, {sum(is_ur)} FFMA with a uniform-register (UR) operand, "
          f"{sum(1 for l in ins if re.search(r'FFMA', l))} FFMA in all, {sum(1 for l in ins if 'FMNMX3' in l)} FMNMX3")
    print(f"# window of {win} instructions at index {at}: {best} UR-operand FFMA; opcode histogram: "
          + ", ".join(f"{k} {v}" for k, v in sorted(ops.items(), key=lambda kv: -kv[1])))
    for l in seg:
        print(re.sub(r"\s+/\* 0x[0-9a-f]{16} \*/\s*$", "", l.rstrip()))
    break
