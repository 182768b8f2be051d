#!/usr/bin/env python
"""Static view of a kernel's SASS: every loop (backward branch) with its instruction count, and the opcode histogram
of the big ones.  Usage: python tools/sass_loops.py <so-or-cubin> <mangled-name-substring> [min-size]"""
import collections
import re
import subprocess
import sys

so, key = sys.argv[1], sys.argv[2]
minsz = int(sys.argv[3]) if len(sys.argv) > 3 else 300
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
for f in funcs[1:]:
    name = f.split("\n", 1)[0].strip()
    if key not in name:
        continue
    ins = []
    for l in f.split("\n"):
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    print(f"== {name}: {len(ins)} SASS instructions")
    for a, t in ins:
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            tgt = int(m.group(1), 16)
            body = [x for x in ins if tgt <= x[0] <= a]
            print(f"loop 0x{tgt:x}..0x{a:x}: {len(body)} instructions")
            if len(body) >= minsz:
                c = collections.Counter()
                for _, x in body:
                    x = re.sub(r"^@!?U?P\d+\s+", "", x)
                    op = x.split()[0]
                    c[op if op.startswith(("MUFU", "IMAD")) else op.split(".")[0]] += 1
                print("   " + ", ".join(f"{v} {k}" for k, v in c.most_common(40)))
