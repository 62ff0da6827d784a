#!/usr/bin/env python
"""Opcode histogram of the EXECUTED warp instructions of an ncu capture (--import-source on): share of each SASS opcode and its
average active lanes -- the evidence for the packed-FP32 claim (FFMA2 / FMUL2 / FADD2 share).
    python tools/sass_opcodes.py prof.ncu-rep [top]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
h = rows[1]
iE, iT, iS = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("Source")
ops = collections.defaultdict(lambda: [0, 0])
tot = 0
for r in rows[2:]:
    if len(r) <= iT:
        continue
    e, t, src = int(r[iE]), int(r[iT]), r[iS].strip()
    op = (src.split()[1] if src.startswith("@") else src.split()[0]).split(".")[0]
    ops[op][0] += e
    ops[op][1] += t
    tot += e
print("executed warp instructions: %d, static instructions: %d" % (tot, len(rows) - 2))
packed = sum(v[0] for k, v in ops.items() if k in ("FFMA2", "FMUL2", "FADD2"))
scalar = sum(v[0] for k, v in ops.items() if k in ("FFMA", "FMUL", "FADD"))
print("packed FP32 (FFMA2+FMUL2+FADD2): %.2f %% of executed instructions; scalar FFMA+FMUL+FADD: %.2f %%; of all FP32 roundings %.1f %% come from packed instructions"
      % (100.0 * packed / tot, 100.0 * scalar / tot, 100.0 * 2 * packed / (2 * packed + scalar)))
for k, v in sorted(ops.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-10s %6.2f %%   active lanes %.1f" % (k, 100.0 * v[0] / tot, v[1] / max(1, v[0])))
