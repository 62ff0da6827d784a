#!/usr/bin/env python
"""Attribute executed instructions / stall samples of an ncu capture to CUDA source lines.

    python tools/sass_lines.py <prof.ncu-rep> <libsail_b200.so> <kernel-name-substring> [top]

ncu's `--page source --csv` lists SASS only; nvdisasm --print-line-info on the same cubin gives the
file:line of every SASS instruction.  The two are joined by instruction order within the kernel."""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def sass_with_lines(so, kernel):
    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
    out = []
    for f in sorted(os.listdir(tmp)):
        if not f.endswith(".cubin"):
            continue
        txt = subprocess.run(["nvdisasm", "--print-line-info", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        cur, line, take = None, ("?", 0), False
        for l in txt.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", l)
            if m:
                take = kernel in m.group(1)
                continue
            if not take:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', l)
            if m:
                line = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
            if m:
                out.append((int(m.group(1), 16), m.group(2).strip(), line))
        if out:
            break
    return out


def main():
    rep, so, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    h = rows[1]
    iE, iS, iT = h.index("Instructions Executed"), h.index("# Samples"), h.index("Thread Instructions Executed")
    prof = [(int(r[iE]), int(r[iS]), int(r[iT]), r[h.index("Source")].strip()) for r in rows[2:] if len(r) > iT]
    sass = sass_with_lines(so, kernel)
    print("ncu instructions: %d, nvdisasm instructions: %d" % (len(prof), len(sass)))
    n = min(len(prof), len(sass))
    agg = collections.defaultdict(lambda: [0, 0, 0])
    for i in range(n):
        a = agg[sass[i][2]]
        a[0] += prof[i][0]; a[1] += prof[i][1]; a[2] += prof[i][2]
    tot = sum(a[0] for a in agg.values()) or 1
    tots = sum(a[1] for a in agg.values()) or 1
    print("%-28s %8s %8s %8s" % ("file:line", "inst%", "stall%", "thr/inst"))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%-28s %7.2f%% %7.2f%% %8.1f" % ("%s:%d" % k, 100.0 * a[0] / tot, 100.0 * a[1] / tots, a[2] / max(a[0], 1)))


if __name__ == "__main__":
    main()
