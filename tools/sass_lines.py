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


FUNC_RE = re.compile(r"^(?:SD|static __device__ __noinline__|__global__|DM_FN|static __device__)\b[^(;]*?(\w+)\s*\(")


def function_index(src_dir):
    idx = {}
    for f in os.listdir(src_dir):
        if not f.endswith((".cuh", ".h", ".cu")):
            continue
        starts = []
        for i, l in enumerate(open(os.path.join(src_dir, f), errors="ignore"), 1):
            m = FUNC_RE.match(l)
            if m:
                starts.append((i, m.group(1)))
        idx[f] = starts
    return idx


def function_of(idx, fl):
    name = "?"
    for start, n in idx.get(fl[0], []):
        if start <= fl[1]:
            name = n
        else:
            break
    return name


def sass_with_lines(so, kernel):
    tmp = tempfile.mkdtemp()
    if so.endswith(".cubin"):  # a run-time compiled kernel dumped with SAIL_B200_JIT_DUMP
        subprocess.check_call(["cp", os.path.abspath(so), tmp])
    else:
        subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
    out = []
    for f in sorted(os.listdir(tmp)):
        if not f.endswith(".cubin"):
            continue
        txt = subprocess.run(["nvdisasm", "--print-line-info-inline", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        cur, line, take, chain, fresh = None, ("?", 0), False, [], True
        for l in txt.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", l)
            if m:
                take = kernel in m.group(1)
                continue
            if not take:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', l)
            if m:
                if fresh:
                    chain, fresh = [], False
                chain.append((os.path.basename(m.group(1)), int(m.group(2))))
                line = chain[0]
                continue
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
            if m:
                fresh = True
                out.append((int(m.group(1), 16), m.group(2).strip(), line, tuple(chain)))
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
    # ---- by function bucket: the outermost inlined frame that is not the kernel / bounce loop itself ----
    idx = function_index(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "sail_b200", "csrc"))
    skip = {"sail_trace_megakernel", "sailMegaBody", "sailWaveBody", "__launch_bounds__", "tracePath", "intersectObjects", "?"}
    fagg = collections.defaultdict(lambda: [0, 0, 0])
    for i in range(n):
        chain = sass[i][3] or (sass[i][2],)
        # nvdisasm names the kernel's file for the outermost frame of an out-of-line device function (sail_sin ...),
        # which live in sail_shapes.cuh above the kernel's first line: look those up there.
        if chain[-1][0] == "sail_mega.cuh" and chain[-1][1] < 90:
            chain = tuple(chain[:-1]) + (("sail_shapes.cuh", chain[-1][1]),)
        names = [function_of(idx, fl) for fl in chain]
        bucket = next((nm for nm in reversed(names) if nm not in skip), names[-1])
        a = fagg[bucket]
        a[0] += prof[i][0]; a[1] += prof[i][1]; a[2] += prof[i][2]
        if os.environ.get("SASS_BUCKET") == bucket:
            print("  %6.3f%% %-40s %s" % (100.0 * prof[i][0] / tot, sass[i][1][:40], " < ".join("%s:%d" % fl for fl in chain)))
    print("%-28s %8s %8s %8s" % ("function", "inst%", "stall%", "thr/inst"))
    for k, a in sorted(fagg.items(), key=lambda kv: -kv[1][0])[:25]:
        print("%-28s %7.2f%% %7.2f%% %8.1f" % (k, 100.0 * a[0] / tot, 100.0 * a[1] / tots, a[2] / max(a[0], 1)))
    print()
    print("%-28s %8s %8s %8s" % ("file:line", "inst%", "stall%", "thr/inst"))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%-28s %7.2f%% %7.2f%% %8.1f" % ("%s:%d" % k, 100.0 * a[0] / tot, 100.0 * a[1] / tots, a[2] / max(a[0], 1)))


if __name__ == "__main__":
    main()
