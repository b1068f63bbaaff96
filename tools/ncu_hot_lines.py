#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares of one kernel from an .ncu-rep (--set full --import-source on).

ncu's SASS page carries per-instruction counters but no line numbers; nvdisasm -g on the same cubin (extracted from the
in-tree .so with cuobjdump) carries the line table.  Joined on the instruction offset inside the kernel.
usage: ncu_hot_lines.py REP KERNEL_REGEX MANGLED_SUBSTR [top_n]"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def line_table(mangled_substr):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "basevarc_b200", "libbvc_gpu.so")], cwd=d, capture_output=True)
    table = {}
    for f in os.listdir(d):
        if not f.endswith(".cubin"):
            continue
        txt = subprocess.run(["nvdisasm", "-g", os.path.join(d, f)], capture_output=True, text=True).stdout
        on, cur = False, None
        for ln in txt.splitlines():
            if ln.startswith(".text."):
                on = mangled_substr in ln
                continue
            if not on:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", ln)
            if m:
                table[int(m.group(1), 16)] = cur
        if table:
            break
    return table


def main():
    rep, kre, mangled = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    table = line_table(mangled)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    data = []
    for r in rows[hi + 1:]:
        if not r or r[0] in ("Kernel Name", "Address"):
            break                                             # first launch only
        data.append(r)
    ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    base = int(data[0][ia], 16)
    inst, samp = collections.Counter(), collections.Counter()
    for r in data:
        key = table.get(int(r[ia], 16) - base)
        inst[key] += int(r[ii]); samp[key] += int(r[isamp])
    ti, ts = sum(inst.values()), max(sum(samp.values()), 1)
    src = {}
    print("%s: %d SASS instructions, %.4g warp instructions executed, %d stall samples" % (kre, len(data), ti, ts))
    print("| file:line | %inst | %samples | source |\n|---|---|---|---|")
    for k, v in sorted(inst.items(), key=lambda kv: -kv[1])[:top]:
        text = ""
        if k:
            path = os.path.join(ROOT, "basevarc_b200", "csrc", k[0])
            if path not in src and os.path.exists(path):
                src[path] = open(path).read().splitlines()
            if path in src and k[1] <= len(src[path]):
                text = src[path][k[1] - 1].strip()[:110]
        print("| %s | %.1f | %.1f | `%s` |" % ("%s:%d" % k if k else "?", 100 * v / ti, 100 * samp[k] / ts, text))


if __name__ == "__main__":
    main()
