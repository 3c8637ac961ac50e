"""Attribute executed warp-instructions of one kernel in an .ncu-rep to CUDA source lines.
ncu's CLI source page is SASS-only, so the line mapping comes from `nvdisasm -g` on the cubin of the
shipped .so (both list a kernel's instructions in address order).

usage: python profiles/sass_by_line.py REPORT.ncu-rep KERNEL_REGEX [launch-index] [top-N]
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.environ.get("PRT_LIB") or os.path.join(ROOT, "parallelraytracing_b200", "libprt_b200.so")   # PRT_LIB: the library the report was captured with


def line_table():
    """{mangled kernel name: [(file, line, inlined_chain) per instruction in address order]}"""
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", SO], cwd=tmp, capture_output=True)
    table = {}
    for f in os.listdir(tmp):
        if not f.endswith(".cubin") or f.count("-"):
            continue
        out = subprocess.run(["nvdisasm", "-g", "-c", f], cwd=tmp, capture_output=True, text=True).stdout
        cur, loc = None, ("?", 0)
        for ln in out.split("\n"):
            m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
            if m:
                cur = m.group(1); table[cur] = []; continue
            m = re.match(r'\s*//## File "(.*)", line (\d+)(.*)', ln)
            if m:
                loc = (os.path.basename(m.group(1)), int(m.group(2))); continue
            if cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
                table[cur].append(loc)
    return table


def main():
    rep, pat = sys.argv[1], sys.argv[2]
    which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat],
                         capture_output=True, text=True).stdout
    blocks = out.split('"Kernel Name",')[1:]
    blk = blocks[which]
    rows = list(csv.reader(io.StringIO('"Kernel Name",' + blk)))
    name = rows[0][1]
    H = rows[1]
    ie, si, te = H.index("Instructions Executed"), H.index("# Samples"), H.index("Thread Instructions Executed")
    insts = [r for r in rows[2:] if len(r) > ie]
    table = line_table()
    short = re.sub(r"\W", "", name.split("(")[0].split("::")[-1].split("<")[0])
    cands = [k for k in table if short in k and len(table[k]) == len(insts)]
    if not cands:
        print("no nvdisasm match for", name, len(insts), {k: len(v) for k, v in table.items() if short in k}); return
    locs = table[cands[0]]
    agg = collections.defaultdict(lambda: [0, 0, 0, 0])
    tot = [0, 0]
    for r, loc in zip(insts, locs):
        a = agg[loc]
        a[0] += int(r[ie]); a[1] += int(r[si]); a[2] += int(r[te]); a[3] += 1
        tot[0] += int(r[ie]); tot[1] += int(r[si])
    print("%s: %d SASS instructions, %d warp-insts executed, %d samples" % (name.split("(")[0], len(insts), tot[0], tot[1]))
    src_cache = {}
    for loc, a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        f = loc[0]
        if f not in src_cache:
            p = os.path.join(ROOT, "parallelraytracing_b200", "csrc", f)
            src_cache[f] = open(p).read().split("\n") if os.path.exists(p) else None
        text = src_cache[f][loc[1] - 1].strip()[:90] if src_cache[f] and loc[1] <= len(src_cache[f]) else ""
        print("%5.1f%% inst %5.1f%% samp  thr/inst %4.1f  n=%3d  %s:%d  %s" % (100.0 * a[0] / max(1, tot[0]), 100.0 * a[1] / max(1, tot[1]),
                                                                      a[2] / max(1, a[0]), a[3], f, loc[1], text))


if __name__ == "__main__":
    main()
