#!/usr/bin/env python
"""Joins the SASS page of an .ncu-rep with nvdisasm line info of the in-tree library and prints, per
CUDA source line (innermost inlined frame), the warp-stall samples and executed instructions.

usage: python tools/ncu_by_line.py gpurun_out/x.ncu-rep [--top 40] [--by func]
The library must be the build the profile was taken from (-lineinfo)."""
import argparse
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ray-tracer-challenge_b200", "librtc_b200.so")


BY = "inner"


def line_table(kernel_substr):
    """offset -> (file, line) for the first kernel whose mangled name contains kernel_substr."""
    tmp = tempfile.mkdtemp()
    # image by image: the merged image's name (all sources joined) is longer than a file name may be, and
    # `cuobjdump -xelf all` gives up on it before extracting the per-source ones
    listing = subprocess.run(["cuobjdump", "-lelf", LIB], capture_output=True, text=True).stdout
    for ln in listing.splitlines():
        name = ln.split(":", 1)[-1].strip()
        if name.endswith(".cubin") and len(name) < 200:
            subprocess.run(["cuobjdump", "-xelf", name, LIB], cwd=tmp, check=False, capture_output=True)
    table, lines_src = {}, {}
    for f in sorted(os.listdir(tmp)):
        if not f.endswith(".cubin"):
            continue
        out = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        cur, infunc, chain, fresh = None, False, [], True
        for ln in out.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", ln)
            if m:
                infunc = kernel_substr in m.group(1)
                continue
            if not infunc:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                if fresh:  # first line-info comment after an instruction starts a new inline chain
                    chain, fresh = [], False
                chain.append((os.path.basename(m.group(1)), int(m.group(2))))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*)", ln)
            if m and chain:
                fresh = True
                cur = chain[0] if BY == "inner" else (chain[-1] if BY == "outer" else
                                                      (" < ".join("%s:%d" % c for c in chain), 0))
                table[int(m.group(1), 16)] = (cur, m.group(2).strip())
        if table:
            break
    return table


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--kernel", default="render_kernel")
    ap.add_argument("--sass", action="store_true", help="print the hottest SASS instructions too")
    ap.add_argument("--by", default="inner", choices=["inner", "outer", "chain"],
                    help="attribute an instruction to its innermost inlined line, the outermost (kernel) line, or the chain")
    a = ap.parse_args()
    global BY
    BY = a.by
    out = subprocess.run(["ncu", "-i", a.rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    col = {h: i for i, h in enumerate(hdr)}
    table = line_table(a.kernel)
    body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    base = int(body[0][col["Address"]], 16) if body[0][col["Address"]].startswith("0x") else int(body[0][col["Address"]])
    agg = collections.defaultdict(lambda: [0, 0, 0, 0])  # samples, inst, thread inst, no_inst samples
    hot = []
    tot_s = tot_i = 0
    for r in body:
        addr = r[col["Address"]]
        off = (int(addr, 16) if addr.startswith("0x") else int(addr)) - base
        s = int(r[col["# Samples"]] or 0)
        i = int(r[col["Instructions Executed"]] or 0)
        ti = int(r[col["Thread Instructions Executed"]] or 0)
        ni = int(r[col["stall_no_inst"]] or 0) if "stall_no_inst" in col else 0
        key, sass = table.get(off, (("?", 0), r[col["Source"]]))
        g = agg[key]
        g[0] += s
        g[1] += i
        g[2] += ti
        g[3] += ni
        tot_s += s
        tot_i += i
        hot.append((s, i, key, sass))
    print("total samples %d, warp instructions %d, SASS lines %d" % (tot_s, tot_i, len(body)))
    byfile = collections.defaultdict(lambda: [0, 0])
    for (f, l), g in agg.items():
        byfile[f][0] += g[0]
        byfile[f][1] += g[1]
    for f, g in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
        print("  %-20s samples %5.1f%%  inst %5.1f%%" % (f, 100.0 * g[0] / max(1, tot_s), 100.0 * g[1] / max(1, tot_i)))
    print("%-28s %8s %8s %8s %8s" % ("source line", "samp%", "inst%", "thr/inst", "noinst%"))
    for (f, l), g in sorted(agg.items(), key=lambda kv: -kv[1][0])[: a.top]:
        print("%-28s %8.2f %8.2f %8.1f %8.1f" % (("%s:%d" % (f, l)) if l else f, 100.0 * g[0] / max(1, tot_s), 100.0 * g[1] / max(1, tot_i),
                                                  g[2] / max(1, g[1]), 100.0 * g[3] / max(1, g[0])))
    if a.sass:
        for s, i, key, sass in sorted(hot, key=lambda t: -t[0])[: a.top]:
            print("%6d %10d  %-24s %s" % (s, i, "%s:%d" % key, sass))


if __name__ == "__main__":
    main()
