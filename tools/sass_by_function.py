#!/usr/bin/env python
"""SASS instruction count of one kernel by the source function the instructions were compiled from (nvdisasm -g line
info of a cubin extracted with `cuobjdump -xelf all lib.so`).  Code size matters here: the render loop is larger than
the instruction cache (stall_no_instruction in the ncu captures).
usage: python tools/sass_by_function.py cubin kernel-substring"""
import collections
import re
import subprocess
import sys


def main():
    cubin, key = sys.argv[1], sys.argv[2]
    txt = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout.splitlines()
    start = next(i for i, l in enumerate(txt) if l.startswith(".text.") and key in l)
    cnt = collections.Counter()
    cur, total = None, 0
    for line in txt[start + 1:]:
        if line.startswith(".text.") or line.startswith(".section"):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', line)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,6}\*/", line):
            total += 1
            cnt[cur] += 1
    print("total SASS instructions:", total)
    per_file = collections.Counter()
    for (f, l), c in cnt.items():
        per_file[f] += c
    print(per_file.most_common())
    for f in per_file:
        lines = sorted((l, c) for (ff, l), c in cnt.items() if ff == f)
        # bucket by 25 source lines
        b = collections.Counter()
        for l, c in lines:
            b[l // 20 * 20] += c
        print(f, " ".join("%d:%d" % (k, v) for k, v in sorted(b.items())))


if __name__ == "__main__":
    main()
