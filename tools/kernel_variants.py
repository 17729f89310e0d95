#!/usr/bin/env python
"""Runs tools/bench_all.py once per kernel/library variant (each in its own process: the switches are read when the
library is loaded) and collects the JSON lines.  A variant = label, library path (or "" for the in-tree build), env.
usage: python tools/kernel_variants.py [--only dragons,teapot] [--out gpurun_out/variants.jsonl] label:lib:K=V,K=V ..."""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="dragons,teapot")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "variants.jsonl"))
    ap.add_argument("--reps", type=int, default=7)
    ap.add_argument("--tool", default="bench_all", choices=["bench_all", "scaling_probe"])
    ap.add_argument("variants", nargs="+")
    a = ap.parse_args()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "a") as out:
        for v in a.variants:
            label, lib, envs = (v.split(":") + ["", ""])[:3]
            env = dict(os.environ)
            if lib:
                env["RTC_B200_LIB"] = os.path.join(ROOT, lib)
            for kv in filter(None, envs.split(",")):
                k, val = kv.split("=")
                env[k] = val
            if a.tool == "bench_all":
                cmd = [sys.executable, os.path.join(ROOT, "tools", "bench_all.py"), "--only", a.only, "--reps", str(a.reps)]
            else:
                cmd = [sys.executable, os.path.join(ROOT, "tools", "scaling_probe.py"), "--scene", a.only.split(",")[0]]
            r = subprocess.run(cmd, env=env, capture_output=True, text=True)
            if r.returncode != 0:
                print(label, "FAILED", r.stderr[-600:], flush=True)
                continue
            for line in r.stdout.splitlines():
                try:
                    d = json.loads(line)
                except ValueError:
                    continue
                d["variant"] = label
                out.write(json.dumps(d) + "\n")
                out.flush()
                if a.tool == "bench_all":
                    print("%-18s %-10s %8.4f ms  %8.1f Mrays/s  checksum %.9g  rays %d" % (
                        label, d["scene"], d["kernel_ms"], d["Mrays_s"], d["checksum"], d["rays"]), flush=True)
                else:
                    print("%-18s n_parts %d  max %.4f  mean %.4f" % (label, d["n_parts"], d["max_ms"], d["mean_ms"]), flush=True)


if __name__ == "__main__":
    main()
