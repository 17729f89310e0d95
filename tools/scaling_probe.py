#!/usr/bin/env python
"""Predicts the multi-GPU kernel scaling on ONE GPU: renders every partition (tile_rows, r, N) of the bench
frame separately (L2 flushed before each) and prints max/mean kernel time over the parts for N = 1, 2, 4, 8.
usage: python tools/scaling_probe.py [--scene dragons] [--tile-rows 8]"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="dragons")
    ap.add_argument("--tile-rows", type=int, default=8)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--parts", default="1,2,4,8")
    ap.add_argument("--kernel", default="", help="pin RTC_KERNEL=mega|wave")
    ap.add_argument("--dealt", action="store_true", help="N scenes on this GPU share the frame tile by tile (rtc_scene_share_tiles) instead of row partitions")
    a = ap.parse_args()
    if a.kernel:
        os.environ["RTC_KERNEL"] = a.kernel
    import torch
    pkg = importlib.import_module("ray-tracer-challenge_b200")
    text, assets = scenes.shipped_scene(a.scene)
    info = pkg.parse_scene(text, assets, a.width, a.height)
    cam = info.camera
    frame = torch.empty((cam.vsize, cam.hsize, 3), dtype=torch.float64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    if a.dealt:
        blob, c = pkg.load_blob(text, assets, a.width, a.height)
        words = int(pkg.lib().rtc_tile_share_words(cam.hsize, cam.vsize))
        for n in [int(x) for x in a.parts.split(",")]:
            bufs = [torch.zeros(2 * words, dtype=torch.int32, device="cuda") for _ in range(n)]
            ptrs = [[bufs[r].data_ptr() + b * words * 4 for r in range(n)] for b in range(2)]
            worlds = [pkg.World(blob, 0) for _ in range(n)]
            for r, w in enumerate(worlds):
                w.share_tiles(r, n, words, ptrs)
            best = [1e9] * n
            for rnd in range(26):
                for r, w in enumerate(worlds):
                    if not a.no_flush:
                        flush.zero_()
                    torch.cuda.synchronize()
                    st = cam.render_device_frame(w, frame.data_ptr(), 5, None)
                    if rnd >= 17:  # (past the library's measurement of its kernel candidates: 3 + 2 x 6 frames)
                        best[r] = min(best[r], st.kernel_ms)
            line = {"scene": a.scene, "n_parts": n, "mode": "dealt", "max_ms": round(max(best), 4),
                    "mean_ms": round(sum(best) / n, 4), "parts": [round(p, 3) for p in best]}
            if os.environ.get("RTC_LPT_ALWAYS") == "1":
                # every frame recorded its tile costs (warp clock ticks / 64 at 1.965 GHz): what the last frame's tiles cost
                torch.cuda.synchronize()
                n_tiles = ((cam.hsize + 7) // 8) * ((cam.vsize + 3) // 4)
                both = bufs[0].view(2, words)[:, :n_tiles].to(torch.float64) * (64.0 / 1965.0)  # us of warp time
                c = both[0] if float(both[0].sum()) > 0 else both[1]
                top = torch.sort(c, descending=True).values
                line["tile_us"] = {"sum_ms": round(float(c.sum()) / 1e3, 2), "max": round(float(top[0]), 1),
                                   "top10": [round(float(x), 1) for x in top[:10]], "over_100us": int((c > 100).sum()),
                                   "over_50us": int((c > 50).sum()), "mean": round(float(c.mean()), 2),
                                   "note": "a heavy tile in the mixed kernel is worked on by four warps: its cost is their sum"}
            if os.environ.get("RTC_DUMP_COSTS"):  # the tile costs of the last recorded frame (RTC_LPT_ALWAYS=1), for offline scheduling studies
                import numpy as np
                torch.cuda.synchronize()
                n_tiles = ((cam.hsize + 7) // 8) * ((cam.vsize + 3) // 4)
                both = bufs[0].view(2, words)[:, :n_tiles].cpu().numpy().astype(np.uint32)
                np.save(os.environ["RTC_DUMP_COSTS"], both[0] if both[0].sum() > 0 else both[1])
            print(json.dumps(line), flush=True)
            for w in worlds:
                w.close()
        return
    for n in [int(x) for x in a.parts.split(",")]:
        per = []
        for r in range(min(n, 8)):
            best = 1e9
            for _ in range(14):
                if not a.no_flush:
                    flush.zero_()
                torch.cuda.synchronize()
                st = cam.render_device_frame(info.world, frame.data_ptr(), 5, (a.tile_rows, r, n))
                best = min(best, st.kernel_ms)
            per.append(best)
        print(json.dumps({"scene": a.scene, "n_parts": n, "tile_rows": a.tile_rows, "max_ms": round(max(per), 4),
                          "mean_ms": round(sum(per) / len(per), 4), "speedup_vs_1": None, "parts": [round(p, 3) for p in per]}),
              flush=True)


if __name__ == "__main__":
    main()
