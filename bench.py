#!/usr/bin/env python
"""bench.py — Mrays/s and frame time of the hot path on BASELINE.json's headline config.

Workload (config.workload): scenes/dragons.json rendered at 1920x1080, recursion depth 5, f64 —
the configuration BASELINE.json quotes its target on.  One "step" = one frame.

  value   Mrays/s (primary + shadow + secondary rays, counted on the device) with the scene resident
          in HBM and the frame left in HBM (rtc_render_device), CUDA-event timed, L2 flushed between
          steps (the scene, ~30 MB, would otherwise sit in the 126 MB L2).
  e2e     the same metric through the reference-facing call Camera.render(world) -> host Canvas
          (rtc_render): camera in, kernel, device->host copy of the f64 frame, all inside the timed
          region.  The world stays uploaded between frames, the reference's own "preheated"
          convention (README.md:122-125, lib.zig:166-190).
  N > 1   one process per GPU (torchrun); rank r renders the interleaved 8-row tiles r, r+N, ... and
          the packed rows are gathered to rank 0 over NCCL.  Strong scaling (same frame).
  --impl reference   the CPU restatement of the reference algorithm (oracle/, all host threads) on a
          bounded sample of the same frame (the reference itself is Zig: unavailable on this box).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
SCENE, WIDTH, HEIGHT, DEPTH, TILE_ROWS = "dragons", 1920, 1080, 5, 8
# algorithmic bytes per ray for dragons.json = 13 two-child f32 nodes x 64 B +
# 4 triangles x 72 B + 6 instance root tests x 64 B
BYTES_PER_RAY = 1504


def scene_source():
    from tests import scenes
    s = scenes.shipped_scene(SCENE)
    if s is None:
        raise SystemExit("bench: assets/_ref/scenes/%s.json is not staged; run __graft_entry__.build() where "
                         "/root/reference is mounted" % SCENE)
    return s


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one render_kernel launch of THIS workload, from the
    committed `ncu --set full` capture (profiles/traffic.json, written by profiles/summarize_ncu.py)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        return float(t["dram_bytes_per_launch"]), t.get("source", "")
    except Exception:
        return None, ""


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def l2_fraction(achieved_gbs):
    """The same figure against the L2-resident copy bandwidth measured on a B200 by tools/peaks (committed)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_alu_l2_peaks.json")) as f:
            l2 = float(json.load(f)["l2_copy_gbs"])
        return {"l2_copy_peak": l2, "frac_of_l2_copy": achieved_gbs / l2}
    except Exception:
        return {}


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons sampled WHILE the timed region runs (NVML, ~1 kHz; falls back to
    nvidia-smi, ~20 Hz, when pynvml is missing)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self.stop_flag = index, [], set(), None, False
        self.h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None

    def sample_nvml(self):
        nv = self.nv
        self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
            else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20),
                          ("hw_thermal_slowdown", 0x40)):
            if r & bit:
                self.reasons.add(name)

    def sample_smi(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
        f = [x.strip() for x in out.strip().split(",")]
        self.samples.append(float(f[0]))
        self.max_mhz = float(f[1])
        for n, v in zip(names, f[2:6]):
            if v.lower().startswith("active"):
                self.reasons.add(n)

    def run(self):
        while not self.stop_flag:
            try:
                if self.h is not None:
                    self.sample_nvml()
                else:
                    self.sample_smi()
            except Exception:
                pass
            time.sleep(0.001 if self.h is not None else 0.05)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "source": "nvml" if self.h is not None else "nvidia-smi"}


def cpu_sample(threads, bands=8, frac=0.125):
    """Oracle on `bands` evenly spaced row bands covering `frac` of the frame -> (rays, seconds, desc)."""
    from oracle import oracle_py
    text, assets = scene_source()
    sc = oracle_py.OracleScene(json_text=text, asset_dir=assets, width=WIDTH, height=HEIGHT)
    rows = max(1, int(HEIGHT * frac / bands))
    rays, secs = 0, 0.0
    import numpy as np
    out = np.zeros((HEIGHT, WIDTH, 3))
    for b in range(bands):
        y0 = int((b + 0.5) * HEIGHT / bands) - rows // 2
        t = time.perf_counter()
        _, c = sc.render(threads=threads, y0=y0, y1=y0 + rows)
        secs += time.perf_counter() - t
        rays += c["primary"] + c["shadow"] + c["secondary"]
    return rays, secs, "%d bands x %d rows of the %dx%d frame (%.1f%% of the pixels)" % (
        bands, rows, WIDTH, HEIGHT, 100.0 * bands * rows / HEIGHT)


def base_line(args, n):
    return {"metric": "Mrays/s", "unit": "Mrays/s", "n_gpus": n, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "scenes/dragons.json 1920x1080 depth 5 (6 x dragon.obj, 140,940 smooth triangles, "
                                   "4 lights); the scene file IS the input, there is no random data",
                       "l2": "flushed between steps (256 MiB memset)", "partition": "interleaved %d-row tiles" % TILE_ROWS,
                       "tile_order": "device-resident frames: tiles handed out in the order of their cost in the previous frame "
                                     "(scheduling only: every ray of every pixel is traced in every step, frames bit-identical; "
                                     "RTC_LPT=0 for the plain order); host canvases: strided row order"}}


def run_reference(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    for _ in range(args.warmup):
        cpu_sample(threads, bands=2, frac=0.004)
    rays = secs = 0
    desc = ""
    for _ in range(args.steps):
        r, s, desc = cpu_sample(threads)
        rays, secs = rays + r, secs + s
    v = rays / secs / 1e6
    line = base_line(args, args.gpus)
    line.update({"impl": "reference", "value": v, "ms_per_step": 1e3 * secs / args.steps,
                 "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": desc},
                 "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "note": "C++ restatement of the reference CPU algorithm (oracle/); the Zig reference cannot be built "
                         "on this box; ms_per_step is the time of the SAMPLE, not of a full frame"})
    emit(line)


def emit(line):
    """The ONE JSON line of the contract goes to the process's real stdout; everything else that libraries write
    to fd 1 (NCCL prints its version banner there) has been sent to stderr by quiet_stdout()."""
    os.write(REAL_STDOUT, (json.dumps(line) + "\n").encode())


REAL_STDOUT = 1


def quiet_stdout():
    global REAL_STDOUT
    sys.stdout.flush()
    REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exchange", default="auto", choices=["auto", "peer", "gather"],
                    help="N > 1: how the framebuffer reaches rank 0")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank)

    import numpy as np
    import torch
    import torch.distributed as dist
    pkg = importlib.import_module("ray-tracer-challenge_b200")
    if not torch.cuda.is_available() or pkg.device_count() == 0:
        raise SystemExit("bench: no CUDA device; this renderer has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    text, assets = scene_source()
    t0 = time.perf_counter()
    blob, cam = pkg.load_blob(text, assets, WIDTH, HEIGHT)
    t1 = time.perf_counter()
    wld = pkg.World(blob, local_rank)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    camera = pkg.Camera(cam)
    part = (TILE_ROWS, rank, world)
    rows = camera.partition_rows(part)
    mg = importlib.import_module("ray-tracer-challenge_b200.multi_gpu")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    # N > 1: every rank's kernel stores its tiles straight into rank 0's canvas over NVLink (PeerFrame);
    # if symmetric memory cannot be set up on this box, packed rows + one NCCL gather (GatherFrame).
    peer = gather = None
    exchange = "none (1 GPU)"
    if world > 1 and args.exchange != "gather":
        try:
            peer = mg.PeerFrame(WIDTH, HEIGHT, rank, world, dev)
            exchange = "peer stores into rank 0's frame (symmetric memory over NVLink) + device barrier"
        except Exception as e:  # noqa: BLE001
            if args.exchange == "peer":
                raise
            if rank == 0:
                print("bench: symmetric memory unavailable (%s); falling back to NCCL gather" % e, file=sys.stderr)
    if peer is None:
        gather = mg.GatherFrame(WIDTH, HEIGHT, TILE_ROWS, rank, world, dev)
        if world > 1:
            exchange = "packed rows + NCCL gather to rank 0 + de-interleave"

    def step_device():
        if peer is not None:
            st = camera.render_device_frame(wld, peer.root_ptr, DEPTH, part, stream)
            peer.finish()
            return st
        st = camera.render_device(wld, gather.local_ptr(), DEPTH, part, stream)
        gather.assemble()
        return st

    def frame_on_rank0():
        return peer.frame if peer is not None else (gather.frame if world > 1 else gather.local)

    # the caller-owned canvas of the e2e leg: page-locked (rtc_host_alloc, what the Zig shim uses), so the kernel
    # stores pixels straight into host memory; a pageable canvas (staging frame + cudaMemcpy) is timed beside it
    host = pkg.pinned_empty((HEIGHT, WIDTH, 3))
    host_pageable = np.zeros((HEIGHT, WIDTH, 3), dtype=np.float64)

    # N > 1: the host canvas is ONE shared-memory segment that every rank maps and page-locks; each rank's public
    # Camera.render(world, partition) stores its rows into it over its own PCIe link (multi_gpu.SharedHostFrame).
    # Fallback (registration refused): assemble on rank 0's GPU as in step_device, then one copy to pinned memory.
    shared = pinned = None
    if world > 1:
        ok = torch.zeros(1, device=dev)
        try:
            shared = mg.SharedHostFrame(WIDTH, HEIGHT, rank, world)
        except Exception as e:  # noqa: BLE001
            ok += 1
            print("bench: shared host canvas unavailable on rank %d (%s)" % (rank, e), file=sys.stderr)
        dist.all_reduce(ok)
        if float(ok) > 0:
            if shared is not None:
                shared.close()
            shared = None
            pinned = torch.empty((HEIGHT, WIDTH, 3), dtype=torch.float64).pin_memory() if rank == 0 else None

    def step_e2e(canvas=None):
        if world == 1:  # the public call: Camera.render(world) -> host Canvas
            return camera.render(wld, DEPTH, None, out=host if canvas is None else canvas).stats
        if shared is not None:  # the same public call on every rank, rows of its partition, then a barrier
            st = camera.render(wld, DEPTH, part, out=shared.array).stats
            dist.barrier()
            return st
        st = step_device()  # render + exchange on the GPUs, then the assembled frame goes to the host on rank 0
        if rank == 0:
            pinned.copy_(frame_on_rank0(), non_blocking=False)
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 6)):  # >= 6: the library measures its kernel candidates on the first four to six frames
        step_device()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    rays = kernel_ms = culled = 0
    launches = 0
    wall0 = time.perf_counter()
    for i in range(args.steps):
        flush.zero_()
        barrier()
        ev[i][0].record()
        st = step_device()
        ev[i][1].record()
        rays += st.rays
        culled += st.shadow_rays_culled
        kernel_ms += st.kernel_ms
        launches += st.launches
    barrier()
    wall = time.perf_counter() - wall0
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    # e2e leg (host buffers, D2H inside the timed region), wall clock bracketed by synchronize
    for _ in range(2):
        step_e2e()
    barrier()
    e0 = time.perf_counter()
    e_rays = 0
    for i in range(args.steps):
        e_rays += step_e2e().rays
    barrier()
    e_secs = time.perf_counter() - e0
    p_secs = None
    if world == 1:
        step_e2e(host_pageable)
        barrier()
        p0 = time.perf_counter()
        for i in range(args.steps):
            step_e2e(host_pageable)
        barrier()
        p_secs = time.perf_counter() - p0
        assert np.array_equal(host, host_pageable), "pinned and pageable canvases differ"
    if shared is not None:
        step_device()
        barrier()
        if rank == 0:  # the host canvas written by N ranks over N links == the frame assembled on rank 0's GPU
            assert np.array_equal(shared.array, frame_on_rank0().cpu().numpy()), "shared host canvas differs from the device frame"
        barrier()
    sampler.stop_flag = True
    sampler.join()
    # what the e2e leg is bounded by: a plain DMA copy of one frame, device -> page-locked host memory
    d2h_ms = None
    if rank == 0:
        src = torch.empty((HEIGHT, WIDTH, 3), dtype=torch.float64, device=dev)
        dst = torch.empty((HEIGHT, WIDTH, 3), dtype=torch.float64).pin_memory()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e9
        for _ in range(6):
            c0.record()
            dst.copy_(src, non_blocking=True)
            c1.record()
            torch.cuda.synchronize()
            best = min(best, c0.elapsed_time(c1))
        d2h_ms = best

    t = torch.tensor([dev_ms, e_secs], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(rays), float(e_rays), float(launches)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    dev_ms, e_secs = float(t[0]), float(t[1])
    rays_all, e_rays_all, launches_all = float(tot[0]), float(tot[1]), int(tot[2])

    if shared is not None:
        shared.close()
    if rank == 0:
        peak, which = peaks()
        value = rays_all / (dev_ms * 1e-3) / 1e6
        # roofline of the render kernel on THIS rank: algorithmic bytes per launch / mean launch time
        k_ms = kernel_ms / args.steps
        achieved = (rays / args.steps) * BYTES_PER_RAY / (k_ms * 1e-3) / 1e9
        traffic, traffic_src = ncu_traffic()
        line = base_line(args, world)
        line.update({
            "value": value, "ms_per_step": dev_ms / args.steps,
            "e2e": {"value": e_rays_all / e_secs / 1e6, "unit": "Mrays/s", "ms_per_step": 1e3 * e_secs / args.steps,
                    "h2d_bytes_per_step": int(ctypes_sizeof_camera(pkg)), "d2h_bytes_per_step": int(HEIGHT * WIDTH * 24),
                    "canvas": "page-locked host memory (rtc_host_alloc): the kernel stores finished pixels into it over "
                              "PCIe while rendering" if world == 1 else (
                                  "one shared-memory host canvas, page-locked in every rank (cudaHostRegister): each rank's "
                                  "Camera.render(world, partition) stores its rows into it over its own PCIe link, then a barrier"
                                  if shared is not None else "rank 0 copies the assembled frame to pinned memory"),
                    "pageable_canvas_ms_per_step": (1e3 * p_secs / args.steps) if p_secs else None,
                    "d2h_copy_floor_ms": d2h_ms,
                    "d2h_copy_floor": "one cudaMemcpy of the f64 frame (49.8 MB) into pinned memory over ONE link of this box"
                                      + (": the e2e step cannot be shorter than max(kernel, this)" if world == 1 else
                                         "; with the shared host canvas each of the N ranks moves 1/N of it over its own link")},
            "gpu_launches": launches_all,
            "rays_per_frame": rays_all / args.steps,
            # shadow rays the reference casts (and `value` counts) but this renderer does not trace because the
            # light is behind the surface: the result is ambient either way (material.zig:55-63); rank 0's share
            "shadow_rays_culled_per_frame": culled / args.steps,
            "kernel_ms": k_ms,
            "wall_s_timed_region": wall,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": which + " (MEASURED_PEAKS.json hbm_gbs)",
                         "model": "%d algorithmic B/ray (SURVEY.md 8d) x rays per launch / CUDA-event launch time; the "
                                  "working set is L2-resident (traffic = measured DRAM bytes per launch: ~1 %% of the "
                                  "algorithmic bytes), so this is a node-fetch figure set against the HBM peak and may "
                                  "exceed it; frac_of_l2_copy sets it against the measured L2-resident copy bandwidth" % BYTES_PER_RAY,
                         **l2_fraction(achieved)},
            "clocks": sampler.summary(),
            "exchange": exchange,
            "scene_setup": {"parse_flatten_s": t1 - t0, "bvh_build_upload_s": t2 - t1, "blob_bytes": len(blob)},
        })
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            r, s, desc = cpu_sample(threads)
            line["cpu_baseline"] = {"value": r / s / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "port",
                                    "sample": desc}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def ctypes_sizeof_camera(pkg):
    import ctypes
    return ctypes.sizeof(pkg.rtc_camera)


if __name__ == "__main__":
    main()
