#!/usr/bin/env python
"""Extracts the metrics DESIGN.md / bench.py quote from an .ncu-rep (run where ncu is installed).
usage: python profiles/summarize_ncu.py gpurun_out/prof.ncu-rep > profiles/<name>.txt"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__thread_inst_executed_per_inst_executed.pct",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "lts__t_bytes.sum.per_second",
        "l1tex__t_bytes.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
        "dram__bytes_write.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum"]


def traffic_json(path, out_path, note):
    """profiles/traffic.json for bench.py's roofline.traffic: DRAM bytes of ONE launch of the profiled kernel."""
    import json
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    rec = dict(zip(hdr, rows[2]))
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot = 0.0
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        tot += float(rec[k].replace(",", "")) * scale[units[hdr.index(k)]]
    with open(out_path, "w") as f:
        json.dump({"dram_bytes_per_launch": tot, "kernel": rec.get("Kernel Name"),
                   "source": "ncu --set full, %s, %s" % (path.split("/")[-1], note)}, f, indent=1)
        f.write("\n")


def roofline_json(path, out_path, note):
    """profiles/r2_roofline_inputs.json for bench.py's roofline block: what ONE launch of the profiled kernel executed
    and moved (warp instructions, DRAM bytes) and how busy the pipes were."""
    import json
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    rec = dict(zip(hdr, rows[2]))
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

    def num(k):
        return float(rec[k].replace(",", ""))
    dram = sum(num(k) * scale[units[hdr.index(k)]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    with open(out_path, "w") as f:
        json.dump({"kernel": rec.get("Kernel Name"), "source": "ncu --set full, %s, %s" % (path.split("/")[-1], note),
                   "warp_instructions_per_launch": num("smsp__inst_executed.sum"),
                   "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                   "fp64_pipe_pct": num("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
                   "threads_per_inst": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
                   "l1_hit_pct": num("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": num("lts__t_sector_hit_rate.pct"),
                   "dram_bytes_per_launch": dram, "dram_pct": num("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
                   "registers_per_thread": num("launch__registers_per_thread"),
                   "warps_active_pct": num("sm__warps_active.avg.pct_of_peak_sustained_active"),
                   "duration_under_ncu_ms": num("gpu__time_duration.sum") * {"us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}.get(
                       units[hdr.index("gpu__time_duration.sum")], 1.0)}, f, indent=1)
        f.write("\n")


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        rec = dict(zip(hdr, r))
        print("kernel:", rec.get("Kernel Name"), " id", rec.get("ID"))
        for k in KEYS:
            if k in rec:
                print("  %-70s %-12s %s" % (k, units[hdr.index(k)], rec[k]))
        stalls = [(float(v), h) for h, v in rec.items() if h.startswith("smsp__average_warps_issue_stalled") and
                  h.endswith("_per_issue_active.ratio") and v not in ("", "n/a")]
        for v, h in sorted(stalls, reverse=True)[:8]:
            print("  stall %-64s %.3f" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2] == "--roofline":
        roofline_json(sys.argv[1], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
    elif len(sys.argv) > 2 and sys.argv[2] == "--traffic":
        traffic_json(sys.argv[1], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
    else:
        main(sys.argv[1])
