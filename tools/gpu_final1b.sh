#!/bin/bash
# final 1-GPU evidence run (end of round 2): tests, parity report, bench line, reference arm, ncu capture of the dragons kernel + launch list, kernel times of all scenes
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/parity_r2.jsonl
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/final_pytest.log
tail -4 gpurun_out/final_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/final_bench1.json 2> gpurun_out/final_bench1.err; echo "bench rc $?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err; echo "ref rc $?"
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 300 $NCU -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2_d_dragons python tools/ncu_one.py dragons 1920 1080 8 > gpurun_out/final_ncu_dragons.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/final_launches.log 2>&1
python tools/bench_all.py --all --reps 5 > gpurun_out/r2_bench_all.jsonl 2>gpurun_out/final_bench_all.err
python tools/scaling_probe.py --scene dragons --parts 2,4,8 --dealt > gpurun_out/r2_scaling_probe_dealt.jsonl 2>/dev/null
python - <<'PY'
import json
d=json.loads(open('gpurun_out/final_bench1.json').read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"],4), "kernel", round(d["kernel_ms"],4), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],4), "rgba8", round(d["e2e_rgba8"]["ms_per_step"],4), "first", d.get("first_frame_ms"), "cold", d.get("cold_first_frame_s"), "cpu", d["cpu_baseline"]["value"], "frac", d["roofline"]["frac"])
for k,v in d.get("configs",{}).items():
    print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ("kernel_ms","value","e2e_ms")})
r=json.loads(open('gpurun_out/final_ref.json').read().strip().splitlines()[-1]); print("ref", r["value"], r["ms_per_step"])
PY
ls -la gpurun_out/r2_d_dragons.ncu-rep
