#!/bin/bash
cd "$(dirname "$0")/.."
RTC_COPY_CTAS=8 RTC_LPT=0 timeout 120 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -1 | sed 's/^/LPT=0 /' | tee gpurun_out/run24_e2e.log
RTC_COPY_CTAS=8 RTC_KERNEL=mega timeout 120 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -1 | sed 's/^/mega /' | tee -a gpurun_out/run24_e2e.log
RTC_COPY_CTAS=8 RTC_TRACE_FRAMES=1 timeout 120 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -12 | tee -a gpurun_out/run24_e2e.log
python - <<'PY' 2>&1 | tee -a gpurun_out/run24_e2e.log
import torch, time
dev = torch.empty((1080,1920,3), dtype=torch.float64, device="cuda")
host = torch.empty((1080,1920,3), dtype=torch.float64).pin_memory()
for _ in range(3): host.copy_(dev, non_blocking=True)
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(20): host.copy_(dev, non_blocking=True)
torch.cuda.synchronize(); print("DMA copy ms", (time.perf_counter()-t)/20*1e3)
PY
