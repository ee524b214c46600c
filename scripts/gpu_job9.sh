#!/bin/bash
# GPU job 9: knob sweep after the allocator / read-buffer / scratch fixes.
mkdir -p gpurun_out
O=gpurun_out/j9
echo "thp: $(cat /sys/kernel/mm/transparent_hugepage/enabled) defrag: $(cat /sys/kernel/mm/transparent_hugepage/defrag)" > ${O}_box.txt
nproc >> ${O}_box.txt; lscpu | grep -E "Model name|L2|L3|Thread|Core|Socket" >> ${O}_box.txt; free -g | head -2 >> ${O}_box.txt
timeout 560 python scripts/server_sweep.py --passes 2 --pass-timeout 60 \
  --points 16:128:1:8:GLIBC_TUNABLES=glibc.malloc.hugetlb=1,16:192:1:8,16:256:1:8,15:128:1:8,16:128:1:8:BT2GPU_PUBLISH=8,16:128:1:8:BT2GPU_PUBLISH=256,16:128:2:8:BT2GPU_FM_LANES=2,16:128:1:16 \
  > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" >> ${O}_box.txt
cat ${O}_box.txt; python - <<'PY'
import json
for ln in open("gpurun_out/j9_sweep.jsonl"):
    d = json.loads(ln); c = d.pop("counters", None) or {}
    print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in d.items()})
    if c: print("   ", {k: round(v / 1e6, 1) for k, v in c.items() if k.startswith("us_")}, "calls", c.get("abi_batches"))
PY
