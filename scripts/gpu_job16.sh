#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/j17
timeout 300 python scripts/server_sweep.py --passes 3 --pass-timeout 60 \
  --points 16:32:1:8,16:48:1:8,16:64:1:8,16:96:1:8,16:64:1:8:BT2GPU_POLL_US=10 \
  > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" > ${O}_box.txt
cat ${O}_box.txt; python - <<'PY'
import json
for ln in open("gpurun_out/j17_sweep.jsonl"):
    d = json.loads(ln); c = d.pop("counters", None) or {}
    print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in d.items()})
    if c: print("   ", {k: round(v / 1e6, 1) for k, v in c.items() if k.startswith("us_")}, "calls", c.get("abi_batches"))
PY
grep -i "failed\|error" ${O}_sweep.err | head -5
