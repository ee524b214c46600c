#!/bin/bash
# GPU job 10: longer passes (1 M reads) -- the steady state without the ramp and the tail of a 0.8 s pass.
mkdir -p gpurun_out
O=gpurun_out/j10
timeout 600 python scripts/server_sweep.py --stock --batch 1048576 --passes 2 --pass-timeout 90 \
  --points 16:128:1:8,16:192:1:8 > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" > ${O}_box.txt
cat ${O}_box.txt; python - <<'PY'
import json
for ln in open("gpurun_out/j10_sweep.jsonl"):
    d = json.loads(ln); c = d.pop("counters", None) or {}
    print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in d.items()})
    if c: print("   ", {k: round(v / 1e6, 1) for k, v in c.items() if k.startswith("us_")}, "calls", c.get("abi_batches"), {k: v for k, v in c.items() if k.startswith("idle")})
PY
grep pass ${O}_sweep.err
