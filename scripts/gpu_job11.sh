#!/bin/bash
# GPU job 11: longer passes for the batched server (ramp and tail of a 0.85 s pass weigh 20-30 %), and 16 clients x 16384 pairs for both servers.
mkdir -p gpurun_out
O=gpurun_out/j11
timeout 300 python scripts/server_sweep.py --batch 1048576 --passes 2 --pass-timeout 60 --points 16:128:1:8 > ${O}_sweep_1m.jsonl 2> ${O}_sweep_1m.err; echo "sweep1m rc=$?" > ${O}_box.txt
timeout 300 python scripts/server_sweep.py --stock --batch 524288 --passes 2 --pass-timeout 45 --points 16:128:1:16 > ${O}_sweep_512k.jsonl 2> ${O}_sweep_512k.err; echo "sweep512k rc=$?" >> ${O}_box.txt
cat ${O}_box.txt; python - <<'PY'
import json
for f in ("gpurun_out/j11_sweep_1m.jsonl", "gpurun_out/j11_sweep_512k.jsonl"):
  for ln in open(f):
    d = json.loads(ln); c = d.pop("counters", None) or {}
    print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in d.items()})
    if c: print("   ", {k: round(v / 1e6, 1) for k, v in c.items() if k.startswith("us_")}, "calls", c.get("abi_batches"), {k: v for k, v in c.items() if k.startswith("idle")})
PY
grep -h "pass\|failed" ${O}_sweep_1m.err ${O}_sweep_512k.err | cut -c1-200
