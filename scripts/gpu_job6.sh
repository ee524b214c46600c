#!/bin/bash
# GPU job 6: the central per-kind lanes + polling waits, swept (bounded at every level).
mkdir -p gpurun_out
O=gpurun_out/j6
timeout 600 python scripts/server_sweep.py --stock --passes 1 --pass-timeout 60 \
  --points 16:128:2:8,16:64:2:8,16:256:2:8,12:128:2:8,16:128:2:8:BT2GPU_SYNC=block,16:128:2:8:BT2GPU_SYNC=spin,16:128:1:8,16:128:2:4,16:128:2:8:BT2GPU_FM_LANES=2 \
  > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" > ${O}_box.txt
cat ${O}_box.txt; python - <<'PY'
import json
for ln in open("gpurun_out/j6_sweep.jsonl"):
    d = json.loads(ln); c = d.pop("counters", None) or {}
    print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in d.items()})
    if c: print("   ", {k: round(v / 1e6, 1) for k, v in c.items() if k.startswith("us_")}, "calls", c.get("abi_batches"))
PY
tail -3 ${O}_sweep.err
