#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/j20
timeout 200 python bench.py --mode local --genome-mbp 250 --nseq 10 --batch 65536 --steps 2 --warmup 1 --pass-timeout 60 > ${O}_bench_local.json 2> ${O}_bench_local.err; echo "bench local rc=$?" > ${O}_box.txt
timeout 300 python scripts/server_sweep.py --passes 3 --pass-timeout 60 \
  --points 16:64:1:8,16:64:1:8:BT2GPU_BATCH_MIN=32:BT2GPU_BATCH_WAIT_US=50,16:64:1:8:BT2GPU_BATCH_MIN=64:BT2GPU_BATCH_WAIT_US=100,16:96:1:8:BT2GPU_BATCH_MIN=64:BT2GPU_BATCH_WAIT_US=100 \
  > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" >> ${O}_box.txt
cat ${O}_box.txt; python - <<'PY'
import json
try:
    b = json.loads(open("gpurun_out/j20_bench_local.json").read())
    print("local: value %.0f e2e %s cpu %s parity %s err %s" % (b["value"], b["e2e"]["value"], b.get("cpu_baseline", {}).get("value"), b["config"]["sam_identical"]["checked_in_this_run"], b["config"]["e2e_path"].get("error")))
except Exception as e:
    print("local line unreadable", e)
for ln in open("gpurun_out/j20_sweep.jsonl"):
    d = json.loads(ln); c = d.pop("counters", None) or {}
    print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in d.items()})
    if c: print("   ", {k: round(v / 1e6, 1) for k, v in c.items() if k.startswith("us_")}, "calls", c.get("abi_batches"))
PY
