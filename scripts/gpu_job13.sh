#!/bin/bash
# GPU job 13: the whole GPU suite after the ABI changes (polling waits, no events, stage waits folded into the batch
# calls), then the batched server again and the default bench.
mkdir -p gpurun_out
O=gpurun_out/j13
timeout 900 python -m pytest tests -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" > ${O}_box.txt
timeout 200 python scripts/server_sweep.py --passes 2 --pass-timeout 60 --points 16:128:1:8,16:64:1:8 > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" >> ${O}_box.txt
timeout 400 python bench.py --steps 3 --warmup 3 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" >> ${O}_box.txt
cat ${O}_box.txt; tail -3 ${O}_pytest.log; python - <<'PY'
import json
for ln in open("gpurun_out/j13_sweep.jsonl"):
    d = json.loads(ln); c = d.pop("counters", None) or {}
    print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in d.items()})
    if c: print("   ", {k: round(v / 1e6, 1) for k, v in c.items() if k.startswith("us_")}, "calls", c.get("abi_batches"))
try:
    b = json.loads(open("gpurun_out/j13_bench.json").read())
    print("bench value %.0f e2e %s cpu %s parity %s" % (b["value"], b["e2e"], b.get("cpu_baseline", {}).get("value"), b["config"]["sam_identical"]["checked_in_this_run"]))
except Exception as e:
    print("bench line unreadable", e)
PY
grep "bench \[" ${O}_bench.err | tail -30
