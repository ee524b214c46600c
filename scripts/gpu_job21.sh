#!/bin/bash
# GPU job 21: the three bench modes with the round's final defaults.
mkdir -p gpurun_out
O=gpurun_out/j21
timeout 420 python bench.py > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" > ${O}_box.txt
timeout 200 python bench.py --mode e2e --genome-mbp 250 --nseq 10 --steps 2 --warmup 1 --pass-timeout 60 > ${O}_bench_e2e.json 2> ${O}_bench_e2e.err; echo "bench e2e-mode rc=$?" >> ${O}_box.txt
timeout 200 python bench.py --mode local --genome-mbp 250 --nseq 10 --batch 65536 --steps 2 --warmup 1 --pass-timeout 60 > ${O}_bench_local.json 2> ${O}_bench_local.err; echo "bench local rc=$?" >> ${O}_box.txt
cat ${O}_box.txt; python - <<'PY'
import json
for f in ("bench", "bench_e2e", "bench_local"):
    try:
        b = json.loads(open("gpurun_out/j21_%s.json" % f).read())
        print(f, "value %.0f e2e %s cpu %s parity %s err %s" % (b["value"], b["e2e"]["value"], b.get("cpu_baseline", {}).get("value"), b["config"]["sam_identical"]["checked_in_this_run"], b["config"]["e2e_path"].get("error")))
    except Exception as e:
        print(f, "line unreadable", e)
PY
grep "bench \[" ${O}_bench.err | tail -14
