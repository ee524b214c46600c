#!/bin/bash
# GPU job 18: the lane-mode test, the default bench (the line the round is judged on), the C3- and C4-shaped modes.
mkdir -p gpurun_out
O=gpurun_out/j18
timeout 300 python -m pytest tests/test_gpu_lane_mode.py tests/test_gpu_fm.py -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" > ${O}_box.txt
timeout 420 python bench.py > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" >> ${O}_box.txt
timeout 240 python bench.py --mode e2e --genome-mbp 250 --nseq 10 --steps 2 --warmup 1 --pass-timeout 60 > ${O}_bench_e2e.json 2> ${O}_bench_e2e.err; echo "bench e2e-mode rc=$?" >> ${O}_box.txt
timeout 240 python bench.py --mode local --genome-mbp 250 --nseq 10 --batch 65536 --steps 2 --warmup 1 --pass-timeout 60 > ${O}_bench_local.json 2> ${O}_bench_local.err; echo "bench local rc=$?" >> ${O}_box.txt
cat ${O}_box.txt; tail -3 ${O}_pytest.log; python - <<'PY'
import json
for f in ("bench", "bench_e2e", "bench_local"):
    try:
        b = json.loads(open("gpurun_out/j18_%s.json" % f).read())
        print(f, "value %.0f e2e %s cpu %s parity %s err %s" % (b["value"], b["e2e"]["value"], b.get("cpu_baseline", {}).get("value"), b["config"]["sam_identical"]["checked_in_this_run"], b["config"]["e2e_path"].get("error")))
    except Exception as e:
        print(f, "line unreadable", e)
PY
tail -4 ${O}_bench_local.err | cut -c1-300
