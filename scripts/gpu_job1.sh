#!/bin/bash
# first GPU job of round 2: tests, the 3.1 Gbp bench, a sweep of the batched server's knobs
mkdir -p gpurun_out
O=gpurun_out/j1
(nvidia-smi --query-gpu=name,memory.total --format=csv; nproc; free -g; df -h /tmp) > ${O}_box.txt 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > ${O}_smoke.log 2>&1; echo "smoke rc=$?" >> ${O}_box.txt
timeout 1800 python -m pytest tests -m gpu -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" >> ${O}_box.txt
timeout 1500 python bench.py --steps 3 --warmup 2 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" >> ${O}_box.txt
timeout 1200 python scripts/server_sweep.py --stock --points 8:128:2:8,16:128:2:8,16:256:2:8,16:128:3:8,16:64:2:8,12:128:2:4,16:128:1:8,16:128:2:12 > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" >> ${O}_box.txt
tail -5 ${O}_pytest.log; cat ${O}_box.txt; cat ${O}_sweep.jsonl | cut -c1-300
