#!/bin/bash
# GPU job 19 (2 GPUs): lane-mode test; one server process driving both GPUs (BT2GPU_DEVICES=0,1); bench.py at N=2.
mkdir -p gpurun_out
O=gpurun_out/j19
timeout 200 python -m pytest tests/test_gpu_lane_mode.py -x -q > ${O}_pytest.log 2>&1; echo "pytest rc=$?" > ${O}_box.txt
timeout 300 python scripts/server_sweep.py --passes 2 --pass-timeout 60 --devices 0,1 --points 16:64:1:8 > ${O}_sweep_2dev.jsonl 2> ${O}_sweep_2dev.err; echo "sweep 2 devices rc=$?" >> ${O}_box.txt
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 > ${O}_bench_2gpu.json 2> ${O}_bench_2gpu.err; echo "bench N=2 rc=$?" >> ${O}_box.txt
cat ${O}_box.txt; tail -2 ${O}_pytest.log; cut -c1-330 ${O}_sweep_2dev.jsonl; nvidia-smi --query-gpu=index,memory.used --format=csv,noheader | head -3
python - <<'PY'
import json
try:
    b = json.loads(open("gpurun_out/j19_bench_2gpu.json").read().strip().splitlines()[-1])
    print("N=2 value %.0f e2e %s n_gpus %s" % (b["value"], b["e2e"]["value"], b["n_gpus"]))
except Exception as e:
    print("2-GPU line unreadable", e)
PY
tail -5 ${O}_bench_2gpu.err | cut -c1-300
