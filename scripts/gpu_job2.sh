#!/bin/bash
# GPU job 2: DP parity tests first (new H-only matrix stream), then everything else, bench and sweep
mkdir -p gpurun_out
O=gpurun_out/j2
(nvidia-smi --query-gpu=name,memory.total --format=csv; nproc; free -g | head -2) > ${O}_box.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_dp.py tests/test_gpu_properties.py tests/test_gpu_pipeline.py tests/test_gpu_pairs.py -x -q > ${O}_pytest_dp.log 2>&1; echo "pytest_dp rc=$?" >> ${O}_box.txt
timeout 2400 python -m pytest tests -m gpu -q --deselect tests/test_gpu_dp.py --deselect tests/test_gpu_properties.py --deselect tests/test_gpu_pipeline.py --deselect tests/test_gpu_pairs.py > ${O}_pytest_rest.log 2>&1; echo "pytest_rest rc=$?" >> ${O}_box.txt
timeout 1500 python bench.py --steps 3 --warmup 2 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" >> ${O}_box.txt
timeout 1500 python scripts/server_sweep.py --stock --points 8:128:2:8,16:128:2:8,16:256:2:8,16:128:3:8,16:64:2:8,12:128:2:4,16:128:1:8,16:128:2:12 > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" >> ${O}_box.txt
tail -4 ${O}_pytest_dp.log; tail -4 ${O}_pytest_rest.log; cat ${O}_box.txt; cut -c1-260 ${O}_sweep.jsonl
