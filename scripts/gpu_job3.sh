#!/bin/bash
# GPU job 3 (bounded at every level): the no-allocation entry point's test alone and after other tests; a sweep of the
# batched server's knobs after the result-flush fix (incl. the 4-client case that deadlocked); the bench.
mkdir -p gpurun_out
O=gpurun_out/j3
timeout 240 python -m pytest tests/test_zz_no_alloc.py -q > ${O}_noalloc_alone.log 2>&1; echo "noalloc_alone rc=$?" > ${O}_box.txt
timeout 400 python -m pytest tests/test_gpu_ungapped.py tests/test_gpu_dp.py tests/test_zz_no_alloc.py -q > ${O}_noalloc_after.log 2>&1; echo "noalloc_after rc=$?" >> ${O}_box.txt
timeout 780 python scripts/server_sweep.py --stock --passes 1 --pass-timeout 60 \
   --points 16:128:2:8,16:128:2:4,16:512:2:8,32:256:2:8,32:128:1:8,8:1024:2:8 > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" >> ${O}_box.txt
timeout 600 python bench.py --steps 2 --warmup 1 --pass-timeout 120 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" >> ${O}_box.txt
tail -2 ${O}_noalloc_alone.log; tail -2 ${O}_noalloc_after.log; cat ${O}_box.txt; cut -c1-330 ${O}_sweep.jsonl; tail -3 ${O}_bench.err; cut -c1-600 ${O}_bench.json
