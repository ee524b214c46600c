#!/bin/bash
# GPU job 4 (bounded at every level): where does the default bench spend its time (progress lines), and where does the
# batched server's time go (us_* counters of the shim) at two worker counts.
mkdir -p gpurun_out
O=gpurun_out/j4
timeout 480 python bench.py --steps 2 --warmup 1 --pass-timeout 90 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?" > ${O}_box.txt
timeout 240 python scripts/server_sweep.py --passes 1 --pass-timeout 60 --points 16:128:2:8,16:512:2:8 > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" >> ${O}_box.txt
timeout 240 python bench.py --steps 2 --warmup 1 --chain-only > ${O}_chain.json 2> ${O}_chain.err; echo "chain rc=$?" >> ${O}_box.txt
cat ${O}_box.txt; tail -25 ${O}_bench.err; cut -c1-1500 ${O}_bench.json; cut -c1-1800 ${O}_sweep.jsonl; tail -12 ${O}_chain.err
