#!/bin/bash
# GPU job 7: sampling profile of the batched server's threads (where do the host threads' 44 us per read go).
mkdir -p gpurun_out
O=gpurun_out/j7
gcc -O1 -shared -fPIC -o /tmp/sprof.so scripts/tools/sprof.c -ldl || exit 1
timeout 300 python scripts/server_sweep.py --passes 2 --pass-timeout 60 \
  --points 16:128:1:8:LD_PRELOAD=/tmp/sprof.so:SPROF_OUT=/tmp/sprof.txt > ${O}_sweep.jsonl 2> ${O}_sweep.err; echo "sweep rc=$?" > ${O}_box.txt
cut -f1,2 /tmp/sprof.txt | sort | uniq -c | sort -rn | head -20
sort /tmp/sprof.txt | uniq -c | sort -rn | head -3000 > ${O}_sprof_top.txt
wc -l /tmp/sprof.txt; cut -c1-300 ${O}_sweep.jsonl; tail -3 ${O}_sweep.err
