#!/bin/bash
# GPU job 15: round-2 ncu evidence of bench.py's own command (after it has exited 0 without ncu), and the reference arm.
mkdir -p gpurun_out
O=gpurun_out/j15
CMD="python bench.py --chain-only --steps 1 --warmup 1"
timeout 240 $CMD > ${O}_chain.json 2> ${O}_chain.err; echo "chain rc=$?" > ${O}_box.txt
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv \
   --log-file ${O}_launches.csv $CMD > ${O}_ncu1.log 2>&1; echo "ncu launches rc=$?" >> ${O}_box.txt
timeout 400 ncu --set full --import-source on --clock-control none -k 'regex:k_dp_(fill|backtrace)' --launch-skip 2 -c 2 -f -o ${O}_dp_full \
   $CMD > ${O}_ncu2.log 2>&1; echo "ncu full rc=$?" >> ${O}_box.txt
timeout 240 python bench.py --impl reference --steps 3 --warmup 3 > ${O}_bench_reference.json 2> ${O}_bench_reference.err; echo "reference rc=$?" >> ${O}_box.txt
cat ${O}_box.txt; ls -la gpurun_out/ | grep j15; tail -3 ${O}_ncu2.log; cut -c1-400 ${O}_bench_reference.json
