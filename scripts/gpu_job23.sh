#!/bin/bash
# GPU job 23: the multi-zone seed kernel against the reference's dumps and the oracle; the -N 1 regression cases through the servers.
mkdir -p gpurun_out
timeout 70 python -m pytest tests/test_zy_gpu_seed_mm.py -x -q > gpurun_out/j23_seedmm.log 2>&1; echo "seedmm rc=$?"
timeout 50 python -m pytest tests/test_gpu_dropin.py -x -q -k "table and (052 or 053 or 054 or 055 or 056 or 092 or 147)" > gpurun_out/j23_dropin.log 2>&1; echo "dropin rc=$?"
tail -15 gpurun_out/j23_seedmm.log | cut -c1-300; tail -5 gpurun_out/j23_dropin.log | cut -c1-300
