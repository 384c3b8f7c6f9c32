#!/bin/bash
# tools/ab.sh v1 v2 ...  -- run quickbench for each variant library on the GPU box (under gpurun), log to gpurun_out/ab.log
mkdir -p gpurun_out
for v in "$@"; do
  BLG_LIB_PATH=biolib_b200/variants/lib$v.so timeout 300 python tools/quickbench.py $QB_ARGS 2>&1 | tail -2 | tee -a gpurun_out/ab.log
done
