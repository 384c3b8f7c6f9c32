#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/r2o_pytest_all.log 2>&1
tail -n 8 gpurun_out/r2o_pytest_all.log
g++ -std=c++17 -O2 -Iinclude tools/shim_bench.cpp -o /tmp/shim_bench -Lbiolib_b200 -lbiolib_gpu -Wl,-rpath,$PWD/biolib_b200 && /tmp/shim_bench > gpurun_out/r2o_shim_bench.txt 2>&1
cat gpurun_out/r2o_shim_bench.txt
python tools/nextbench.py > gpurun_out/r2o_nextbench.json 2> gpurun_out/r2o_nextbench.err; tail -n 5 gpurun_out/r2o_nextbench.err; cat gpurun_out/r2o_nextbench.json
