#!/bin/bash
# session r4v: C++ shim: streaming form of super_kmer_view (6-byte records, positions rebuilt by the iterator): test + rate
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "cpp_shim" 2>&1 | tail -3
g++ -std=c++17 -O2 -Iinclude tools/shim_bench.cpp -o gpurun_out/shim_bench -Lbiolib_b200 -lbiolib_gpu -Wl,-rpath,$PWD/biolib_b200 &&
./gpurun_out/shim_bench 2>&1 | tee gpurun_out/r4v_shim_bench.txt; rm -f gpurun_out/shim_bench
