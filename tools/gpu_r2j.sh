#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_fullsize.py > gpurun_out/r2j_pytest.log 2>&1
tail -n 5 gpurun_out/r2j_pytest.log
timeout 900 python -m pytest tests/test_gpu_fullsize.py -x -q -m gpu > gpurun_out/r2j_pytest_fullsize.log 2>&1
tail -n 5 gpurun_out/r2j_pytest_fullsize.log
python bench.py --steps 20 --warmup 3 > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err; tail -n 3 gpurun_out/r2j_bench.err; cat gpurun_out/r2j_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2j_bench_ref.json 2>&1; cat gpurun_out/r2j_bench_ref.json
