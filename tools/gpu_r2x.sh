#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_async.py tests/test_gpu_sk4.py -x -q -m gpu > gpurun_out/r2x_pytest.log 2>&1
tail -n 6 gpurun_out/r2x_pytest.log
python bench.py --steps 50 --warmup 3 > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err; tail -n 3 gpurun_out/r2x_bench.err; cut -c1-900 gpurun_out/r2x_bench.json
