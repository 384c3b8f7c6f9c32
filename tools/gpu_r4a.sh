#!/bin/bash
# session r4a: whole GPU suite + bench line on the re-created container's tree
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -x -q -m gpu ) > gpurun_out/r4a_pytest.log 2>&1
tail -n 6 gpurun_out/r4a_pytest.log
python bench.py > gpurun_out/r4a_bench.json 2> gpurun_out/r4a_bench.err
tail -c 600 gpurun_out/r4a_bench.err; cat gpurun_out/r4a_bench.json
