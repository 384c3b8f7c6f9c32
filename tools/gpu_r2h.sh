#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2h_pytest_all.log 2>&1
tail -n 5 gpurun_out/r2h_pytest_all.log
{
for a in "--k 31 --m 19" "--k 21 --m 11" "--k 31 --m 19 --reads 150"; do
    echo -n "default: "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -n 1 | cut -c50-190
done
} > gpurun_out/r2h_ab.txt 2>&1
cat gpurun_out/r2h_ab.txt
python bench.py --steps 20 --warmup 3 > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err; tail -n 2 gpurun_out/r2h_bench.json | cut -c1-300
