#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sk4.py -x -q -m gpu > gpurun_out/r2v_pytest_sk4.log 2>&1
tail -n 12 gpurun_out/r2v_pytest_sk4.log
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/r2v_pytest_all.log 2>&1
tail -n 5 gpurun_out/r2v_pytest_all.log
{
for a in "--k 31 --m 19" "--k 21 --m 11" "--k 63 --m 31" "--k 63 --m 31 --n 1000000000 --steps 3" "--k 31 --m 19 --reads 150"; do
    echo -n "default: "; timeout 300 python tools/quickbench.py $a 2>&1 | tail -n 1 | cut -c50-190
done
} > gpurun_out/r2v_ab.txt 2>&1
cat gpurun_out/r2v_ab.txt
