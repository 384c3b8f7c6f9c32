#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "syncmers or kmers" > gpurun_out/r3b_pytest.log 2>&1
tail -n 6 gpurun_out/r3b_pytest.log
{
for a in "--k 31 --syncmer 19" "--k 31 --syncmer 19 --canon 0" "--k 21 --syncmer 11" "--k 21 --syncmer 11 --canon 0" "--k 31 --syncmer 19 --reads 150" "--k 31 --syncmer 19 --breaks 0.001" "--k 31 --syncmer 15"; do
  timeout 200 python tools/kmerbench.py $a 2>&1 | tail -n 1
done
} > gpurun_out/r3b_kmerbench.txt 2>&1
cat gpurun_out/r3b_kmerbench.txt
