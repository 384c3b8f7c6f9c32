#!/bin/bash
# session r4d: hand-lowered hash in every kernel + rewritten FASTX parser: GPU suite, kernel rates, next-row rates
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -x -q -m gpu ) > gpurun_out/r4d_pytest.log 2>&1
tail -n 6 gpurun_out/r4d_pytest.log
for a in "--k 31 --m 19" "--k 31 --m 19 --reads 150" "--k 63 --m 31" "--k 21 --m 11" "--k 21 --m 11 --reads 20000" "--k 40 --m 9 --reads 150" "--k 25 --m 22"; do
  timeout 200 python tools/quickbench.py $a 2>&1 | tail -1 | cut -c50-200
done | tee gpurun_out/r4d_quickbench.txt
python tools/auxbench.py > gpurun_out/r4d_auxbench.txt 2>&1; tail -n 12 gpurun_out/r4d_auxbench.txt
python tools/nextbench.py > gpurun_out/r4d_nextbench.json 2> gpurun_out/r4d_nextbench.err; head -c 600 gpurun_out/r4d_nextbench.json; tail -n 3 gpurun_out/r4d_nextbench.err
