#!/bin/bash
# session r4b: ncu --set full capture of sk4_kernel (source page imported) on the current tree
mkdir -p gpurun_out
python tools/quickbench.py --k 31 --m 19 --steps 3 > gpurun_out/r4b_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sk4_kernel -s 2 -c 1 -o gpurun_out/r4b_sk4 python tools/quickbench.py --k 31 --m 19 --steps 3 > gpurun_out/r4b_ncu.log 2>&1
tail -n 2 gpurun_out/r4b_plain.log gpurun_out/r4b_ncu.log
python tools/nextbench.py > gpurun_out/r4b_nextbench.json 2> gpurun_out/r4b_nextbench.err; tail -c 1500 gpurun_out/r4b_nextbench.json
