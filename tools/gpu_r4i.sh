#!/bin/bash
# session r4i: ncu --set full capture of sk4_kernel on the tree with the hand-lowered hash
mkdir -p gpurun_out
python tools/quickbench.py --k 31 --m 19 --steps 3 > gpurun_out/r4i_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sk4_kernel -s 2 -c 1 -o gpurun_out/r4i_sk4 python tools/quickbench.py --k 31 --m 19 --steps 3 > gpurun_out/r4i_ncu.log 2>&1
tail -n 2 gpurun_out/r4i_plain.log gpurun_out/r4i_ncu.log
