#!/bin/bash
# sy4 (syncmer filter in the sk4 frame): one full ncu capture of the canonical kernel on the 100 Mbp workload
mkdir -p gpurun_out
python tools/kmerbench.py --k 31 --syncmer 19 --steps 3 > gpurun_out/r3e_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sy4_kernel -s 2 -c 1 -o gpurun_out/r3e_sy4 python tools/kmerbench.py --k 31 --syncmer 19 --steps 3 > gpurun_out/r3e_ncu.log 2>&1
tail -n 2 gpurun_out/r3e_plain.log gpurun_out/r3e_ncu.log
