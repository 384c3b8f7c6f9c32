#!/bin/bash
mkdir -p gpurun_out
python tools/jaccard_probe.py --reps 2 > gpurun_out/r3i_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rs_scatter -s 20 -c 1 -o gpurun_out/r3i_rs_scatter python tools/jaccard_probe.py --reps 2 > gpurun_out/r3i_ncu.log 2>&1
tail -n 2 gpurun_out/r3i_plain.log gpurun_out/r3i_ncu.log
