#!/bin/bash
mkdir -p gpurun_out
python tools/jaccard_probe.py --reps 3 > gpurun_out/r3f_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r3f_launches.csv python tools/jaccard_probe.py --reps 1 > gpurun_out/r3f_ncu.log 2>&1
cat gpurun_out/r3f_plain.log; tail -n 2 gpurun_out/r3f_ncu.log
