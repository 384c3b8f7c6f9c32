#!/bin/bash
mkdir -p gpurun_out
python tools/timing_probe.py > gpurun_out/r2i_probe.txt 2>&1
cat gpurun_out/r2i_probe.txt
python tools/quickbench.py --k 31 --m 19 --steps 3 > gpurun_out/r2i_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2i_launches.csv python tools/quickbench.py --k 31 --m 19 --steps 3 > gpurun_out/r2i_ncu.log 2>&1
grep -i "sk4" gpurun_out/r2i_launches.csv | tail -n 4 | cut -c1-60,200-400
