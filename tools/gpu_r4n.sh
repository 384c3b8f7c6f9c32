#!/bin/bash
# session r4n: ncu --set full of the two FASTX text passes (FASTA input)
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:fx_s -c 3 -o gpurun_out/r4n_fx python tools/fastx_probe.py --reps 1 > gpurun_out/r4n_ncu.log 2>&1
tail -n 3 gpurun_out/r4n_ncu.log
