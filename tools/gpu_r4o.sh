#!/bin/bash
# session r4o: FASTX block scans without serial walks over the warps: tests + times
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ingest.py -x -q -m gpu 2>&1 | tail -2
python tools/fastx_probe.py 2>&1 | tail -2 | tee gpurun_out/r4o_fastx.txt
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:fx_ --csv --log-file gpurun_out/r4o_fx_launches.csv python tools/fastx_probe.py --reps 1 > /dev/null 2>&1
grep -o 'fx_[a-z]*\|"[0-9]*"$' gpurun_out/r4o_fx_launches.csv | paste - - | tail -6
