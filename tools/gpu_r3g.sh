#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_jaccard.py tests/test_multi_gpu.py tests/test_payloads.py -m gpu -x -q 2>&1 | tail -n 3
python tools/jaccard_probe.py --reps 3 > gpurun_out/r3g_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r3g_launches.csv python tools/nextbench.py > gpurun_out/r3g_ncu.log 2>&1
cat gpurun_out/r3g_plain.log; tail -n 2 gpurun_out/r3g_ncu.log
