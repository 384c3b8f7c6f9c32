#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 > gpurun_out/r3m_bench_plain.json 2> gpurun_out/r3m_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r3m_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r3m_ncu.log 2>&1
tail -n 2 gpurun_out/r3m_plain.err; tail -c 300 gpurun_out/r3m_ncu.log; wc -l gpurun_out/r3m_launches.csv
