#!/bin/bash
# session r4t: BLG_F_RUNS for blg_kmers: tests + end-to-end time
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "kmers" 2>&1 | tail -4
python tools/nextbench.py > gpurun_out/r4t_nextbench.json 2> gpurun_out/r4t_nextbench.err; tail -n 3 gpurun_out/r4t_nextbench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r4t_nextbench.json'))
for k in ("kmers_runs_e2e_k31","kmers_e2e_k31"): print(k, d[k])
PY
