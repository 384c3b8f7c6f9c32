#!/bin/bash
# session r4g: BLG_F_TIGHT (6-byte records): tests, bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tight or runs" 2>&1 | tail -5
python bench.py --steps 20 --warmup 3 > gpurun_out/r4g_bench.json 2> gpurun_out/r4g_bench.err
tail -c 800 gpurun_out/r4g_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r4g_bench.json'))
print("value", d["value"], "frac", d["roofline"]["frac"])
for k in d:
    if k.startswith("e2e"): print(k, round(d[k]["value"],2), d[k]["ms_per_step"], d[k]["h2d_bytes_per_step"], d[k]["d2h_bytes_per_step"])
PY
