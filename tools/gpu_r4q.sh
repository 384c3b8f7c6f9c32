#!/bin/bash
# session r4q: final tree: GPU suite, smoke, bench line, launch list, ncu --set full of sk4 (bulk-staged)
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -x -q -m gpu ) > gpurun_out/r4q_pytest.log 2>&1
tail -n 5 gpurun_out/r4q_pytest.log
python __graft_entry__.py smoke > gpurun_out/r4q_smoke.log 2>&1; tail -n 1 gpurun_out/r4q_smoke.log
python bench.py > gpurun_out/r4q_bench.json 2> gpurun_out/r4q_bench.err; tail -c 300 gpurun_out/r4q_bench.err
python bench.py --impl reference > gpurun_out/r4q_bench_reference.json 2>> gpurun_out/r4q_bench.err
python bench.py --steps 2 --warmup 1 > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r4q_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r4q_ncu.log 2>&1
python tools/quickbench.py --k 31 --m 19 --steps 3 > gpurun_out/r4q_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sk4_kernel -s 2 -c 1 -o gpurun_out/r4q_sk4 python tools/quickbench.py --k 31 --m 19 --steps 3 > gpurun_out/r4q_ncu2.log 2>&1
python - <<'PY'
import json
d=json.load(open('gpurun_out/r4q_bench.json'))
print("value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "kernel_ms", d["roofline"]["kernel_ms"])
for k in d:
    if k.startswith("e2e"): print(k, round(d[k]["value"],2), round(d[k]["ms_per_step"],3))
PY
