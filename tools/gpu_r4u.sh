#!/bin/bash
# session r4u: full GPU suite + smoke + bench on the tree with blg_kmers BLG_F_RUNS
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -x -q -m gpu ) > gpurun_out/r4u_pytest.log 2>&1
tail -n 5 gpurun_out/r4u_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -n 1 | cut -c1-120
python bench.py > gpurun_out/r4u_bench.json 2> gpurun_out/r4u_bench.err; tail -c 300 gpurun_out/r4u_bench.err
python -c "
import json
d=json.load(open('gpurun_out/r4u_bench.json'))
print('value', d['value'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'parity', d['parity_checked'])
print({k: v.get('Gbases_per_s', v.get('ms_max_over_ranks')) for k, v in d['extra'].items()})"
