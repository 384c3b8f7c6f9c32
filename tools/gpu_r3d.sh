#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r3d_pytest.log 2>&1
tail -n 4 gpurun_out/r3d_pytest.log
python tools/nextbench.py > gpurun_out/r3d_nextbench.json 2> gpurun_out/r3d_nextbench.err
tail -n 3 gpurun_out/r3d_nextbench.err
python bench.py --steps 10 --warmup 3 > gpurun_out/r3d_bench.json 2> gpurun_out/r3d_bench.err
tail -n 3 gpurun_out/r3d_bench.err
python -c "
import json
d=json.loads(open('gpurun_out/r3d_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['extra']['jaccard_k31_two_200mbp_genomes'])
n=json.load(open('gpurun_out/r3d_nextbench.json'))
print({k:round(v['ms'],2) for k,v in n.items() if 'e2e_k31' in k})
"
