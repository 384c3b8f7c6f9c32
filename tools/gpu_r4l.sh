#!/bin/bash
# session r4l: the bench line under torchrun (N = number of GPUs of the call), as the driver launches it; both arms
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/r4l_bench_reference_n$N.json 2> gpurun_out/r4l_ref.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r4l_bench_n$N.json 2> gpurun_out/r4l_bench.err
tail -c 400 gpurun_out/r4l_bench.err; head -c 300 gpurun_out/r4l_bench_reference_n$N.json; echo
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r4l_bench_n$N.json') if l.startswith('{')][-1])
print("value", d["value"], "ms", d["ms_per_step"], "n", d["n_gpus"])
for k in d:
    if k.startswith("e2e"): print(k, round(d[k]["value"],2), round(d[k]["ms_per_step"],3))
for k,v in (d.get("extra") or {}).items(): print(k, {a:b for a,b in v.items() if a in ("Gbases_per_s","ms_max_over_ranks","kernel_ms_max_over_ranks","G_kmers_per_s")})
PY
