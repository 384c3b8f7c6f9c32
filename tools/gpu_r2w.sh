#!/bin/bash
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2w_bench_n2.json 2> gpurun_out/r2w_bench_n2.err
tail -n 3 gpurun_out/r2w_bench_n2.err; cat gpurun_out/r2w_bench_n2.json | cut -c1-3000
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2w_bench_ref_n2.json 2>&1
cat gpurun_out/r2w_bench_ref_n2.json | cut -c1-600
