#!/bin/bash
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 50 --warmup 3 > gpurun_out/r2y_bench_n8.json 2> gpurun_out/r2y_bench_n8.err
tail -n 2 gpurun_out/r2y_bench_n8.err; cut -c1-400 gpurun_out/r2y_bench_n8.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 4 --steps 50 --warmup 3 --no-extra > gpurun_out/r2y_bench_n4.json 2> gpurun_out/r2y_bench_n4.err
cut -c1-300 gpurun_out/r2y_bench_n4.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 tools/run_configs.py 5 > gpurun_out/r2y_config5_n8.jsonl 2> gpurun_out/r2y_config5_n8.err
cat gpurun_out/r2y_config5_n8.jsonl | cut -c1-600
