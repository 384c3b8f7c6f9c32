#!/bin/bash
# session r4k: the numbers of the tree with the hand-lowered hash, the new FASTX parser and BLG_F_TIGHT: GPU suite, bench line,
# launch list of the bench command, kernel-rate tools
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -x -q -m gpu ) > gpurun_out/r4k_pytest.log 2>&1
tail -n 5 gpurun_out/r4k_pytest.log
python bench.py > gpurun_out/r4k_bench.json 2> gpurun_out/r4k_bench.err; tail -c 300 gpurun_out/r4k_bench.err
python bench.py --impl reference > gpurun_out/r4k_bench_reference.json 2>> gpurun_out/r4k_bench.err
python bench.py --steps 2 --warmup 1 > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r4k_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r4k_ncu.log 2>&1
wc -l gpurun_out/r4k_launches.csv
{
for a in "--k 31" "--k 63" "--k 31 --syncmer 19" "--k 31 --syncmer 19 --canon 0" "--k 21 --syncmer 11" "--k 31 --syncmer 15" "--k 31 --hash-sample 0.1"; do
  timeout 200 python tools/kmerbench.py $a 2>&1 | tail -n 1
done
} > gpurun_out/r4k_kmerbench.txt 2>&1; cat gpurun_out/r4k_kmerbench.txt | cut -c1-200
python tools/auxbench.py 2>&1 | tail -n 6 > gpurun_out/r4k_auxbench.txt
python tools/nextbench.py > gpurun_out/r4k_nextbench.json 2> gpurun_out/r4k_nextbench.err
python tools/fastx_probe.py 2>&1 | tail -n 2 > gpurun_out/r4k_fastx.txt
./tools/ubench 2>&1 | tail -n 4 > gpurun_out/r4k_ubench_hash.txt; cat gpurun_out/r4k_ubench_hash.txt | cut -c1-250
python tools/run_configs.py > gpurun_out/r4k_configs_n1.jsonl 2> gpurun_out/r4k_configs.err; tail -c 400 gpurun_out/r4k_configs.err; wc -l gpurun_out/r4k_configs_n1.jsonl
