#!/bin/bash
mkdir -p gpurun_out
{
for a in "--k 31 --m 19" "--k 21 --m 11" "--k 31 --m 19 --reads 150"; do
    echo -n "default(mb5): "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -n 1 | cut -c50-190
done
} > gpurun_out/r2e_ab.txt 2>&1
cat gpurun_out/r2e_ab.txt
python tools/quickbench.py --k 21 --m 11 --steps 3 > gpurun_out/r2e_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sk4_kernel -s 2 -c 1 -o gpurun_out/r2e_sk4 python tools/quickbench.py --k 21 --m 11 --steps 3 > gpurun_out/r2e_ncu.log 2>&1
tail -n 2 gpurun_out/r2e_ncu.log
