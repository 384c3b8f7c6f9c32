#!/bin/bash
# session r4x: ncu --set full of kmer2_kernel (kmer_view, k = 31 canonical, 100 Mbp) on the final tree
mkdir -p gpurun_out
python tools/kmerbench.py --k 31 --steps 3 > gpurun_out/r4x_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:kmer2_kernel -s 2 -c 1 -o gpurun_out/r4x_kmer2 python tools/kmerbench.py --k 31 --steps 3 > gpurun_out/r4x_ncu.log 2>&1
tail -n 1 gpurun_out/r4x_plain.log gpurun_out/r4x_ncu.log | cut -c1-200
