#!/bin/bash
# session r4c: A/B of hash lowering variants (SK4_HASH_CHAIN) over three benchmark families
mkdir -p gpurun_out
for v in "$@"; do
  export BLG_LIB_PATH=biolib_b200/variants/lib$v.so
  for a in "--k 31 --m 19" "--k 31 --m 19 --reads 150" "--k 63 --m 31" "--k 21 --m 11"; do
    echo -n "$v: "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -1 | cut -c1-200
  done
done | tee gpurun_out/r4c_ab.txt
