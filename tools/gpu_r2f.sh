#!/bin/bash
mkdir -p gpurun_out
{
for v in nb1 nb10_2; do
  export BLG_LIB_PATH=biolib_b200/variants/lib$v.so
  for a in "--k 31 --m 19" "--k 21 --m 11"; do
    echo -n "$v: "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -n 1 | cut -c50-190
  done
done
} > gpurun_out/r2f_ab.txt 2>&1
cat gpurun_out/r2f_ab.txt
