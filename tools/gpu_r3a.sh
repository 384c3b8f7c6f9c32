#!/bin/bash
mkdir -p gpurun_out
{
for v in default c6kb2 c5kb2; do
  if [ $v = default ]; then unset BLG_LIB_PATH; else export BLG_LIB_PATH=biolib_b200/variants/lib$v.so; fi
  for a in "--k 31 --m 19" "--k 21 --m 11" "--k 31 --m 19 --reads 150" "--k 31 --m 19 --breaks 0.001"; do
    echo -n "$v: "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -n 1 | cut -c50-190
  done
done
} > gpurun_out/r3a_ab.txt 2>&1
cat gpurun_out/r3a_ab.txt
