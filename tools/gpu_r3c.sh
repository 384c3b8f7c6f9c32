#!/bin/bash
mkdir -p gpurun_out
{
for v in default sy5; do
  if [ $v = default ]; then unset BLG_LIB_PATH; else export BLG_LIB_PATH=biolib_b200/variants/lib$v.so; fi
  for a in "--k 31 --syncmer 19" "--k 31 --syncmer 19 --canon 0" "--k 21 --syncmer 11" "--k 21 --syncmer 11 --canon 0"; do
    echo -n "$v: "; timeout 200 python tools/kmerbench.py $a 2>&1 | tail -n 1 | cut -c1-150
  done
done
} > gpurun_out/r3c_ab.txt 2>&1
cat gpurun_out/r3c_ab.txt
