#!/bin/bash
mkdir -p gpurun_out
{
for v in noexact noexact5 noexact5fma; do
  export BLG_LIB_PATH=biolib_b200/variants/lib$v.so
  for a in "--k 31 --m 19" "--k 21 --m 11" "--k 31 --m 19 --reads 150"; do
    echo -n "$v: "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -1 | cut -c50-190
  done
done
} > gpurun_out/r2c_ab.txt 2>&1
cat gpurun_out/r2c_ab.txt
export BLG_LIB_PATH=biolib_b200/variants/libnoexact5.so
python tools/quickbench.py --k 31 --m 19 --steps 3 > gpurun_out/r2c_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sk4_kernel -s 2 -c 1 -o gpurun_out/r2c_sk4 python tools/quickbench.py --k 31 --m 19 --steps 3 > gpurun_out/r2c_ncu.log 2>&1
tail -n 2 gpurun_out/r2c_ncu.log
