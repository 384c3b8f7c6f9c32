#!/bin/bash
# round 2, call A: microbenchmarks, sk4 parity, variant A/B
mkdir -p gpurun_out
./tools/ubench > gpurun_out/r2a_ubench.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_sk4.py -x -q -m gpu > gpurun_out/r2a_pytest_sk4.log 2>&1
tail -15 gpurun_out/r2a_pytest_sk4.log
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2a_pytest_all.log 2>&1
tail -5 gpurun_out/r2a_pytest_all.log
{
for v in default fma kb2 mb5 mb5fma mb6kb2 nb1; do
  if [ $v = default ]; then unset BLG_LIB_PATH; else export BLG_LIB_PATH=biolib_b200/variants/lib$v.so; fi
  for a in "--k 31 --m 19" "--k 21 --m 11" "--k 31 --m 19 --reads 150"; do
    echo -n "$v: "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -1 | cut -c50-190
  done
done
unset BLG_LIB_PATH
export BLG_SK=3
for a in "--k 31 --m 19" "--k 21 --m 11" "--k 31 --m 19 --reads 150"; do
  echo -n "sk3: "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -1 | cut -c50-190
done
} > gpurun_out/r2a_ab.txt 2>&1
cat gpurun_out/r2a_ab.txt
