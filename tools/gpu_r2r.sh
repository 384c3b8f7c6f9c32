#!/bin/bash
mkdir -p gpurun_out
{
for v in default k2tma; do
  if [ $v = default ]; then unset BLG_LIB_PATH; else export BLG_LIB_PATH=biolib_b200/variants/lib$v.so; fi
  echo "== $v"
  timeout 300 python tools/kmerbench.py 2>&1 | tail -n 12
done
export BLG_LIB_PATH=biolib_b200/variants/libk2tma.so
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "kmers or config1 or syncmer or hash_sampled" 2>&1 | tail -n 3
} > gpurun_out/r2r_k2tma.txt 2>&1
cat gpurun_out/r2r_k2tma.txt
