#!/bin/bash
# session r4p: sk4 / sy4 with the next tile's ASCII staged by one cp.async.bulk per tile (SK4_BULK) against the register prefetch
mkdir -p gpurun_out
cp biolib_b200/libbiolib_gpu.so biolib_b200/variants/libcur.so
export BLG_LIB_PATH=biolib_b200/variants/libbulk.so
timeout 600 python -m pytest tests/test_gpu_sk4.py tests/test_gpu_parity.py -x -q -m gpu -k "not pipelined and not fullsize" 2>&1 | tail -3
for v in cur bulk cur bulk; do
  export BLG_LIB_PATH=biolib_b200/variants/lib$v.so
  for a in "--k 31 --m 19" "--k 31 --m 19 --reads 150" "--k 63 --m 31" "--k 21 --m 11"; do
    echo -n "$v: "; timeout 120 python tools/quickbench.py $a 2>&1 | tail -1 | cut -c50-200
  done
  echo -n "$v: "; timeout 120 python tools/kmerbench.py --k 31 --syncmer 19 2>&1 | tail -n 1 | cut -c1-150
done | tee gpurun_out/r4p_bulk_ab.txt
