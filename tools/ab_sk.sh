#!/bin/bash
# tools/ab_sk.sh v1 v2 ... -- tools/quickbench.py over the benchmark families (single sequence and reads) for each variant library
for v in "$@"; do
  export BLG_LIB_PATH=biolib_b200/variants/lib$v.so
  for a in "--k 31 --m 19" "--k 31 --m 19 --reads 150" "--k 63 --m 31" "--k 21 --m 11" "--k 21 --m 11 --reads 20000" "--k 40 --m 9 --reads 150"; do
    echo -n "$v: "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -1 | cut -c50-175
  done
done
