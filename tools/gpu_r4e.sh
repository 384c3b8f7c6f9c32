#!/bin/bash
# session r4e: reordered tile loop A/B; FASTX parser kernel times
mkdir -p gpurun_out
cp biolib_b200/libbiolib_gpu.so biolib_b200/variants/libcur.so
for v in cur ro1 cur ro1; do
  export BLG_LIB_PATH=biolib_b200/variants/lib$v.so
  for a in "--k 31 --m 19" "--k 31 --m 19 --reads 150" "--k 63 --m 31" "--k 21 --m 11"; do
    echo -n "$v: "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -1 | cut -c50-200
  done
done | tee gpurun_out/r4e_ab.txt
unset BLG_LIB_PATH
python tools/fastx_probe.py 2>&1 | tail -3 | tee gpurun_out/r4e_fastx.txt
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:fx_ --csv --log-file gpurun_out/r4e_fx_launches.csv python tools/fastx_probe.py --reps 2 > /dev/null 2>&1
cut -d, -f5,12- gpurun_out/r4e_fx_launches.csv | tail -12
