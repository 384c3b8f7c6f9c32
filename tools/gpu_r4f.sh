#!/bin/bash
# session r4f: FASTX parser (8 KB blocks, chunked scan, no host copies) tests + times; tile-loop reorder variants
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ingest.py -x -q -m gpu 2>&1 | tail -3
python tools/fastx_probe.py 2>&1 | tail -3 | tee gpurun_out/r4f_fastx.txt
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:fx_ --csv --log-file gpurun_out/r4f_fx_launches.csv python tools/fastx_probe.py --reps 2 > /dev/null 2>&1
grep -o 'fx_[a-z]*\|"[0-9]*"$' gpurun_out/r4f_fx_launches.csv | paste - - | tail -12
cp biolib_b200/libbiolib_gpu.so biolib_b200/variants/libcur.so
for v in cur dbg ro2 ro3; do
  export BLG_LIB_PATH=biolib_b200/variants/lib$v.so
  for a in "--k 31 --m 19" "--k 31 --m 19 --reads 150"; do
    echo -n "$v: "; timeout 200 python tools/quickbench.py $a 2>&1 | tail -1 | cut -c50-200
  done
done | tee gpurun_out/r4f_ab.txt
