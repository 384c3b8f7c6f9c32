#!/bin/bash
# session r4m: chunked host path with a smaller LAST piece (PIPE_LAST_DIV), 6-byte records
mkdir -p gpurun_out
cp biolib_b200/libbiolib_gpu.so biolib_b200/variants/libcur.so
for v in cur ld2 ld4 ld8 cur; do
  export BLG_LIB_PATH=biolib_b200/variants/lib$v.so
  for sh in "--chunks 12 --divs 4" "--chunks 8 --divs 4" "--chunks 6 --divs 4"; do
    echo -n "$v: "; python tools/e2e.py --layouts tight $sh --reps 12 2>&1 | tail -1 | cut -c1-110
  done
done | tee gpurun_out/r4m_last_piece.txt
