#!/bin/bash
# tools/abk.sh v1 v2 ...  -- tools/kmerbench.py for the default library and each variant library (run under gpurun)
for v in default "$@"; do
  if [ "$v" = default ]; then unset BLG_LIB_PATH; else export BLG_LIB_PATH=biolib_b200/variants/lib$v.so; fi
  for a in "--k 31" "--k 63" "--k 31 --breaks 0.01" "--k 31 --reads 150"; do
    echo -n "$v: "; timeout 200 python tools/kmerbench.py $a $KB_ARGS 2>&1 | tail -1 | cut -c1-160
  done
done
