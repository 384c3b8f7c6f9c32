#!/bin/bash
# tools/build_variant.sh NAME "EXTRA nvcc flags"  ->  biolib_b200/variants/libNAME.so (development A/B builds)
set -e
cd "$(dirname "$0")/../biolib_b200/csrc"
make -j8 BUILD=build_$1 OUT=../variants/lib$1.so EXTRA="$2" 2>&1 | grep -E "error|sk2_kernel|Used" | head -8 || true
ls -la ../variants/lib$1.so
