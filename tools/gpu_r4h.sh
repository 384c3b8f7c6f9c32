#!/bin/bash
# session r4h: tight-layout tests again; pipeline shapes with the tight layout
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tight or runs" 2>&1 | tail -3
python tools/e2e.py --layouts tight --chunks 3,4,6,8,12,16 --divs 1,2,4,8 2>&1 | tee gpurun_out/r4h_e2e_shapes.txt | cut -c1-120
