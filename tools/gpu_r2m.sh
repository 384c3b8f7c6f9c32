#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_jaccard.py -x -q -m gpu > gpurun_out/r2m_pytest_jaccard.log 2>&1
tail -n 30 gpurun_out/r2m_pytest_jaccard.log
