#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ingest.py -x -q -m gpu > gpurun_out/r2k_pytest_ingest.log 2>&1
tail -n 25 gpurun_out/r2k_pytest_ingest.log
