#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_payloads.py -x -q -m gpu > gpurun_out/r2l_pytest_payloads.log 2>&1
tail -n 25 gpurun_out/r2l_pytest_payloads.log
