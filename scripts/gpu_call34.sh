#!/bin/bash
# GPU call 34 (1 GPU): BDF2 / static Newton / quasi-static Newton parity
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_implicit.py -m gpu -q > gpurun_out/pytest34.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest34.log; tail -25 gpurun_out/pytest34.log
