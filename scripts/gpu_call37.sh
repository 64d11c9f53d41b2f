#!/bin/bash
# GPU call 37 (1 GPU): facade picking scenario + live edits
timeout 600 python -m pytest tests/test_gpu_facade.py tests/test_live_edits.py -m gpu -q 2>&1 | tail -25
