#!/bin/bash
# GPU call 36 (1 GPU): live topology edits (picking sequence) through the C ABI against the oracle
timeout 600 python -m pytest tests/test_live_edits.py -m gpu -q -x 2>&1 | tail -25
