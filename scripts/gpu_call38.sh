#!/bin/bash
# GPU call 38 (1 GPU): headless runner on loaded meshes against the oracle
timeout 600 python -m pytest tests/test_meshio.py -m gpu -q 2>&1 | tail -30
