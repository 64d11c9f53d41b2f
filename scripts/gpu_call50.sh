#!/bin/bash
# GPU call 50 (1 GPU): the tests not covered by call 49 after the arena / gather-list change, and smoke()
timeout 300 python -m pytest tests/test_meshio.py tests/test_gpu_facade.py tests/test_gpu_elements.py -m gpu -q 2>&1 | tail -3
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
