#!/bin/bash
timeout 600 python -m pytest tests/test_padded_rows_gpu.py tests/test_parity_gpu.py tests/test_scene_gpu.py tests/test_abi.py tests/test_fused_gpu.py tests/test_io_gpu.py -m gpu -x -q --timeout 200 2>&1 | tail -2
LATS="96,96,96 112,112,112 100,100,100 120,120,120" bash tools/r02_small3d.sh
