#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_mtx_parse_gpu.py -q -m gpu -x -s 2>&1 | tail -12
timeout 600 python -m pytest tests -q -m gpu -x 2>&1 | tail -4
