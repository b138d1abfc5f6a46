#!/bin/bash
# the no-allocation ingest entry point: alone in a fresh process, then at the end of the whole GPU suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | tail -15
timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -6
