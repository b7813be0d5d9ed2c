#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_regimes.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -12 | cut -c1-220
