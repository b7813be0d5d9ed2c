#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_shim.py -m gpu -q -p no:cacheprovider 2>&1 | tail -15 | cut -c1-220
