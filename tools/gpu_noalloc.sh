#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/noalloc_alone.log 2>&1
echo "alone rc=$?"; tail -5 gpurun_out/noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/noalloc_full.log 2>&1
echo "full rc=$?"; tail -5 gpurun_out/noalloc_full.log
