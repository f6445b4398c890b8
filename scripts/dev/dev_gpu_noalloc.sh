#!/bin/bash
mkdir -p gpurun_out
{
echo "=== alone"; timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -6
echo "=== full suite"; timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
} 2>&1 | tee gpurun_out/noalloc.log
