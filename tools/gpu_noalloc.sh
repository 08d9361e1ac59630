#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -8 ) > gpurun_out/h_noalloc_alone.txt
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 ) > gpurun_out/h_pytest_all.txt
echo done
