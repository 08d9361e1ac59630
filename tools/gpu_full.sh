#!/bin/bash
# the whole -m gpu suite, as the driver runs it at round end
mkdir -p gpurun_out
timeout 2700 python -m pytest tests -x -q -m gpu > gpurun_out/full_pytest.txt 2>&1
tail -5 gpurun_out/full_pytest.txt
