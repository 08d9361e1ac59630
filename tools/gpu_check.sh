#!/bin/bash
# regression check after the round's feature work: the whole -m gpu suite, smoke(), the default bench line and the fox run
mkdir -p gpurun_out
timeout 2700 python -m pytest tests -x -q -m gpu > gpurun_out/chk_pytest.txt 2>&1
tail -3 gpurun_out/chk_pytest.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/chk_smoke.txt 2>&1
tail -1 gpurun_out/chk_smoke.txt
( time timeout 1500 python bench.py > gpurun_out/chk_bench.json 2> gpurun_out/chk_bench.err ) 2> gpurun_out/chk_bench.time
tail -c 600 gpurun_out/chk_bench.json
timeout 900 python tools/bench_fox.py > gpurun_out/chk_fox.json 2> gpurun_out/chk_fox.err
tail -c 400 gpurun_out/chk_fox.json
