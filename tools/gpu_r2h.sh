#!/bin/bash
# round 2h: the no-alloc entry point (tests/test_zz_no_alloc.py) alone in a fresh process, then the whole GPU suite
# (which ends with it), smoke(), and one default bench run.
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -s 2>&1 | tail -8 | tee gpurun_out/r2h_noalloc_alone.log
python -m pytest tests -x -q -m gpu 2>&1 | tail -6 | tee gpurun_out/r2h_pytest_gpu_tail.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 | tee gpurun_out/r2h_smoke.log
python bench.py > gpurun_out/r2h_bench_cfg3.json 2> gpurun_out/r2h_bench_cfg3.err; tail -c 600 gpurun_out/r2h_bench_cfg3.json
