#!/bin/bash
# round 2t: tests of the TMA forward (own file) + the whole GPU suite + default bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_unproject_tma.py -x -q 2>&1 | tail -15 | tee gpurun_out/r2t_tma_tests.log
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -4 | tee gpurun_out/r2t_pytest_gpu_tail.log
