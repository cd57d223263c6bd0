#!/bin/bash
set -x
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/t_noalloc_alone.log 2>&1; tail -3 gpurun_out/t_noalloc_alone.log
python -m pytest tests -x -q -m gpu > gpurun_out/t_full2.log 2>&1; tail -3 gpurun_out/t_full2.log
