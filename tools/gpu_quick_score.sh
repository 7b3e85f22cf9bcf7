#!/bin/bash
O=gpurun_out/${1:-r02u}; mkdir -p $O
timeout -s KILL 900 python -m pytest tests/test_gpu_scoring.py tests/test_zz_no_alloc.py tests/test_gpu_guards.py -x -q > $O/pytest_score.log 2>&1
echo "rc=$?" >> $O/pytest_score.log
timeout -s KILL 300 python tools/score_bench.py > $O/score_bench.txt 2>&1
tail -n 4 $O/pytest_score.log; cat $O/score_bench.txt
