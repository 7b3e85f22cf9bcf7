#!/bin/bash
O=gpurun_out/${1:-r02m}; mkdir -p $O
timeout -s KILL 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_guards.py -x -q > $O/pytest_tc.log 2>&1
echo "rc=$?" >> $O/pytest_tc.log
timeout -s KILL 300 python tools/tc_time.py lmfe > $O/time_lmfe.log 2>&1
tail -n 4 $O/pytest_tc.log; cat $O/time_lmfe.log
