#!/bin/bash
O=gpurun_out/${1:-r02r}; mkdir -p $O
timeout -s KILL 900 python -m pytest tests/test_gpu_batch.py tests/test_gpu_features.py tests/test_gpu_guards.py -x -q > $O/pytest_cm.log 2>&1
echo "rc=$?" >> $O/pytest_cm.log
timeout -s KILL 300 python tools/tc_time.py lmfe_cmvnw one > $O/time_lmfe_cmvnw.log 2>&1
tail -n 4 $O/pytest_cm.log; cat $O/time_lmfe_cmvnw.log
