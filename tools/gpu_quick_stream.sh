#!/bin/bash
O=gpurun_out/${1:-r02e}; mkdir -p $O
timeout -s KILL 600 python -m pytest tests/test_gpu_batch.py -x -q -k "single_pass or ragged or nan or large_tile" > $O/pytest_stream.log 2>&1
echo "rc=$?" >> $O/pytest_stream.log
timeout -s KILL 300 python tools/tc_time.py lmfe_cmvnw > $O/time_lmfe_cmvnw.log 2>&1
tail -n 4 $O/pytest_stream.log; cat $O/time_lmfe_cmvnw.log
