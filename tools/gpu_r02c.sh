#!/bin/bash
# round 2: single-pass cmvnw first run
O=gpurun_out/r02c; mkdir -p $O
timeout -s KILL 600 python -m pytest tests/test_gpu_batch.py -x -q -k "single_pass or ragged or nan or large_tile" > $O/pytest_stream.log 2>&1
echo "rc=$?" >> $O/pytest_stream.log
timeout -s KILL 300 python tools/tc_time.py lmfe_cmvnw > $O/time_lmfe_cmvnw.log 2>&1
timeout -s KILL 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1
echo "rc=$?" >> $O/pytest_gpu.log
timeout -s KILL 900 python bench.py --steps 5 --warmup 3 --cpu-utts 256 > $O/bench_full.json 2> $O/bench_full.err
echo "rc=$?" >> $O/bench_full.err
for f in $O/pytest_stream.log $O/time_lmfe_cmvnw.log $O/pytest_gpu.log $O/bench_full.err; do echo "== $f"; tail -n 5 $f; done
