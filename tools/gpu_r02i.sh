#!/bin/bash
O=gpurun_out/r02i; mkdir -p $O
timeout -s KILL 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1
echo "rc=$?" >> $O/pytest_gpu.log
timeout -s KILL 300 python tools/tc_time.py lmfe_cmvnw > $O/time_lmfe_cmvnw.log 2>&1
timeout -s KILL 900 python bench.py --steps 5 --warmup 3 > $O/bench_full.json 2> $O/bench_full.err
echo "rc=$?" >> $O/bench_full.err
timeout -s KILL 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
echo "rc=$?" >> $O/bench_ref.err
for f in $O/pytest_gpu.log $O/time_lmfe_cmvnw.log $O/bench_full.err $O/bench_ref.err; do echo "== $f"; tail -n 8 $f; done
