#!/bin/bash
# Round-2 measurement pass on one B200: GPU tests, bench (both arms), launch list, ncu --set full of the hot kernels.
TAG=${1:-r02q}; O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > $O/smi.txt 2>&1
timeout -s KILL 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
timeout -s KILL 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "rc=$?" >> $O/bench_ref.err
timeout -s KILL 900 python bench.py --steps 10 --warmup 3 > $O/bench_full.json 2> $O/bench_full.err; echo "rc=$?" >> $O/bench_full.err
timeout -s KILL 300 python tools/tc_time.py lmfe > $O/time_lmfe.log 2>&1
timeout -s KILL 300 python tools/tc_time.py lmfe_cmvnw > $O/time_lmfe_cmvnw.log 2>&1
timeout -s KILL 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > $O/ncu_launches.log 2>&1
for K in features_tc2 cmvnw_resident cmvnw_big cosine_gemm; do
  timeout -s KILL 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 2 -o $O/ncu_$K -f python tools/ncu_workload.py > $O/ncu_$K.log 2>&1
  echo "ncu $K rc=$?"
done
for f in $O/pytest_gpu.log $O/bench_ref.err $O/bench_full.err $O/time_lmfe.log $O/time_lmfe_cmvnw.log; do echo "== $f"; tail -n 7 $f; done
