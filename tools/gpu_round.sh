#!/bin/bash
# One gpurun call: parity tests, bench lines, ncu launch list and full captures.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_round.sh <tag>'
set -u
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/${TAG}_smoke.log
python bench.py --steps 5 --warmup 3 > $O/${TAG}_bench_full.json 2> $O/${TAG}_bench_full.err; echo "bench rc=$?"
tail -c 3000 $O/${TAG}_bench_full.json
python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "ref rc=$?"
SMALL="python bench.py --utts 4000 --steps 2 --warmup 3 --no-cpu-baseline --e2e-fraction 64"
$SMALL > $O/${TAG}_bench_small.json 2> $O/${TAG}_bench_small.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv $SMALL > $O/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
$SMALL > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'features_tc|features_fast512|cmvnw_big_kernel|cosine_gemm' -s 6 -c 6 -o $O/${TAG}_prof -f $SMALL > $O/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $O
python tools/score_bench.py > $O/${TAG}_score_bench.txt 2>&1; echo "score bench rc=$?"
python bench.py --utts 20000 --steps 3 --warmup 3 --no-cpu-baseline --e2e-fraction 64 > $O/${TAG}_bench_20k.json 2>&1
