#!/bin/bash
# usage: bash tools/ncu_kernel.sh <tag> <kernel-regex> [skip] [count]   (under gpurun)
TAG=$1; PAT=$2; SKIP=${3:-8}; CNT=${4:-2}
CMD="python bench.py --utts 4000 --steps 2 --warmup 3 --no-cpu-baseline --e2e-fraction 64"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:"$PAT" -s $SKIP -c $CNT -o gpurun_out/${TAG}_prof -f $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/${TAG}_ncu.log
