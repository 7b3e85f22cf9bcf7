#!/bin/bash
# usage: bash tools/ncu_extra.sh <tag> <kernel-regex> [skip] [count]   (under gpurun) -- ncu over tools/bench_extra.py
TAG=$1; PAT=$2; SKIP=${3:-3}; CNT=${4:-1}
CMD="python tools/bench_extra.py"
$CMD > gpurun_out/${TAG}_plain.jsonl 2> gpurun_out/${TAG}_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:"$PAT" -s $SKIP -c $CNT -o gpurun_out/${TAG}_prof -f $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/${TAG}_ncu.log
