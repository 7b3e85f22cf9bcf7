#!/bin/bash
# usage (under gpurun): bash tools/sweep_chunks.sh   -- per-kernel ms at 20000 utterances for launch-group / tile sizes
B="python bench.py --utts 20000 --steps 3 --warmup 3 --no-cpu-baseline --e2e-fraction 64"
run() { echo -n "group=$1 rmax=$2 : "; SPB_GROUP_ROWS=$1 SPB_BIG_RMAX=$2 $B 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']
print(round(d['value'],1), round(d['ms_per_step'],3), {k:round(v,3) for k,v in r['kernel_ms_per_step'].items()}, r['launches_per_step']['features'], 'e2e', round(d['e2e']['value'],1))"; }
for c in 196608 1048576 2097152 4194304 8388608 33554432; do run $c 960; done
for r in 480 720; do run 4194304 $r; done
