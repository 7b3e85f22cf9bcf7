#!/bin/bash
# ncu --set full capture of one launch of a kernel (regex $1) inside tools/tc_time.py-like workload
K=${1:-cmvnw_stream}; TAG=${2:-r02d}; O=gpurun_out/$TAG; mkdir -p $O
cat > /tmp/one_call.py <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, ".")
from speaker_verification_b200 import _lib, batch
rng = np.random.default_rng(0)
U, N = 4874, 128000
big = torch.from_numpy((np.round(rng.normal(0, 0.1, N * 64) * 32768) / 32768).astype(np.float32)).cuda().repeat(U // 64 + 1)[: U * N]
fb = batch.FeatureBatch(np.arange(U, dtype=np.int64) * N, np.full(U, N, dtype=np.int64))
out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
for _ in range(3):
    fb.lmfe_cmvnw_dev(big.data_ptr(), big.numel(), out.data_ptr())
fb.ctx.sync()
PY
timeout -s KILL 600 python /tmp/one_call.py > $O/one_call.log 2>&1 || { echo "plain run failed"; tail $O/one_call.log; exit 1; }
timeout -s KILL 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 2 -c 1 -o $O/ncu_$K -f python /tmp/one_call.py > $O/ncu_$K.log 2>&1
echo "ncu rc=$?"; tail -3 $O/ncu_$K.log
