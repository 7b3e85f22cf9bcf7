#!/bin/bash
# round 2, first GPU pass: new tensor-core kernel + exactness guard
O=gpurun_out/r02a; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout -s KILL 300 python -c "
import numpy as np, sys, torch
sys.path.insert(0,'.')
from speaker_verification_b200 import _lib, batch, synth
from oracle import speechpy_oracle as o
x = synth.audio(1, 48000, np.float32)
arena, starts, lens = batch.pack([x, synth.audio(2, 100000, np.float32)])
fb = batch.FeatureBatch(starts, lens)
ref = np.concatenate([o.lmfe(arena[s:s+n], 16000) for s, n in zip(starts, lens)])
for opts in (dict(tc_version=1), dict(feature_kernel=2), dict(tc_epilogue=0), dict(tc_epilogue=1), dict(), dict(redo=0)):
    d_in = torch.from_numpy(arena).cuda(); d_out = torch.zeros((fb.total_frames, 40), device='cuda')
    with _lib.options(**opts):
        fb.lmfe_dev(d_in.data_ptr(), d_in.numel(), d_out.data_ptr()); fb.ctx.sync()
    got = d_out.cpu().numpy()
    print(opts, 'max err', float(np.abs(got-ref).max()), flush=True)
" > $O/first.log 2>&1
echo "first rc=$?" >> $O/first.log
timeout -s KILL 900 python -m pytest tests/test_gpu_tc.py -x -q > $O/pytest_tc.log 2>&1
echo "rc=$?" >> $O/pytest_tc.log
timeout -s KILL 600 python tools/guard_calibrate.py $O/calib.npz > $O/calib.log 2>&1
timeout -s KILL 300 python tools/tc_time.py lmfe > $O/time_lmfe.log 2>&1
timeout -s KILL 300 python tools/tc_time.py lmfe_cmvnw > $O/time_lmfe_cmvnw.log 2>&1
timeout -s KILL 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1
echo "rc=$?" >> $O/pytest_gpu.log
tail -3 $O/first.log $O/pytest_tc.log $O/time_lmfe.log $O/pytest_gpu.log
