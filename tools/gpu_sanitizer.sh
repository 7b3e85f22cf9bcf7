#!/bin/bash
# compute-sanitizer (memcheck / racecheck / initcheck / synccheck) over a small-size subset of the parity suite
O=gpurun_out/r02_san; mkdir -p $O
SEL='tests/test_gpu_tc.py::test_tc_matches_fft_kernel_and_oracle_on_group_boundaries tests/test_gpu_tc.py::test_tc_falls_back_for_samples_it_cannot_split tests/test_gpu_tc.py::test_tc_silence_and_gain tests/test_gpu_batch.py::test_lmfe_cmvnw_ragged_batch tests/test_gpu_batch.py::test_single_pass_cmvnw_matches_two_pass_kernels_and_oracle tests/test_gpu_batch.py::test_mfcc_delta_batch tests/test_gpu_scoring.py::test_cosine_shapes tests/test_zz_no_alloc.py'
for tool in memcheck racecheck initcheck synccheck; do
  timeout -s KILL 1500 compute-sanitizer --tool $tool --error-exitcode 77 --print-limit 20 python -m pytest $SEL -q -x -p no:cacheprovider > $O/$tool.log 2>&1
  echo "exit=$?" >> $O/$tool.log
  echo "== $tool"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|exit=" $O/$tool.log | tail -4
done
