#!/bin/bash
# N=2: GPU tests, then the 2-rank bench (strong + weak + sharded cfg5) and the reference arm under torchrun
O=gpurun_out/r02j; mkdir -p $O
timeout -s KILL 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1
echo "rc=$?" >> $O/pytest_gpu.log
timeout -s KILL 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err
echo "rc=$?" >> $O/bench_n2.err
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/bench_ref_n2.json 2> $O/bench_ref_n2.err
echo "rc=$?" >> $O/bench_ref_n2.err
for f in $O/pytest_gpu.log $O/bench_n2.err $O/bench_ref_n2.err; do echo "== $f"; tail -n 6 $f; done
cat $O/bench_ref_n2.json
