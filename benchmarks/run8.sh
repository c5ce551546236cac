#!/bin/bash
# One consolidated multi-GPU session (an N-GPU call is charged N x): tests, tuning sweeps, bench both arms.
# usage: benchmarks/run8.sh <ngpu>
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo$N.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
(timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi$N.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi$N.log)
timeout 600 $TR --master-port 29731 benchmarks/algo_sweep.py --sizes 1048576,16777216,268435456,1073741824 --algos twoshot,nvls --out gpurun_out/sweep$N.json > gpurun_out/sweep$N.log 2>&1
timeout 600 $TR --master-port 29732 benchmarks/algo_sweep.py --sym --sizes 1048576,16777216,268435456,1073741824 --algos twoshot,nvls --out gpurun_out/sweep${N}_sym.json > gpurun_out/sweep${N}_sym.log 2>&1
(timeout 600 python benchmarks/recovery_time.py --gpus $N --repeat 2 --out gpurun_out/recovery_gpu$N.json > gpurun_out/recovery_gpu$N.log 2>&1; echo "rc=$?" >> gpurun_out/recovery_gpu$N.log)
(timeout 600 $TR --master-port 29733 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench$N.log)
(timeout 900 $TR --master-port 29734 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/bench${N}_ref.log 2>&1; echo "rc=$?" >> gpurun_out/bench${N}_ref.log)
tail -3 gpurun_out/pytest_multi$N.log
grep "^{" gpurun_out/sweep$N.log | cut -c1-200
grep "^{" gpurun_out/sweep${N}_sym.log | cut -c1-200
tail -6 gpurun_out/recovery_gpu$N.log | cut -c1-200
tail -2 gpurun_out/bench$N.log | cut -c1-600
tail -2 gpurun_out/bench${N}_ref.log | cut -c1-400
