#!/bin/bash
# One consolidated multi-GPU evidence run (an N-GPU call is charged N x its box time): tests, chunk-plan sweep,
# dispatch table, both bench arms, the 1 GB lazy checkpoint and the recovery profile. Everything lands in gpurun_out/.
#   usage: scripts/gpu8_suite.sh <ngpus> [tag]
N=${1:-8}
TAG=${2:-g$N}
OUT=gpurun_out
mkdir -p $OUT
echo "== tests"; timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -k "(no_fault and auto) or streamed_allreduce_no_fault or death_inside or local_models or recovery_after_kills or host_buffers or nccl_engine or zero_copy or lazy_checkpoint" > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> $OUT/${TAG}_tests.log; tail -4 $OUT/${TAG}_tests.log | cut -c1-600
echo "== chunk plans"; for p in 64M:512M 512M:512M 32M:128M; do timeout 120 python -m rabit_b200.parallel.launcher -n $N --gpus python benchmarks/stream_diag.py plans=$p mib=256 2>&1 | grep DIAG; done > $OUT/${TAG}_plans.log; cut -c1-500 $OUT/${TAG}_plans.log
echo "== dispatch table"; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 benchmarks/algo_sweep.py --emit-table $OUT/${TAG}_dispatch.txt > $OUT/${TAG}_sweep.log 2>&1; grep TABLE $OUT/${TAG}_sweep.log; cat $OUT/${TAG}_dispatch.txt | cut -c1-700
echo "== bench ours"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --steps 20 --warmup 5 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; grep -i "error\|assert" $OUT/${TAG}_bench.err | head -5
echo "== bench reference"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --impl reference --gpus $N --steps 5 --warmup 2 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "ref rc=$?"; head -c 600 $OUT/${TAG}_bench_ref.json
echo "== lazy 1GB + recovery profile"; RABIT_RECOVER_PROFILE=1 timeout 600 python - > $OUT/${TAG}_recover.log 2>&1 <<PY
import sys
sys.path.insert(0, ".")
from rabit_b200.parallel.launcher import launch
n = $N
res = launch(["build/bin/lazy_big", "268435456", "mock=1,1,0,0", "rabit_debug=1"], n, capture=True, timeout=300, per_rank_gpu=True)
print("lazy_big", res.ok, res.restarts, round(res.wall, 3))
for o in res.outputs.values():
    print("\n".join(l for l in o.splitlines() if "LAZY_LOAD" in l or "moved over NVLink" in l))
res = launch(["build/bin/local_recover", "2000000", "rabit_debug=1", "mock=0,1,1,0", "mock=%d,2,1,0" % (n - 1)], n, capture=True, timeout=300, per_rank_gpu=True)
print("local_recover", res.ok, res.restarts, round(res.wall, 3))
lines = [l for o in res.outputs.values() for l in o.splitlines() if "local state complete" in l]
print("\n".join(lines[-n:]))
for sb in (True, False):
    base = launch(["build/bin/device_recover_cuda", "200000", "rabit_bootstrap_cache=1"], n, capture=True, timeout=300, per_rank_gpu=True, standby=sb)
    res = launch(["build/bin/device_recover_cuda", "200000", "rabit_bootstrap_cache=1", "mock=0,0,1,0", "mock=1,1,1,0"], n, capture=True, timeout=300, per_rank_gpu=True, standby=sb)
    print("standby", sb, res.ok, res.restarts, "wall", round(res.wall, 3), "base", round(base.wall, 3), "per kill", round((res.wall - base.wall) / 2, 3))
    for i, o in sorted(res.outputs.items()):
        print("\n".join(l for l in o.splitlines() if "re-join" in l or "attach detail" in l or "imported" in l))
PY
tail -60 $OUT/${TAG}_recover.log | cut -c1-400
