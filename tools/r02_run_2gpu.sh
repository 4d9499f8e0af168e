#!/bin/bash
# round-2 multi-GPU check (run under gpurun --gpus N): multi-rank parity tests, no-alloc tests on device 1, N-GPU bench line, reference arm under torchrun
N=${1:-2}
out=gpurun_out/r02/mgpu$N
mkdir -p $out
timeout -s KILL 400 python -m pytest tests/test_multirank.py tests/test_drivers.py tests/test_zz_no_alloc.py -q -m gpu -rs -s --timeout 150 > $out/pytest_mgpu.log 2>&1; echo "pytest rc=$?"; grep -E "MP_OK|passed|failed|SKIP" $out/pytest_mgpu.log | tail -30
timeout -s KILL 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus $N --steps 20 --warmup 5 > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"; tail -3 $out/bench.err
timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29742 bench.py --impl reference --gpus $N --steps 20 --warmup 5 > $out/bench_reference.json 2> $out/bench_reference.err; echo "ref rc=$?"
