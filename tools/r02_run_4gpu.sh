#!/bin/bash
# round-2 check on N GPUs (gpurun --gpus N): multi-rank parity (incl. dynamic CFL), bench line with also:{}, the reference's
# own MultiGPU programs at np = N through the MPI stand-in, dynamic-CFL overhead on one GPU
N=${1:-4}
out=gpurun_out/r02/mgpu$N
mkdir -p $out
timeout -s KILL 90 python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1 || { echo "smoke FAILED or hung"; tail -5 $out/smoke.log; exit 1; }
timeout -s KILL 400 python -m pytest tests/test_multirank.py -q -m gpu -rs -s --timeout 180 > $out/pytest_mgpu.log 2>&1; echo "pytest rc=$?"; grep -E "MP_OK|passed|failed|SKIP|Error|error" $out/pytest_mgpu.log | tail -30
timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus $N --steps 20 --warmup 5 > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"; tail -3 $out/bench.err
REFMGPU_ONLY=Burgers3d,Diffusion3d timeout -s KILL 400 python tests/baselines/ref_mgpu_baseline.py $N > $out/ref_mgpu_baseline_np$N.json 2> $out/ref_mgpu_baseline.err; echo "ref mgpu rc=$?"; tail -c 600 $out/ref_mgpu_baseline_np$N.json
timeout -s KILL 120 python tests/baselines/cfl_overhead.py > $out/cfl_overhead.json 2>&1; cat $out/cfl_overhead.json
