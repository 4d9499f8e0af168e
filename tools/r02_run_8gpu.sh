#!/bin/bash
# round-2 record on 8 GPUs (gpurun --gpus 8): bench line with also:{} and halo probe, reference arm under torchrun,
# the reference's own MultiGPU programs at np = 8 through the MPI stand-in
N=${1:-8}
out=gpurun_out/r02/mgpu$N
mkdir -p $out
timeout -s KILL 90 python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1 || { echo "smoke FAILED or hung"; tail -5 $out/smoke.log; exit 1; }
timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus $N --steps 20 --warmup 5 > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"; tail -2 $out/bench.err | cut -c1-200
timeout -s KILL 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29742 bench.py --impl reference --gpus $N --steps 20 --warmup 5 > $out/bench_reference.json 2> $out/bench_reference.err; echo "ref rc=$?"
REFMGPU_ONLY=Burgers3d,Diffusion3d timeout -s KILL 300 python tests/baselines/ref_mgpu_baseline.py $N > $out/ref_mgpu_baseline_np$N.json 2> $out/ref_mgpu_baseline.err; echo "ref mgpu rc=$?"; tail -c 400 $out/ref_mgpu_baseline_np$N.json
nvidia-smi topo -m > $out/topo.txt 2>&1; nproc >> $out/topo.txt
