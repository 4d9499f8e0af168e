#!/bin/bash
# A/B on N GPUs: one launch per stage (default) against the multi-launch schedule (ADV_HALO_MULTILAUNCH=1)
N=${1:-2}
run() { wl=$1; shift; env "$@" timeout -s KILL 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29751 bench.py --gpus $N --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --no-also --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('%-22s %-28s %.0f  %.4f ms  launches %d' % (sys.argv[1], ' '.join(sys.argv[2:]), d['value'], d['ms_per_step'], d['gpu_launches']))" $wl "$@"; }
for wl in heat3d_weak512 heat3d_strong1024 burgers3d_weak512; do
  run $wl A=fused
  run $wl ADV_HALO_MULTILAUNCH=1
  run $wl A=fused
done
