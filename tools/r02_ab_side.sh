#!/bin/bash
# A/B on N GPUs: heat boundary launches on a side stream (ADV_HALO_SIDE_STREAM=1) against the one-stream schedule
N=${1:-4}
run() { wl=$1; shift; env "$@" timeout -s KILL 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29751 bench.py --gpus $N --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --no-also --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('%-22s %-28s %.0f  %.4f ms  launches %d' % (sys.argv[1], ' '.join(sys.argv[2:]), d['value'], d['ms_per_step'], d['gpu_launches']))" $wl "$@"; }
ADV_HALO_SIDE_STREAM=1 timeout -s KILL 200 python -m pytest tests/test_multirank.py -q -m gpu --timeout 150 -k "four_ranks" 2>&1 | tail -2
for rep in 1 2; do
  run heat3d_weak512 A=one-stream
  run heat3d_weak512 ADV_HALO_SIDE_STREAM=1
done
run heat3d_strong1024 A=one-stream
run heat3d_strong1024 ADV_HALO_SIDE_STREAM=1
