#!/bin/bash
# A/B of the heat 3-D chunk rule (cost model vs round-1 rule) and of CUDA-graph replay at 256^3, one GPU
run() { wl=$1; shift; env "$@" timeout -s KILL 60 python bench.py --workload $wl --steps 60 --warmup 6 --no-cpu-baseline --no-also --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('%-18s %-24s %.0f  %.4f ms  frac %.3f' % (sys.argv[1], ' '.join(sys.argv[2:]), d['value'], d['ms_per_step'], d['roofline']['frac']))" $wl "$@"; }
timeout -s KILL 90 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout -s KILL 200 python -m pytest tests/test_gpu_parity.py -q -m gpu -x --timeout 60 2>&1 | tail -2
for wl in heat3d_256 heat3d_sgpu_256 heat3d_weak512 burgers3d_256; do
  run $wl A=default
  run $wl ADV_NO_GRAPH=1
  run $wl ADV_HEAT_CHUNK_OLD=1 ADV_NO_GRAPH=1
done
