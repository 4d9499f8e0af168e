#!/bin/bash
run() { wl=$1; shift; env "$@" timeout -s KILL 60 python bench.py --workload $wl --steps 50 --warmup 5 --no-cpu-baseline --no-also --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('%-18s %-24s %.0f  %.4f ms  frac %.3f' % (sys.argv[1], ' '.join(sys.argv[2:]), d['value'], d['ms_per_step'], d['roofline']['frac']))" $wl "$@"; }
for n in 10 12 14 16 18 20 24 28 32 37 43 52; do run heat3d_weak512 ADV_HEAT_NCH=$n; done
for n in 9 10 11 13 14 16 20 24 32; do run heat3d_256 ADV_HEAT_NCH=$n; done
