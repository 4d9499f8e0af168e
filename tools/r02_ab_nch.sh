#!/bin/bash
# A/B of the pair kernel's z-chunk count at 512^3 (one GPU)
run() { env "$@" timeout -s KILL 60 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-also --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('%-28s %.0f  %.3f ms' % (' '.join(sys.argv[1:]), d['value'], d['ms_per_step']))" "$@"; }
for n in 4 5 7 10 12 14 16 20 26 37; do run ADV_BURGERS_NCH=$n; done
run ADV_BURGERS_NCH=4
