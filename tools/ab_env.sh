#!/bin/bash
# generic A/B: each argument is an environment assignment list, e.g. "ADV_BURGERS_BARSYNC=1 ADV_BURGERS_RING=32"
W=${WORKLOAD:-burgers3d_weak512}
for e in "$@"; do
  echo "[$e] smoke: $(env $e timeout -s KILL 60 python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -1)"
  echo "[$e] $W $(env $e timeout -s KILL 90 python bench.py --workload $W --steps 10 --warmup 3 --no-cpu-baseline --no-e2e | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"]), "Mcell-upd/s", round(d["ms_per_step"],3), "ms/step")')"
done
