#!/bin/bash
# A/B of the heat3d stage-kernel variants on one B200 (device-resident timing only).
for v in ${VARIANTS:-tma direct}; do
  unset ADV_HEAT_DIRECT; if [ "$v" = direct ]; then export ADV_HEAT_DIRECT=1; fi
  timeout -s KILL 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
  for w in ${WORKLOADS:-heat3d_256 heat3d_weak512}; do
    echo "heat_variant=$v workload=$w $(timeout -s KILL 60 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-e2e | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"]), "Mcell-upd/s", round(d["ms_per_step"],3), "ms/step", round(d["roofline"]["achieved"]), "GB/s", round(d["roofline"]["frac"],3))')"
  done
done
