#!/bin/bash
# A/B of the Burgers stage-kernel variants on one B200 (device-resident timing only).
mkdir -p gpurun_out
for v in direct 0 1 2; do
  if [ "$v" = direct ]; then export ADV_BURGERS_DIRECT=1; unset ADV_BURGERS_VARIANT; else unset ADV_BURGERS_DIRECT; export ADV_BURGERS_VARIANT=$v; fi
  for w in burgers3d_weak512 burgers3d_256; do
    echo "variant=$v workload=$w $(python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-e2e | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"]), "Mcell-upd/s", round(d["ms_per_step"],3), "ms/step", d["clocks"])')"
  done
done
