#!/bin/bash
# A/B of the Burgers stage-kernel variants on one B200 (device-resident timing only).
mkdir -p gpurun_out
run() { for w in burgers3d_weak512 burgers3d_256; do
    echo "$1 workload=$w $(python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-e2e | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(round(d["value"]), "Mcell-upd/s", round(d["ms_per_step"],3), "ms/step", d["clocks"]["sm_mhz"])')"; done; }
for v in ${VARIANTS:-ring direct}; do
  unset ADV_BURGERS_DIRECT
  case $v in
    direct) export ADV_BURGERS_DIRECT=1;;
    ring) ;;
  esac
  python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
  run "variant=$v"
done
