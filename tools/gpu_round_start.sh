#!/bin/bash
# One gpurun call that re-establishes the state of the build on a fresh B200 box (about two minutes):
#   /usr/local/graft/bin/gpurun --timeout 600 -- 'bash tools/gpu_round_start.sh'
# Everything lands in gpurun_out/round_start/.  No profiler here (ncu goes in its own call, after this one is green).
out=gpurun_out/round_start
mkdir -p $out
timeout -s KILL 300 python -m pytest tests -x -q -m gpu > $out/pytest_gpu.log 2>&1; echo "pytest -m gpu: rc=$? $(tail -1 $out/pytest_gpu.log)"
timeout -s KILL 60 python -c 'import __graft_entry__ as g; g.smoke()' > $out/smoke.log 2>&1; echo "smoke: rc=$? $(tail -1 $out/smoke.log)"
timeout -s KILL 150 python tests/baselines/heat2d_tma_check.py > $out/heat2d_tma.log 2>&1; echo "heat2d TMA (experimental): rc=$? $(tail -3 $out/heat2d_tma.log | tr '\n' ' ')"
for w in burgers3d_weak512 heat3d_weak512 burgers2d_weak8192 heat2d_weak8192 burgers3d_256 heat3d_256; do
  timeout -s KILL 150 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > $out/bench_$w.json 2> $out/bench_$w.err
  python - "$out/bench_$w.json" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print("%-22s %8.0f Mcell-updates/s  %.3f ms/step  hbm frac %.3f  e2e %.0f  clocks %s" % (
        d["config"]["workload"], d["value"], d["ms_per_step"], d["roofline"]["frac"], (d.get("e2e") or {}).get("value", 0), d["clocks"]["sm_mhz"]))
except Exception as e:
    print(sys.argv[1], "no bench line:", e)
PY
done
