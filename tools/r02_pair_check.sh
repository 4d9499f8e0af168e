#!/bin/bash
# bring-up of a changed Burgers 3-D kernel: parity tests that exercise it, then the default bench line (no CPU baseline, no extra records)
mkdir -p gpurun_out/r02
# a hung kernel must cost seconds, not the call's limit: smoke first, under a short timeout, and stop if it does not pass
timeout -s KILL 90 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02/smoke_pair.log 2>&1 || { echo "smoke FAILED or hung"; tail -5 gpurun_out/r02/smoke_pair.log; exit 1; }
tail -1 gpurun_out/r02/smoke_pair.log
timeout -s KILL 240 python -m pytest tests/test_gpu_parity.py tests/test_recorded_runs.py tests/test_zz_no_alloc.py tests/test_drivers.py -q -m gpu -x --timeout 45 > gpurun_out/r02/pytest_pair.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02/pytest_pair.log
for v in "" "ADV_BURGERS_RING=1"; do
  env $v timeout -s KILL 60 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-also --no-e2e > gpurun_out/r02/bench_pair_$v.json 2> gpurun_out/r02/bench_pair_$v.err; echo "bench [$v] rc=$?"
  python - "gpurun_out/r02/bench_pair_$v.json" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print("  %.0f Mcell-updates/s  %.3f ms/step  clocks %s" % (d["value"], d["ms_per_step"], d["clocks"]["sm_mhz"]))
except Exception as e:
    print("  no bench line:", e)
PY
done
timeout -s KILL 60 python bench.py --workload burgers3d_sgpu_256 --steps 20 --no-cpu-baseline --no-e2e 2>&1 | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('sgpu256', round(d['value']))"
timeout -s KILL 60 python bench.py --workload burgers3d_256 --steps 20 --no-cpu-baseline --no-e2e 2>&1 | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('mgpu256', round(d['value']))"
timeout -s KILL 60 python bench.py --workload burgers3d_256 --precision f32 --steps 20 --no-cpu-baseline --no-e2e 2>&1 | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('mgpu256 f32', round(d['value']))"
