#!/bin/bash
# quick 1-GPU check: smoke under a short timeout, the named tests, the default bench line (kernel-timed only)
mkdir -p gpurun_out/r02
timeout -s KILL 90 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02/smoke_quick.log 2>&1 || { echo "smoke FAILED or hung"; tail -5 gpurun_out/r02/smoke_quick.log; exit 1; }
tail -1 gpurun_out/r02/smoke_quick.log
timeout -s KILL 240 python -m pytest ${1:-tests} -q -m gpu -x --timeout 60 > gpurun_out/r02/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02/pytest_quick.log
timeout -s KILL 60 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-also --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('bench %.0f  %.3f ms' % (d['value'], d['ms_per_step']))"
