#!/bin/bash
# A/B of the pair kernel's launch shape, then one ncu --set full capture of three consecutive stage launches
mkdir -p gpurun_out/r02
run() { env "$@" timeout -s KILL 60 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-also --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('%-50s %.0f  %.3f ms' % (' '.join(sys.argv[1:]), d['value'], d['ms_per_step']))" "$@"; }
run A=0
run ADV_BURGERS_WAVES=4
run ADV_BURGERS_WAVES=8
run ADV_BURGERS_WAVES=12
run ADV_BURGERS_WAVES=16 ADV_BURGERS_MINCHUNK=8
run ADV_BURGERS_WAVES=3
run A=0
timeout -s KILL 300 ncu --set full --clock-control none --import-source on --kernel-name regex:burgers3d_stage_pair --launch-skip 9 --launch-count 3 -o gpurun_out/r02/pair_512 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-also --no-e2e > gpurun_out/r02/ncu_pair.log 2>&1; echo "ncu rc=$?"
timeout -s KILL 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02/launches_burgers512.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-also --no-e2e > /dev/null 2>&1; echo "launch list rc=$?"
