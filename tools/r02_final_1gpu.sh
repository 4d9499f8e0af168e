#!/bin/bash
# round-2 final record on one B200: whole GPU suite, default bench line (with also / e2e / cpu_baseline), reference arm,
# the other workloads' bench lines, then the ncu launch list and one --set full capture of three consecutive stage launches
out=gpurun_out/r02/final
mkdir -p $out
timeout -s KILL 90 python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1 || { echo "smoke FAILED or hung"; tail -5 $out/smoke.log; exit 1; }
tail -1 $out/smoke.log
timeout -s KILL 500 python -m pytest tests -q -m gpu -x --timeout 120 -rs > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 $out/pytest_gpu.log
timeout -s KILL 300 python bench.py --steps 20 --warmup 5 > $out/bench_default.json 2> $out/bench_default.err; echo "bench rc=$?"
timeout -s KILL 300 python bench.py --impl reference --steps 20 --warmup 5 > $out/bench_reference.json 2> $out/bench_reference.err; echo "ref rc=$?"
for w in burgers3d_256 burgers3d_sgpu_256 heat3d_256 heat3d_sgpu_256 heat2d_512 burgers2d_weak8192 heat2d_weak8192; do
  timeout -s KILL 150 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $out/bench_$w.json 2> $out/bench_$w.err; echo "bench $w rc=$?"
done
timeout -s KILL 100 python bench.py --workload burgers3d_256 --precision f32 --steps 20 --no-cpu-baseline > $out/bench_burgers3d_256_f32.json 2>/dev/null
timeout -s KILL 100 python bench.py --workload heat3d_256 --precision f32 --steps 20 --no-cpu-baseline > $out/bench_heat3d_256_f32.json 2>/dev/null
timeout -s KILL 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $out/launches_burgers512.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-also --no-e2e > /dev/null 2>&1; echo "launch list rc=$?"
timeout -s KILL 300 ncu --set full --clock-control none --import-source on --kernel-name regex:burgers3d_stage_pair --launch-skip 9 --launch-count 3 -o $out/pair_512 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-also --no-e2e > $out/ncu_pair.log 2>&1; echo "ncu rc=$?"
# the reports are ~45 MB each and gpurun brings back at most 64 MiB: summarise on the box, keep the text
python tools/ncu_summary.py $out/pair_512.ncu-rep 1 > $out/burgers_pair_512_ncu.txt 2>&1
python tools/ncu_traffic.py burgers3d_weak512 burgers3d_stage_pair $out/pair_512.ncu-rep 'burgers3d_stage_pairIdLi0ELi1ELi[123]ELb0ELb0E' > $out/traffic_burgers.log 2>&1
rm -f $out/pair_512.ncu-rep
timeout -s KILL 200 ncu --set full --clock-control none --import-source on --kernel-name regex:heat3d_stage_tma --launch-skip 9 --launch-count 3 -o $out/heat_512 -f python bench.py --workload heat3d_weak512 --steps 2 --warmup 3 --no-cpu-baseline --no-also --no-e2e > $out/ncu_heat.log 2>&1; echo "ncu heat rc=$?"
python tools/ncu_summary.py $out/heat_512.ncu-rep 1 > $out/heat_tma_512_ncu.txt 2>&1
python tools/ncu_traffic.py heat3d_weak512 heat3d_stage_tma $out/heat_512.ncu-rep 'heat3d_stage_tmaIdLi8ELi6ELi[123]ELb0E' > $out/traffic_heat.log 2>&1
rm -f $out/heat_512.ncu-rep
cp profiles/traffic.json $out/traffic.json
