#!/bin/bash
# round-2 GPU check: full GPU suite (stop at first failure), default bench line (with also:{}), reference arm, heat2d TMA bring-up
mkdir -p gpurun_out/r02
timeout -s KILL 600 python -m pytest tests -q -m gpu -x > gpurun_out/r02/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02/pytest_gpu.log
timeout -s KILL 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r02/bench_default.json 2> gpurun_out/r02/bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/r02/bench_default.err
timeout -s KILL 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02/bench_reference.json 2> gpurun_out/r02/bench_reference.err; echo "ref rc=$?"; tail -3 gpurun_out/r02/bench_reference.err
timeout -s KILL 150 python tests/baselines/heat2d_tma_check.py > gpurun_out/r02/heat2d_tma.log 2>&1; echo "heat2d tma rc=$?"; tail -6 gpurun_out/r02/heat2d_tma.log
