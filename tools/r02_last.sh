#!/bin/bash
# the driver's two round-end commands on one B200, final build
out=gpurun_out/r02/last
mkdir -p $out
timeout -s KILL 300 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $out/ref.json 2> $out/ref.err; echo "ref rc=$?"
timeout -s KILL 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
def load(f): return json.loads([l for l in open(f) if l.startswith("{")][-1])
d, r = load("gpurun_out/r02/last/bench.json"), load("gpurun_out/r02/last/ref.json")
h = d["also"]["heat3d_weak512"]
print("burgers %.0f (%.3f ms) hbm frac %.3f traffic %s fp64 frac %.3f e2e %.0f launches %d clocks %s" % (
    d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["traffic"], d["roofline_fp64"]["frac"], d["e2e"]["value"], d["gpu_launches"], d["clocks"]["sm_mhz"]))
print("heat %.0f (%.4f ms) frac %.3f traffic %s" % (h["value"], h["ms_per_step"], h["roofline"]["frac"], h["roofline"]["traffic"]))
print("reference arm %.1f (%.1f ms/step), same metric %s, same config %s, ratio value %.0f e2e %.0f" % (
    r["value"], r["ms_per_step"], r["metric"] == d["metric"], r["config"] == d["config"], d["value"] / r["value"], d["e2e"]["value"] / r["value"]))
PY
