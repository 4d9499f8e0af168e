#!/usr/bin/env python
"""Cost of the dynamic CFL time step against the fixed one, 512^3 double, one GPU:
the same number of steps through adv_step (fixed dt) and adv_run_until_cfl (max|u| folded into stage 3,
dt computed on the device).  Prints one JSON line; run on a GPU box."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import multigpu_advectiondiffusion_b200 as adv  # noqa: E402

n, steps = 512, 20
with adv.Solver("burgers3d_mgpu", n, n, n, coef=0.3) as s:
    s.init_reference_ic_device()
    s.step(3); s.sync()
    s.step(steps); s.sync()
    fixed_ms = s.elapsed_ms() / steps
    info = s.run_until_cfl(1e9, max_steps=3); s.sync()
    info = s.run_until_cfl(1e9, max_steps=steps); s.sync()
    cfl_ms = s.elapsed_ms() / info["it"]
print(json.dumps({"grid": [n, n, n], "steps": steps, "fixed_dt_ms_per_step": fixed_ms, "dynamic_cfl_ms_per_step": cfl_ms,
                  "overhead": cfl_ms / fixed_ms - 1.0, "dynamic_steps_taken": info["it"],
                  "note": "dynamic run: 2 extra no-op steps are enqueued past the end and are inside its time"}))
