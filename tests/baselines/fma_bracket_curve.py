"""Rounding bracket of the reference's recorded runs (hard part 3).

The CPU restatement is built twice (oracle/Makefile): contraction off (the parity target) and
-mfma -ffp-contract=fast (what a compiler that fuses a*b+c does to the SAME source).  This script records
max|u_fma - u_off| / max|u_off| against the step count on the reference's recorded multi-GPU Burgers runs
(MultiGPU/Burgers2d_Baseline/Run.m:4-14: 400 x 406, tEnd 0.40, CFL 0.40, 200 steps;
 MultiGPU/Burgers3d_Baseline/Run.m:4-14 on a thin 400 x 20 x 26 grid: tEnd 0.40, CFL 0.30, 267 steps).
Writes profiles/r02_fma_bracket.json.  CPU only; about a minute.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as orc


def curve(run, steps_list):
    rows = []
    for n in steps_list:
        a, ia = run(n)
        with orc.fma_variant():
            b, ib = run(n)
        assert ia["it"] == ib["it"] == n
        rows.append({"steps": n, "t": ia["t"], "rel_diff_fma_vs_off": orc.rel_err(b, a)})
        print(rows[-1], flush=True)
    return rows


def main():
    u2 = orc.ic_burgers2d_mgpu(400, 400)
    r2 = curve(lambda n: orc.burgers2d_mgpu(u2, 0.40, 0.40, max_steps=n), [25, 50, 75, 100, 125, 150, 160, 170, 175, 180, 185, 190, 200])
    u3 = orc.ic_burgers3d_mgpu(400, 20, 20, 2.0, 2.0, 4.0)
    r3 = curve(lambda n: orc.burgers3d_mgpu(u3, 0.40, 0.30, 2.0, 2.0, 4.0, max_steps=n), [50, 100, 133, 150, 200, 230, 250, 267])
    out = {"what": "max|u_fma - u_off|/max|u_off| of oracle/advdiff_oracle.c built with and without FP contraction",
           "burgers2d_mgpu_400x400_cfl0.40": r2, "burgers3d_mgpu_400x20x20_cfl0.30_H4": r3}
    with open(os.path.join(ROOT, "profiles", "r02_fma_bracket.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
