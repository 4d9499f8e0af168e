import numpy as np, sys
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__)))))
import multigpu_advectiondiffusion_b200 as adv
from oracle import oracle as orc
h0 = orc.ic_heat3d_sgpu(32, 32, 32, dtype=np.float32)
want, _ = orc.heat3d(h0, 10, "baseline")
got, _ = adv.solve_host("heat3d_sgpu_baseline", h0, 10, lx=10, ly=10, lz=10, coef=1.0, precision=adv.F32)
print("rel", orc.rel_err(got, want))
