#!/usr/bin/env python3
"""Launch one kernel class a few times on the 3D cavity (for ncu captures).
usage: run_kernel.py N {jacobian|residual|spmv|mass|poisson_rhs|correction_rhs} [reps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nsb200                      # noqa: E402
from nsb200 import _lib as L       # noqa: E402

n, what = int(sys.argv[1]), sys.argv[2]
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
V = nsb200.VelocityBCType
mesh, markers = nsb200.hyper_cube(3, n)
ts = nsb200.BDFTimeStepping(0.0, 1e3, order=2, desired_start_time_step=1e-2)
s = nsb200.IPCSSolver(mesh, markers, "standard", ts)
s.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=0.01))
s.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3, 5, 6)] + [(V.constant, 4, (1.0, 0.0, 0.0))])
s._setup_boundary_conditions()
s._update_time_stepping_coefficients()
s._assemble_constant()
dev = s.dev
u = s._intermediate_velocity
u.uniform_(-1.0, 1.0)
s._pressure.uniform_(-1.0, 1.0)
x, y = dev.zeros(u.numel()), dev.zeros(u.numel())
x.uniform_(-1, 1)
for _ in range(reps):
    if what == "jacobian":
        dev.assemble_ipcs_momentum(L.MAT, u, A=s._A_mom)
    elif what == "residual":
        dev.assemble_ipcs_momentum(L.VEC, u, s._velocities[1], s._velocities[2], s._pressure, b=s._bV)
    elif what == "spmv":
        dev.spmv(s._jac, x, y)
    elif what == "mass":
        dev.assemble_correction(L.MAT, A=s._A_mass)
    elif what == "poisson_rhs":
        dev.assemble_poisson(L.VEC, pn=s._pressure, ustar=u, b=s._bQ)
    elif what == "correction_rhs":
        dev.assemble_correction(L.VEC, ustar=u, pnew=s._pressure, pold=s._old_pressure, b=s._bV)
torch.cuda.synchronize()
print("ok", what, n)
