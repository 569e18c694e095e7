#!/usr/bin/env python3
"""Would a block-scalar preconditioner do?  Replace every 3x3 block of the single-precision copy of the
momentum Jacobian by (mean of its diagonal) * I and compare GMRES iteration counts (development probe).
usage: probe_scalar_pc.py N [dt]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nsb200                      # noqa: E402
from nsb200 import _lib as L       # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dt = float(sys.argv[2]) if len(sys.argv) > 2 else 1e-2
V = nsb200.VelocityBCType
mesh, markers = nsb200.hyper_cube(3, n)
ts = nsb200.BDFTimeStepping(0.0, 1e3, order=2, desired_start_time_step=dt)
solver = nsb200.IPCSSolver(mesh, markers, "standard", ts, preconditioner_precision="f32")
solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / 100.0))
solver.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3, 5, 6)] + [(V.constant, 4, (1.0, 0.0, 0.0))])
dev = solver.dev
for s in range(3):
    ts.update_coefficients()
    solver.solve()
    print("step %d newton=%s krylov=%s" % (s, solver.newton_iterations[-1], solver.krylov_iterations[-1]), flush=True)
    ts.advance_time()
    solver.advance_time()
ts.update_coefficients()
solver._update_time_stepping_coefficients()
r0 = solver._residual()
dev.assemble_ipcs_momentum(L.MAT, solver._intermediate_velocity, A=solver._A_mom)
dev.bc_apply_matrix("P2P2", solver._A_mom, L.BSR, 3, solver._vbc_dofs)
dev.inv_diagonal(solver._jac, solver._dinvV)
lmax = 1.1 * dev.lambda_max(solver._jac, solver._dinvV, iters=8, work=solver._work)
for what in ("full blocks", "scalar blocks"):
    dev.to_f32(solver._A_mom, solver._A_mom32)
    if what == "scalar blocks":
        blk = solver._A_mom32.view(-1, 3, 3)
        s = (blk[:, 0, 0] + blk[:, 1, 1] + blk[:, 2, 2]) / 3.0
        blk.zero_()
        blk[:, 0, 0] = s; blk[:, 1, 1] = s; blk[:, 2, 2] = s
    for deg, ratio in ((4, 30.0), (8, 30.0), (10, 30.0), (12, 30.0)):
        pc = dev.precond(solver._dinvV, L.PC_CHEBYSHEV, deg, lmax / ratio, lmax, lowp=True)
        solver._dxV.zero_()
        torch.cuda.synchronize()
        t = time.time()
        its, res = dev.gmres(solver._jac, pc, solver._bV, solver._dxV, rtol=1e-12, restart=40, work=solver._work)
        torch.cuda.synchronize()
        print("%-13s degree %2d: %3d its %7.1f ms res %.2e" % (what, deg, its, (time.time() - t) * 1e3, res), flush=True)
