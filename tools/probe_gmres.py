#!/usr/bin/env python3
"""Sweep of the momentum-system preconditioner (Jacobi vs Chebyshev degree / interval ratio) on
the 3D cavity Jacobian after a few IPCS steps (development probe).  usage: probe_gmres.py N"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nsb200                      # noqa: E402
from nsb200 import _lib as L       # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
V = nsb200.VelocityBCType
mesh, markers = nsb200.hyper_cube(3, n)
ts = nsb200.BDFTimeStepping(0.0, 1e3, order=2, desired_start_time_step=1e-2)
solver = nsb200.IPCSSolver(mesh, markers, "standard", ts)
solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / 100.0))
solver.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3, 5, 6)] + [(V.constant, 4, (1.0, 0.0, 0.0))])
dev = solver.dev
for s in range(3):
    torch.cuda.synchronize()
    t = time.time()
    ts.update_coefficients()
    solver.solve()
    torch.cuda.synchronize()
    print("step %d: %.1f ms newton=%s krylov=%s" % (s, (time.time() - t) * 1e3, solver.newton_iterations[-1],
                                                    solver.krylov_iterations[-1]), flush=True)
    ts.advance_time()
    solver.advance_time()
# momentum system of the next step's first Newton iteration
ts.update_coefficients()
solver._update_time_stepping_coefficients()
r0 = solver._residual()
dev.assemble_ipcs_momentum(L.MAT, solver._intermediate_velocity, A=solver._A_mom)
dev.bc_apply_matrix("P2P2", solver._A_mom, L.BSR, 3, solver._vbc_dofs)
dev.inv_diagonal(solver._jac, solver._dinvV)
t = time.time()
lmax = dev.lambda_max(solver._jac, solver._dinvV, iters=8, work=solver._work)
torch.cuda.synchronize()
print("lambda_max(8 its) = %.4f (%.1f ms); 30 its: %.4f" % (lmax, (time.time() - t) * 1e3,
                                                          dev.lambda_max(solver._jac, solver._dinvV, iters=30)))
ref = None
lowp = solver._pc_lowp
if solver._A_mom32 is not None:
    dev.to_f32(solver._A_mom, solver._A_mom32)
if solver._S_mom32 is not None:
    dev.block_scalar_f32(solver._A_mom, 3, solver._S_mom32)
print("single-precision polynomial / flexible GMRES:", lowp)
for kind, deg, ratio in [("jacobi", 1, 0)] + [("cheb", d, r) for d in (3, 4, 5, 6, 8, 10) for r in (15.0, 30.0, 60.0)]:
    pc = dev.precond(solver._dinvV) if kind == "jacobi" else \
        dev.precond(solver._dinvV, L.PC_CHEBYSHEV, deg, 1.1 * lmax / ratio, 1.1 * lmax, lowp=lowp)
    solver._dxV.zero_()
    torch.cuda.synchronize()
    t = time.time()
    try:
        its, res = dev.gmres(solver._jac, pc, solver._bV, solver._dxV, rtol=1e-12, restart=40, work=solver._work)
    except RuntimeError as e:
        print(kind, deg, ratio, "FAILED", e)
        continue
    torch.cuda.synchronize()
    el = (time.time() - t) * 1e3
    x = solver._dxV.clone()
    if ref is None:
        ref = x
    err = float((x - ref).abs().max() / ref.abs().max())
    print("%-7s degree %d ratio %5.0f: %3d its (%4d spmv) %8.1f ms  res %.2e  diff vs jacobi %.1e"
          % (kind, deg, ratio, its, its * deg, el, res, err), flush=True)
