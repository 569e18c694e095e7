#!/usr/bin/env python3
"""Development probe: Chebyshev- vs Jacobi-GMRES on the 2D channel momentum systems, Newton iteration
by Newton iteration (prints lambda_max, iterations and residuals).  usage: probe_cheb2d.py [N]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nsb200                      # noqa: E402
from nsb200 import _lib as L       # noqa: E402
from tools.run_partitioned import make_problem   # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 12
mesh, markers, bcs = make_problem(nsb200, 2, n, "channel")
ts = nsb200.BDFTimeStepping(0.0, 1e3, order=2, desired_start_time_step=1e-2)
solver = nsb200.IPCSSolver(mesh, markers, "standard", ts, momentum_preconditioner="jacobi")
solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=0.02))
solver.set_boundary_conditions(bcs)
dev, d = solver.dev, 2
for step in range(2):
    ts.update_coefficients()
    solver._setup_boundary_conditions() if not solver._bc_ready else None
    solver._set_time(ts.next_time)
    solver._update_bc_values()
    solver._update_time_stepping_coefficients()
    if not solver._have_constant:
        solver._assemble_constant()
    ustar = solver._intermediate_velocity
    r0 = r = solver._residual()
    for it in range(6):
        dev.assemble_ipcs_momentum(L.MAT, ustar, A=solver._A_mom)
        dev.bc_apply_matrix("P2P2", solver._A_mom, L.BSR, d, solver._vbc_dofs)
        dev.inv_diagonal(solver._jac, solver._dinvV)
        lm8 = dev.lambda_max(solver._jac, solver._dinvV, iters=8, work=solver._work)
        lm40 = dev.lambda_max(solver._jac, solver._dinvV, iters=40, work=solver._work)
        dinv = solver._dinvV.cpu().numpy()
        print("step %d newton %d: r=%.3e r0=%.3e lmax8=%.4f lmax40=%.4f dinv range [%.3e, %.3e] bnorm=%.3e"
              % (step, it, r, r0, lm8, lm40, dinv.min(), dinv.max(), dev.norm(solver._bV)), flush=True)
        sols = {}
        for kind in ("jacobi", "cheb4", "cheb2"):
            pc = dev.precond(solver._dinvV) if kind == "jacobi" else \
                dev.precond(solver._dinvV, L.PC_CHEBYSHEV, int(kind[4:]), 1.1 * lm8 / 30.0, 1.1 * lm8)
            x = dev.zeros(solver._nV_own)
            for maxit in (5, 40, 400):
                x.zero_()
                try:
                    its, res = dev.gmres(solver._jac, pc, solver._bV, x, rtol=1e-12, atol=1e-12 * r0, maxit=maxit,
                                         restart=40, work=solver._work)
                    print("   %-6s maxit %3d: converged its=%d res=%.3e" % (kind, maxit, its, res), flush=True)
                    break
                except RuntimeError as e:
                    y = dev.zeros(solver._nV_own)
                    dev.spmv(solver._jac, x, y)
                    tr = float((solver._bV - y).norm())
                    print("   %-6s maxit %3d: FAILED (%s) true residual %.3e" % (kind, maxit, str(e)[-40:], tr), flush=True)
            sols[kind] = x.clone()
        dev.axpy(-1.0, sols["jacobi"], ustar)
        r = solver._residual()
        if r < max(1e-10, 1e-9 * r0):
            break
    # finish the step with the regular path
    solver._intermediate_velocity.copy_(solver._velocities[1])
    solver.solve()
    ts.advance_time()
    solver.advance_time()
