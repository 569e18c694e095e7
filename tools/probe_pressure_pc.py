#!/usr/bin/env python3
"""Development probe: Jacobi- vs Chebyshev-polynomial-preconditioned CG on the C5 pressure Poisson system
(iterations, SpMVs, milliseconds).  usage: probe_pressure_pc.py [N]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nsb200                      # noqa: E402
from nsb200 import _lib as L       # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 73
V = nsb200.VelocityBCType
mesh, markers = nsb200.hyper_cube(3, n)
ts = nsb200.BDFTimeStepping(0.0, 1e3, order=2, desired_start_time_step=1e-2)
solver = nsb200.IPCSSolver(mesh, markers, "standard", ts)
solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / 100.0))
solver.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3, 5, 6)] + [(V.constant, 4, (1.0, 0.0, 0.0))])
dev = solver.dev
for s in range(3):
    ts.update_coefficients()
    solver.solve()
    ts.advance_time()
    solver.advance_time()
print("krylov", solver.krylov_iterations[-1], flush=True)
# pressure system of the next step
ts.update_coefficients()
solver._update_time_stepping_coefficients()
dev.assemble_poisson(L.VEC, pn=solver._old_pressure, ustar=solver._intermediate_velocity, b=solver._bQ)
lmax = dev.lambda_max(solver._lap, solver._dinvQ, iters=30, work=solver._work)
print("lambda_max(D^-1 L) =", lmax, flush=True)
x0 = solver._pressure.clone()
ref = None
cfgs = [("jacobi", 1, 0.0)] + [("cheb", d, r) for d in (4, 8, 12, 16) for r in (30.0, 100.0, 300.0, 1000.0)]
for kind, deg, ratio in cfgs:
    pc = dev.precond(solver._dinvQ) if kind == "jacobi" else \
        dev.precond(solver._dinvQ, L.PC_CHEBYSHEV, deg, 1.1 * lmax / ratio, 1.1 * lmax)
    best = None
    for rep in range(2):
        x = x0.clone()
        torch.cuda.synchronize()
        t = time.perf_counter()
        try:
            its, res = dev.cg(solver._lap, pc, solver._bQ, x, rtol=1e-12, nullspace=not solver._pressure_dirichlet,
                              work=solver._work)
        except RuntimeError as e:
            print(kind, deg, ratio, "FAILED", e, flush=True)
            its = None
            break
        torch.cuda.synchronize()
        el = (time.perf_counter() - t) * 1e3
        best = el if best is None else min(best, el)
    if its is None:
        continue
    if ref is None:
        ref = x.clone()
    err = float((x - ref).abs().max() / ref.abs().max())
    print("%-7s degree %2d ratio %6.0f: %4d its %5d spmv %8.2f ms  res %.2e  diff vs jacobi %.1e"
          % (kind, deg, ratio, its, its * max(deg, 1), best, res, err), flush=True)
