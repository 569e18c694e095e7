#!/usr/bin/env python3
"""Development probe: K16 (cuSOLVER sparse QR) time per reordering on the DFG mixed Jacobian.  usage: probe_direct.py [m]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nsb200                      # noqa: E402
from nsb200 import _lib as L       # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 24
V = nsb200.VelocityBCType
mesh, markers, ids = nsb200.channel_with_cylinder(m)
h = 4.1
ts = nsb200.BDFTimeStepping(0.0, 80.0, order=2, desired_start_time_step=0.005)
solver = nsb200.ImplicitBDFSolver(mesh, markers, "standard", ts)
solver.pin_pressure = False
solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=0.01))
solver.set_boundary_conditions([(V.function, ids["inlet"], lambda x: np.stack([6.0 * x[:, 1] / h * (1.0 - x[:, 1] / h), 0.0 * x[:, 0]], axis=1)),
                                (V.no_slip, ids["cylinder"], None), (V.no_slip, ids["upper wall"], None),
                                (V.no_slip, ids["lower wall"], None)])
solver._setup_boundary_conditions()
ts.update_coefficients()
solver._update_time_stepping_coefficients()
r = solver._residual()
solver._jacobian()
dev = solver.dev
print("mixed dofs", solver._nW, "nnz", solver._pat.nnz, "residual", r, flush=True)
ref = None
for reorder in (1, 2, 3, 0):
    x = dev.empty(solver._nW)
    t = time.perf_counter()
    try:
        sec = dev.direct_solve("WW", solver._A, solver._b, x, reorder)
    except RuntimeError as e:
        print("reorder", reorder, "FAILED", e, flush=True)
        continue
    if ref is None:
        ref = x.clone()
    print("reorder %d: %.2f s  diff %.2e" % (reorder, sec, float((x - ref).abs().max() / ref.abs().max())), flush=True)
