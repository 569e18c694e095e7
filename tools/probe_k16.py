#!/usr/bin/env python3
"""GPU probe of the hand-written K16 solver on the DFG cylinder mesh: GMRES / inner CG iteration counts per Newton update
for a few preconditioner settings and mesh sizes."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch    # noqa: E402
import nsb200   # noqa: E402

V = nsb200.VelocityBCType
h = 4.1


def inlet(x, t=0.0):
    return np.stack([6.0 * x[:, 1] / h * (1.0 - x[:, 1] / h), 0.0 * x[:, 0]], axis=1)


for m in [int(a) for a in sys.argv[1:]] or [16, 40]:
    for opts in (dict(), dict(inner_rtol=1e-4), dict(vel_ratio=8.0), dict(vel_degree=12), dict(inner_rtol=1e-3, vel_ratio=8.0, vel_degree=12)):
        mesh, markers, ids = nsb200.channel_with_cylinder(m)
        ts = nsb200.BDFTimeStepping(0.0, 80.0, order=2, desired_start_time_step=0.005)
        s = nsb200.ImplicitBDFSolver(mesh, markers, "standard", ts, linear_solver="gmres")
        s.saddle_options.update(opts)
        s.pin_pressure = False
        s.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=0.01))
        s.set_boundary_conditions([(V.function, ids["inlet"], inlet), (V.no_slip, ids["cylinder"], None),
                                   (V.no_slip, ids["upper wall"], None), (V.no_slip, ids["lower wall"], None)])
        s.set_initial_conditions(dict(velocity=(0.0, 0.0)))
        out = dict(m=m, dofs=s.dev.spaces.ndofs_W, options=opts)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        try:
            for _ in range(3):
                ts.update_coefficients()
                s.solve()
                ts.advance_time()
                s.advance_time()
            out["ok"] = True
        except RuntimeError as e:
            out["error"] = str(e)[:100]
        torch.cuda.synchronize()
        out.update(seconds=time.perf_counter() - t0, newton=s.newton_iterations, gmres=s.krylov_iterations, inner=s.inner_iterations,
                   vel_lmax=s._vel_lmax, newton_history=["%.2e" % r for r in s.newton_history])
        print(json.dumps(out), flush=True)
        s.dev.close() if hasattr(s.dev, "close") else None
