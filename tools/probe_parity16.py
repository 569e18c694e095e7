#!/usr/bin/env python3
"""GPU probe: which Krylov tolerance sets the last digit of the 16^3 cavity parity (fields after 2 IPCS steps against
the C oracle at rtol 1e-14)?  One line per solver setting."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nsb200   # noqa: E402
import oracle   # noqa: E402

V = nsb200.VelocityBCType
n, Re, dt = int(os.environ.get("N", "16")), 100.0, 1e-2
om = oracle.mesh.hyper_rectangle((0.0, 0.0, 0.0), (1.0, 1.0, 1.0), (n, n, n))
vel = [(w, None, (0.0, 0.0, 0.0)) for w in (1, 2, 3, 5, 6)] + [(4, None, (1.0, 0.0, 0.0))]
orc = oracle.solvers.IPCSOracle(om, Re, vel, [], assembler="c")
orc.linear_solver, orc.krylov_rtol = "krylov-c", 1e-14
ots = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=dt)
for _ in range(2):
    ots.update_coefficients()
    a = list(ots.coefficients(1)) + [0.0]
    orc.solve(tuple(a[:3]), ots.get_next_step_size(), ots.next_time)
    ots.advance_time()
    orc.advance_time()
ou, op = orc.velocities[1], orc.old_pressure - orc.old_pressure.mean()
for kw in (dict(correction_rtol=1e-12), dict(correction_rtol=1e-12, linear_atol_factor=0.01), dict(),
           dict(linear_rtol=1e-13), dict(linear_rtol=1e-11, correction_rtol=1e-13)):
    mesh, markers = nsb200.hyper_cube(3, n)
    ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=dt)
    s = nsb200.IPCSSolver(mesh, markers, "standard", ts, **kw)
    s.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
    s.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3, 5, 6)] + [(V.constant, 4, (1.0, 0.0, 0.0))])
    for _ in range(2):
        ts.update_coefficients()
        s.solve()
        ts.advance_time()
        s.advance_time()
    u, p = s._velocities[1].cpu().numpy(), s._old_pressure.cpu().numpy()
    print(json.dumps(dict(settings=kw, velocity=float(np.abs(u - ou).max() / np.abs(ou).max()),
                          pressure=float(np.abs(p - p.mean() - op).max() / np.abs(op).max()), newton=s.newton_iterations,
                          krylov=s.krylov_iterations, newton_residual=s.newton_residual)), flush=True)
    s.dev.close()
