#!/usr/bin/env python3
"""Timing of BASELINE.json configs C1-C4 at their named sizes on one GPU (bench.py covers C5).

  python tools/run_configs.py [c1] [c2] [c3] [c4]  [--steps K]

One JSON line per config: steps/s (or solves/s), Newton iterations, seconds of the hand-written
path and -- reported separately, as BASELINE.json's north star asks -- of the K16 library leg
(cuSOLVER sparse QR of the monolithic Newton system).  Parity of the same configs against the
oracle is tests/test_gpu_configs.py (sizes the LU oracle finishes in seconds).
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import torch                                # noqa: E402
import nsb200                               # noqa: E402

V, P = nsb200.VelocityBCType, nsb200.PressureBCType


def _sync():
    torch.cuda.synchronize()


def _report(name, solver, wall, units, extra):
    classes = solver.dev.timer_report()
    kernel_ms = sum(v[0] for v in classes.values())
    try:
        dsec, dn = solver.dev.direct_seconds()
    except Exception:
        dsec, dn = 0.0, 0
    if getattr(solver, "krylov_iterations", None) and not isinstance(solver.krylov_iterations[-1], dict):
        extra = dict(extra, k16_gmres_iterations=solver.krylov_iterations, k16_inner_cg_iterations=solver.inner_iterations)
    line = {"config": name, "mixed_dofs": solver.dev.spaces.ndofs_W, "cells": solver.dev.mesh.num_cells,
            "units": units, "wall_s": wall, "units_per_s": units / wall,
            "handwritten_kernel_ms": kernel_ms, "k16_direct_solve_s": dsec, "k16_solves": dn,
            "kernel_classes_ms": {k: round(v[0], 3) for k, v in classes.items() if v[1]},
            "newton_iterations": solver.newton_iterations}
    line.update(extra)
    print(json.dumps(line), flush=True)


def c1(steps):
    mesh, markers = nsb200.hyper_cube(2, 25)
    solver = nsb200.StationarySolver(mesh, markers, "standard")
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / 100.0))
    solver.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3)] + [(V.constant, 4, (1.0, 0.0))])
    solver.dev.timer_enable(True)
    _sync()
    t = time.perf_counter()
    r = solver.solve()
    _sync()
    _report("C1 stationary cavity Re=100, 25x25", solver, time.perf_counter() - t, 1, {"unit": "solves", "residual": r})


def _k16_timing(solver):
    """seconds of ONE Newton update of the current state by (a) the hand-written block-preconditioned GMRES, (b) cuSOLVER
    sparse QR (library leg, only below 100 k dofs: 26 s at 200 k), (c) SuperLU (scipy splu) on the host -- the
    reference's LU stand-in -- on the SAME Jacobian and right-hand side; relative difference of (a) against (c)"""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla
    dev = solver.dev
    r0 = solver._residual()
    solver._jacobian()
    pat = dev.patterns["WW"]
    out = {}
    _sync()
    t = time.perf_counter()
    solver._gmres_update(r0)
    _sync()
    out["gmres_s"] = time.perf_counter() - t
    out["gmres_iterations"] = solver.krylov_iterations[-1]
    x_g = solver._dx.cpu().numpy().copy()
    A = sp.csr_matrix((solver._A.cpu().numpy(), pat.col_host, pat.rowptr_host), shape=(pat.nrows, pat.nrows))
    b = solver._b.cpu().numpy()
    t = time.perf_counter()
    x_lu = spla.splu(A.tocsc()).solve(b)
    out["host_superlu_s"] = time.perf_counter() - t
    out["gmres_vs_lu_rel_diff"] = float(np.abs(x_g - x_lu).max() / np.abs(x_lu).max())
    if pat.nrows < 100000:
        out["cusolver_qr_s"] = dev.direct_solve("WW", solver._A, solver._b, solver._dx, 1)
    return out


def _time_steps(solver, ts, steps):
    solver.dev.timer_enable(True)
    for s in range(steps + 1):
        if s == 1:                           # first step (BDF-1 start-up + setup) is warm-up
            solver.dev.timer_report()
            if hasattr(solver.dev, "direct_seconds"):
                solver.dev.direct_seconds(reset=True)
            _sync()
            t = time.perf_counter()
        ts.update_coefficients()
        solver.solve()
        ts.advance_time()
        solver.advance_time()
    _sync()
    return time.perf_counter() - t


def c2(steps):
    def inflow(x, t):
        return np.stack([6.0 * x[:, 1] * (1.0 - x[:, 1]) * (1.0 + 0.5 * np.sin(np.pi * t)), 0.0 * x[:, 0]], axis=1)
    mesh, markers = nsb200.hyper_rectangle((0.0, 0.0), (10.0, 1.0), (470, 47))
    ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=0.01)
    solver = nsb200.ImplicitBDFSolver(mesh, markers, "standard", ts)
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=0.1))
    solver.set_boundary_conditions([(V.function, 1, nsb200.Expression(inflow)), (V.no_slip, 3, None),
                                    (V.no_slip, 4, None), (P.constant, 2, 0.0)])
    wall = _time_steps(solver, ts, steps)
    ts.update_coefficients()
    solver._set_time(ts.next_time)
    solver._update_bc_values()
    solver._update_time_stepping_coefficients()
    _report("C2 channel BDF-2 monolithic, 470x47", solver, wall, steps, {"unit": "steps", "k16_one_newton_update": _k16_timing(solver)})


def c3(steps, n=666):
    g, Re = 2.0 * np.pi, 100.0

    def u_ex(x, t):
        f = np.exp(-2.0 * g ** 2 * t / Re)
        return np.stack([np.cos(g * x[:, 0]) * np.sin(g * x[:, 1]) * f, -np.sin(g * x[:, 0]) * np.cos(g * x[:, 1]) * f], axis=1)

    def p_ex(x, t):
        return -0.25 * (np.cos(2 * g * x[:, 0]) + np.cos(2 * g * x[:, 1])) * np.exp(-4.0 * g ** 2 * t / Re)
    mesh, markers = nsb200.hyper_cube(2, n)
    ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=1e-3)
    solver = nsb200.IPCSSolver(mesh, markers, "standard", ts)
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
    solver.set_boundary_conditions([(V.function, w, nsb200.Expression(u_ex)) for w in (1, 2, 3, 4)] +
                                   [(P.mean_value, None, 0.0)])
    solver.set_initial_conditions(dict(velocity=nsb200.Expression(u_ex, degree=3), pressure=nsb200.Expression(p_ex, degree=3)))
    wall = _time_steps(solver, ts, steps)
    t = (steps + 1) * 1e-3
    eu = solver.dev.errornorm(lambda x: u_ex(x, t), solver._velocities[1], "V")
    ep = solver.dev.errornorm(lambda x: p_ex(x, t), solver._old_pressure, "Q")
    _report("C3 Taylor-Green IPCS, %dx%d" % (n, n), solver, wall, steps,
            {"unit": "steps", "errornorm_velocity": eu, "errornorm_pressure": ep, "cfl": solver.cfl_number(1e-3),
             "krylov_iterations": solver.krylov_iterations[-1], "chebyshev_fallbacks": solver.chebyshev_fallbacks})


def c4(steps, m=40):
    h = 4.1
    mesh, markers, ids = nsb200.channel_with_cylinder(m)

    def inlet(x, t=0.0):
        return np.stack([6.0 * x[:, 1] / h * (1.0 - x[:, 1] / h), 0.0 * x[:, 0]], axis=1)
    ts = nsb200.BDFTimeStepping(0.0, 80.0, order=2, desired_start_time_step=0.005)
    solver = nsb200.ImplicitBDFSolver(mesh, markers, "standard", ts)
    solver.pin_pressure = False
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=0.01))
    solver.set_boundary_conditions([(V.function, ids["inlet"], inlet), (V.no_slip, ids["cylinder"], None),
                                    (V.no_slip, ids["upper wall"], None), (V.no_slip, ids["lower wall"], None)])
    solver.set_initial_conditions(dict(velocity=(0.0, 0.0)))
    wall = _time_steps(solver, ts, steps)
    F = solver.boundary_force(solver._solutions[1], ids["cylinder"])
    ts.update_coefficients()
    solver._set_time(ts.next_time)
    solver._update_bc_values()
    solver._update_time_stepping_coefficients()
    _report("C4 DFG 2D-2 BDF-2 on the synthetic cylinder mesh (m=%d)" % m, solver, wall, steps,
            {"unit": "steps", "drag_coefficient": 2.0 * F[0], "lift_coefficient": 2.0 * F[1],
             "k16_one_newton_update": _k16_timing(solver)})


if __name__ == "__main__":
    argv = sys.argv[1:]
    args = [a for i, a in enumerate(argv) if not a.startswith("--") and not (i > 0 and argv[i - 1] == "--steps")]
    steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 2
    todo = args or ["c1", "c4", "c3", "c2"]
    for name in todo:
        fn = {"c1": c1, "c2": c2, "c3": c3, "c4": c4}[name]
        try:
            fn(steps)
        except Exception as e:      # report and go on: one config must not hide the others
            print(json.dumps({"config": name, "error": "%s: %s" % (type(e).__name__, e)}), flush=True)
