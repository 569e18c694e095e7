#!/usr/bin/env python3
"""Per-kernel CUDA-event timing of the libnsb200 hot path on a 3D cavity mesh (development probe).
usage: python tools/probe_kernels.py N [steps] [dt]"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nsb200                      # noqa: E402
from nsb200 import _lib as L       # noqa: E402


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(min(ts))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    dt = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-2
    V = nsb200.VelocityBCType
    t0 = time.time()
    mesh, markers = nsb200.hyper_cube(3, n)
    ts = nsb200.BDFTimeStepping(0.0, 1e3, order=2, desired_start_time_step=dt)
    solver = nsb200.IPCSSolver(mesh, markers, "standard", ts)
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / 100.0))
    solver.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3, 5, 6)] + [(V.constant, 4, (1.0, 0.0, 0.0))])
    dev, sp = solver.dev, solver.dev.spaces
    print("setup %.1fs  n=%d cells=%d dofW=%d nV=%d nQ=%d" % (time.time() - t0, n, mesh.num_cells, sp.ndofs_W,
                                                              sp.ndofs_V, sp.ndofs_Q), flush=True)
    p22, p11 = dev.patterns["P2P2"], dev.patterns["P1P1"]
    nV, nQ, d = sp.ndofs_V, sp.ndofs_Q, 3
    out = {}
    # step a few times to get a realistic state
    for s in range(steps):
        torch.cuda.synchronize()
        t = time.time()
        ts.update_coefficients()
        solver.solve()
        torch.cuda.synchronize()
        el = time.time() - t
        ts.advance_time()
        solver.advance_time()
        print("step %d: %.1f ms newton=%s krylov=%s" % (s, el * 1e3, solver.newton_iterations[-1],
                                                        solver.krylov_iterations[-1]), flush=True)
    u, p = solver._velocities[0], solver._pressure
    bV, bQ = solver._bV, solver._bQ
    hbm = 6546.6

    def rep(name, fn, bytes_alg):
        med, best = timed(fn)
        out[name] = dict(ms=med, ms_best=best, gbs=bytes_alg / med / 1e6, frac=bytes_alg / med / 1e6 / hbm,
                         mbytes=bytes_alg / 1e6)
        print("%-22s %8.3f ms (best %8.3f)  alg %9.1f MB  %7.1f GB/s  %.3f of measured HBM" %
              (name, med, best, bytes_alg / 1e6, bytes_alg / med / 1e6, bytes_alg / med / 1e6 / hbm), flush=True)

    geo = 8 * 3 * mesh.num_vertices + 4 * 4 * mesh.num_cells + 4 * 10 * mesh.num_cells
    rep("jacobian", lambda: dev.assemble_ipcs_momentum(L.MAT, u, A=solver._A_mom),
        8 * 9 * p22.nnz + geo + 8 * nV)
    rep("residual", lambda: dev.assemble_ipcs_momentum(L.VEC, u, solver._velocities[1], solver._velocities[2], p, b=bV),
        8 * nV + geo + 8 * (3 * nV + nQ))
    rep("mass_matrix", lambda: dev.assemble_correction(L.MAT, A=solver._A_mass), 8 * p22.nnz + geo)
    rep("laplace_matrix", lambda: dev.assemble_poisson(L.MAT, A=solver._A_lap),
        8 * p11.nnz + 8 * 3 * mesh.num_vertices + 16 * mesh.num_cells)
    rep("poisson_rhs", lambda: dev.assemble_poisson(L.VEC, pn=p, ustar=u, b=bQ), 8 * nQ + geo + 8 * (nV + nQ))
    rep("correction_rhs", lambda: dev.assemble_correction(L.VEC, ustar=u, pnew=p, pold=solver._old_pressure, b=bV),
        8 * nV + geo + 8 * (nV + 2 * nQ))
    x, y = dev.zeros(nV), dev.zeros(nV)
    x.uniform_(-1, 1)
    rep("spmv_bsr3", lambda: dev.spmv(solver._jac, x, y), 8 * 9 * p22.nnz + 4 * p22.nnz + 16 * nV + 4 * p22.nrows)
    rep("spmv_mass_kron", lambda: dev.spmv(solver._mass, x, y), 12 * p22.nnz + 16 * nV + 4 * p22.nrows)
    xq, yq = dev.zeros(nQ), dev.zeros(nQ)
    xq.uniform_(-1, 1)
    rep("spmv_laplace", lambda: dev.spmv(solver._lap, xq, yq), 12 * p11.nnz + 20 * nQ)
    rep("bc_rows_bsr", lambda: dev.bc_apply_matrix("P2P2", solver._A_mom, L.BSR, d, solver._vbc_dofs),
        8 * 9 * 30 * solver._vbc_dofs.numel() / 3)
    rep("ipcs_advance", lambda: dev.ipcs_advance(solver._velocities[0], solver._velocities[1], solver._velocities[2],
                                                 solver._pressure, solver._old_pressure), 8 * (4 * nV + 2 * nQ))
    rep("dot", lambda: dev.dot(x, y), 16 * nV)
    rep("axpy", lambda: dev.axpy(0.5, x, y), 24 * nV)
    out["fp64_tflops"] = dev.fp64_peak()
    print("fp64 peak %.1f TFLOP/s" % out["fp64_tflops"])
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/probe_n%d.json" % n, "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
