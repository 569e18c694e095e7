#!/usr/bin/env python3
"""Jacobian assembly (k_jacobian_rowgroups) under different scatter-schedule settings (development probe).
usage: python tools/probe_jac.py N "SPLIT,BANKS,SLACK" ...   -> one JSON line per setting in gpurun_out/probe_jac.jsonl"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nsb200                      # noqa: E402
from nsb200 import _lib as L       # noqa: E402
from probe_kernels import timed    # noqa: E402


def run(n, setting, ref):
    sp_, bk, sl = setting.split(",")
    os.environ.update(NSB_SCHED_SPLIT=sp_, NSB_SCHED_BANKS=bk, NSB_SCHED_SLACK=sl)
    V = nsb200.VelocityBCType
    mesh, markers = nsb200.hyper_cube(3, n)
    ts = nsb200.BDFTimeStepping(0.0, 1e3, order=2, desired_start_time_step=1e-2)
    solver = nsb200.IPCSSolver(mesh, markers, "standard", ts)
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / 100.0))
    solver.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3, 5, 6)] + [(V.constant, 4, (1.0, 0.0, 0.0))])
    dev, sp = solver.dev, solver.dev.spaces
    g = torch.Generator(device="cuda").manual_seed(7)
    u = dev.zeros(sp.ndofs_V)
    u.copy_(torch.rand(sp.ndofs_V, generator=g, device="cuda", dtype=torch.float64) - 0.5)
    dev.set_coefficients(1e-2, (1.5, -2.0, 0.5), c_conv=1.0, c_p=1.0, c_visc=0.01)
    A = solver._A_mom
    med, best = timed(lambda: dev.assemble_ipcs_momentum(L.MAT, u, A=A), reps=7)
    x, y = dev.zeros(sp.ndofs_V), dev.zeros(sp.ndofs_V)
    x.copy_(torch.rand(sp.ndofs_V, generator=g, device="cuda", dtype=torch.float64) - 0.5)
    dev.spmv(dev.matrix("P2P2", A, L.BSR, 3, None), x, y)
    torch.cuda.synchronize()
    yh = y.cpu().numpy().copy()
    rec = dict(n=n, split=int(sp_), banks=int(bk), slack=int(sl), ms=med, ms_best=best, y_l2=float(np.linalg.norm(yh)))
    if ref is not None:
        rec["rel_diff_vs_first"] = float(np.linalg.norm(yh - ref) / np.linalg.norm(ref))
    print(json.dumps(rec), flush=True)
    with open("gpurun_out/probe_jac.jsonl", "a") as f:
        f.write(json.dumps(rec) + "\n")
    del solver, dev, A
    torch.cuda.empty_cache()
    return yh


if __name__ == "__main__":
    os.makedirs("gpurun_out", exist_ok=True)
    n = int(sys.argv[1])
    ref = None
    for s in sys.argv[2:]:
        y = run(n, s, ref)
        ref = y if ref is None else ref
