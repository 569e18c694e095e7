#!/usr/bin/env python3
"""Newton residual histories of the C5 cavity, with and without the forcing term of the first update.
usage: python tools/probe_newton.py N STEPS FORCING [FORCING ...]"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                       # noqa: E402


def main():
    n, steps = int(sys.argv[1]), int(sys.argv[2])
    ref = None
    for f in sys.argv[3:]:
        solver, ts = bench.make_cavity_solver(n, newton_forcing=float(f))
        hist = []
        for s in range(steps):
            ts.update_coefficients()
            solver.solve()
            hist.append(dict(res=solver.newton_residuals, kry=solver.krylov_iterations[-1]))
            ts.advance_time()
            solver.advance_time()
        torch.cuda.synchronize()
        u = solver._velocities[0].cpu().numpy().copy()
        rec = dict(n=n, forcing=float(f), steps=hist, checksums=solver.checksums())
        if ref is not None:
            rec["rel_diff_u_vs_first"] = float(np.linalg.norm(u - ref) / np.linalg.norm(ref))
        ref = u if ref is None else ref
        print(json.dumps(rec), flush=True)
        os.makedirs("gpurun_out", exist_ok=True)
        with open("gpurun_out/probe_newton.jsonl", "a") as fh:
            fh.write(json.dumps(rec) + "\n")
        del solver
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
