#!/usr/bin/env python3
"""Step time of the C5 cavity against the degree / interval of the momentum Chebyshev preconditioner (development probe).
usage: python tools/probe_cheb_sweep.py N STEPS "degree,ratio" ..."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                       # noqa: E402


def main():
    n, steps = int(sys.argv[1]), int(sys.argv[2])
    for spec in sys.argv[3:]:
        deg, ratio = spec.split(",")
        solver, ts = bench.make_cavity_solver(n, chebyshev_degree=int(deg), chebyshev_ratio=float(ratio))
        kry = []
        for s in range(steps + 6):
            if s == 6:
                torch.cuda.synchronize()
                t0 = time.time()
            ts.update_coefficients()
            solver.solve()
            kry.append(solver.krylov_iterations[-1]["momentum"])
            ts.advance_time()
            solver.advance_time()
        torch.cuda.synchronize()
        rec = dict(n=n, degree=int(deg), ratio=float(ratio), ms_per_step=1e3 * (time.time() - t0) / steps,
                   momentum_iterations=kry[6:], fallbacks=solver.chebyshev_fallbacks, checksums=solver.checksums())
        print(json.dumps(rec), flush=True)
        os.makedirs("gpurun_out", exist_ok=True)
        with open("gpurun_out/probe_cheb_sweep.jsonl", "a") as fh:
            fh.write(json.dumps(rec) + "\n")
        del solver
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
