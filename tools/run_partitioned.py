#!/usr/bin/env python3
"""Mesh-partitioned IPCS run, one rank per GPU, checked against the single-GPU run.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29517 tools/run_partitioned.py [--dim 3] [--cells 8] [--steps 2] [--flow cavity|channel]

Every rank advances its partition; rank 0 then runs the same problem un-partitioned on its own
GPU, gathers the owned velocity / pressure entries of all ranks by global id and prints ONE JSON
line with the relative field differences (bar: 1e-8, BASELINE.json north star) and the Newton /
Krylov iteration counts of both runs.  Exit code 1 if the bar is missed.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def make_problem(nsb200, dim, n, flow):
    V, P = nsb200.VelocityBCType, nsb200.PressureBCType
    if flow == "cavity":
        mesh, markers = nsb200.hyper_cube(dim, n)
        walls = (1, 2, 3) if dim == 2 else (1, 2, 3, 5, 6)
        lid = (1.0, 0.0) if dim == 2 else (1.0, 0.0, 0.0)
        bcs = [(V.no_slip, w, None) for w in walls] + [(V.constant, 4, lid)]
    else:
        p1 = (2.0, 1.0) if dim == 2 else (2.0, 1.0, 1.0)
        nn = (2 * n, n) if dim == 2 else (2 * n, n, n)
        mesh, markers = nsb200.hyper_rectangle((0.0,) * dim, p1, nn)

        def inflow(x, t):
            u = np.zeros((len(x), dim))
            u[:, 0] = 6.0 * x[:, 1] * (1.0 - x[:, 1]) * (1.0 + 0.5 * np.sin(np.pi * t))
            return u
        walls = (3, 4) if dim == 2 else (3, 4, 5, 6)
        bcs = [(V.function, 1, nsb200.Expression(inflow))] + [(V.no_slip, w, None) for w in walls] + \
              [(P.constant, 2, 0.0)]
    return mesh, markers, bcs


def run(nsb200, mesh, markers, bcs, steps, device_index, partition=None, pc="chebyshev"):
    ts = nsb200.BDFTimeStepping(0.0, 1.0e6, order=2, desired_start_time_step=0.01)
    solver = nsb200.IPCSSolver(mesh, markers, "standard", ts, device_index=device_index, partition=partition,
                               momentum_preconditioner=pc)
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=0.02))
    solver.set_boundary_conditions(bcs)
    for _ in range(steps):
        ts.update_coefficients()
        solver.solve()
        ts.advance_time()
        solver.advance_time()
    return solver


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dim", type=int, default=3)
    ap.add_argument("--cells", dest="n", type=int, default=8)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--flow", default="cavity", choices=["cavity", "channel"])
    ap.add_argument("--axis", type=int, default=None)
    ap.add_argument("--pc", default="chebyshev", choices=["chebyshev", "jacobi"])
    ap.add_argument("--single-first", action="store_true", help="debug: run the un-partitioned problem first")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")           # host plumbing only: NCCL id broadcast, plan exchange, gathers
    import __graft_entry__
    if rank == 0:
        __graft_entry__.build()
    dist.barrier()
    import nsb200
    from nsb200 import partition as part
    mesh, markers, bcs = make_problem(nsb200, args.dim, args.n, args.flow)
    if args.single_first and rank == 0:
        ref = run(nsb200, mesh, markers, bcs, args.steps, local, pc=args.pc)
        print("single-GPU run ok:", ref.newton_iterations, ref.krylov_iterations, flush=True)
    p = part.connect_distributed(part.build_partition(mesh, rank, world, axis=args.axis))
    solver = run(nsb200, mesh, markers, bcs, args.steps, local, partition=p, pc=args.pc)
    d = mesh.dim
    nVo, nQo = solver._nV_own, solver._nQ_own
    mine = dict(node_gid=p.nodes.owned_gid, vert_gid=p.verts.owned_gid,
                u=solver._velocities[1][:nVo].cpu().numpy(), p=solver._old_pressure[:nQo].cpu().numpy(),
                newton=solver.newton_iterations, krylov=solver.krylov_iterations,
                launches=solver.dev.launches, cells=p.mesh.num_cells, ghosts=p.nodes.n_ghost)
    gathered = [None] * world
    dist.gather_object(mine, gathered if rank == 0 else None, dst=0)
    ok = True
    if rank == 0:
        ref = run(nsb200, mesh, markers, bcs, args.steps, local, pc=args.pc)
        uref, pref = ref._velocities[1].cpu().numpy(), ref._old_pressure.cpu().numpy()
        u, pr = np.full_like(uref, np.nan), np.full_like(pref, np.nan)
        for g in gathered:
            u.reshape(-1, d)[g["node_gid"]] = g["u"].reshape(-1, d)
            pr[g["vert_gid"]] = g["p"]
        eu = float(np.abs(u - uref).max() / np.abs(uref).max())
        if not ref._pressure_dirichlet:          # all-Neumann: both runs return the zero-mean representative
            pr, pref = pr - pr.mean(), pref - pref.mean()
        ep = float(np.abs(pr - pref).max() / np.abs(pref).max())
        ok = bool(eu < 1e-8 and ep < 1e-8)
        print(json.dumps({"check": "partitioned IPCS vs single GPU", "ranks": world, "dim": d, "n": args.n,
                          "flow": args.flow, "steps": args.steps, "rel_err_velocity": eu, "rel_err_pressure": ep,
                          "ok": ok, "newton": {"partitioned": gathered[0]["newton"], "single": ref.newton_iterations},
                          "krylov": {"partitioned": gathered[0]["krylov"], "single": ref.krylov_iterations},
                          "cells_per_rank": [g["cells"] for g in gathered],
                          "ghost_nodes_per_rank": [g["ghosts"] for g in gathered],
                          "launches_per_rank": [g["launches"] for g in gathered],
                          "halo_transport": solver.dev.halo_transport}), flush=True)
    solver.dev.close()
    flag = [ok]
    dist.broadcast_object_list(flag, src=0)
    dist.destroy_process_group()
    sys.exit(0 if flag[0] else 1)


if __name__ == "__main__":
    main()
