#!/usr/bin/env python3
"""Latency of the halo exchange on a partitioned cavity mesh, per transport (NSB_HALO=p2p|nccl).

  [NSB_HALO=nccl] python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29519 tools/probe_comm.py [--cells 73] [--reps 2000]

Back-to-back exchanges on one stream, CUDA events, max over ranks; one JSON line from rank 0.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cells", dest="n", type=int, default=73)
    ap.add_argument("--reps", type=int, default=2000)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    import nsb200
    from nsb200 import partition as part, _lib as L
    from nsb200.device import Device
    mesh, _ = nsb200.hyper_cube(3, args.n)
    p = part.connect_distributed(part.build_partition(mesh, rank, world))
    dev = Device(p.mesh, local, partition=p)
    nn, nv = p.nodes.n_owned + p.nodes.n_ghost, p.verts.n_owned + p.verts.n_ghost
    u, q = dev.zeros(3 * nn), dev.zeros(nv)
    u[:3 * p.nodes.n_owned] = float(rank + 1)
    q[:p.verts.n_owned] = float(rank + 1)
    out = {"transport": dev.halo_transport, "ranks": world, "cells_per_edge": args.n, "reps": args.reps,
           "ghost_nodes": p.nodes.n_ghost, "ghost_verts": p.verts.n_ghost}
    for name, which, bs, x in (("nodes_bs3", L.HALO_NODES, 3, u), ("verts_bs1", L.HALO_VERTS, 1, q)):
        for _ in range(20):
            dev.halo_exchange(which, bs, x)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            dev.halo_exchange(which, bs, x)
        e1.record()
        torch.cuda.synchronize()
        us = torch.tensor([e0.elapsed_time(e1) * 1e3 / args.reps], dtype=torch.float64)
        dist.all_reduce(us, op=dist.ReduceOp.MAX)
        out[name + "_us"] = float(us)
    # ghosts hold the owners' values: every ghost of this test vector is another rank's (rank + 1)
    g = q[p.verts.n_owned:]
    out["ghost_values_ok"] = bool(((g >= 1) & (g <= world) & (g != rank + 1)).all().item()) if g.numel() else True
    dev.halo_status()
    flags = [None] * world
    dist.all_gather_object(flags, out["ghost_values_ok"])
    out["ghost_values_ok"] = all(flags)
    if rank == 0:
        print(json.dumps(out), flush=True)
    dev.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
