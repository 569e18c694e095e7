"""Host side of the mesh-partitioned multi-GPU path: partition invariants in one
process, the halo plans and the owned-dof reductions across two ``gloo`` ranks on CPU, and (GPU,
>= 2 devices) the partitioned IPCS solver against the single-GPU solver through tools/run_partitioned.py."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [(2, (6, 5), 3, None), (2, (6, 5), 2, 0), (3, (4, 3, 5), 2, None), (3, (3, 3, 8), 4, None)]


def _mesh(dim, n):
    import nsb200
    return nsb200.hyper_rectangle((0.0,) * dim, (1.0,) * dim, n)[0]


def _field(gid):
    return np.sin(0.37 * gid) + gid


@pytest.mark.parametrize("dim,n,nranks,axis", CASES)
def test_partition_invariants(dim, n, nranks, axis):
    import nsb200
    from nsb200 import partition as P
    mesh = _mesh(dim, n)
    parts = P.connect([P.build_partition(mesh, r, nranks, axis=axis) for r in range(nranks)])
    gsp = nsb200.TaylorHoodSpaces(mesh)
    gx = gsp.node_coordinates()
    gn = np.concatenate([mesh.cells, mesh.num_vertices + mesh.cell_edges], axis=1)
    for name, nglob in (("nodes", gsp.n_nodes), ("verts", gsp.n_vertices)):
        owned = np.concatenate([getattr(p, name).owned_gid for p in parts])
        assert len(owned) == nglob and len(np.unique(owned)) == nglob          # every index owned exactly once
        for p in parts:
            pl = getattr(p, name)
            assert pl.n_global == nglob and np.all(np.diff(pl.owned_gid) > 0)
            ghost = np.full(pl.n_ghost, np.nan)
            for i, q in enumerate(pl.peers):                                    # what the peers would send
                qp = getattr(parts[q], name)
                j = list(qp.peers).index(p.rank)
                idx = qp.send_idx[qp.send_ptr[j]:qp.send_ptr[j + 1]]
                assert np.all(idx < qp.n_owned)
                ghost[pl.recv_ptr[i]:pl.recv_ptr[i + 1]] = _field(qp.owned_gid[idx])
            assert np.array_equal(ghost, _field(pl.ghost_gid))
    for p in parts:
        sp = nsb200.TaylorHoodSpaces(p.mesh, p.node_perm)
        assert np.allclose(sp.node_coordinates(), gx[p.node_gid])               # owned-first numbering is consistent
        assert np.array_equal(np.sort(p.node_perm), np.arange(sp.n_nodes))
        # owned rows are complete: every global cell touching an owned node is local
        touching = np.isin(gn, p.nodes.owned_gid).any(axis=1)
        assert np.array_equal(np.nonzero(touching)[0], p.cell_gid)
        # markers: real boundary facets keep theirs, cut faces get 0, ids are the global set
        assert p.mesh.boundary_ids() == mesh.boundary_ids()
        assert set(np.unique(p.mesh.bfacet_marker)) <= mesh.boundary_ids() | {0}
        # Dirichlet tables agree with the global ones on the owned dofs
        bcs = [(1, None, lambda x: x[:, :dim] * 2.0 + 1.0), (4, 0, 3.0)]
        gd, gv = gsp.dirichlet_table(bcs, "V")
        ld, lv = sp.dirichlet_table(bcs, "V")
        keep = ld < dim * p.nodes.n_owned
        lg = dim * p.node_gid[ld[keep] // dim] + ld[keep] % dim
        sel = np.isin(gd, dim * p.nodes.owned_gid[:, None] + np.arange(dim)[None, :])
        order = np.argsort(lg)
        assert np.array_equal(lg[order], gd[sel]) and np.allclose(lv[keep][order], gv[sel])


@pytest.mark.parametrize("dim,n,nranks,axis", CASES + [(3, (3, 3, 16), 8, None)])
def test_peer_memory_tables_emulated_push(dim, n, nranks, axis):
    """The tables handed to nsb_p2p_connect, checked by emulating the push kernel on the host: every rank
    stores its boundary values into its neighbours' ring slots at (offset, count) and every rank then finds
    exactly its ghosts there -- for interleaved widths 3 and 2 (velocity nodes / float4-padded nodes)."""
    from nsb200 import partition as P
    mesh = _mesh(dim, n)
    parts = P.connect([P.build_partition(mesh, r, nranks, axis=axis) for r in range(nranks)])
    for name in ("nodes", "verts"):
        plans = [getattr(p, name) for p in parts]
        layouts = [pl.peer_memory_layout() for pl in plans]
        for bs in (3, 2, 1):
            rings = [np.full(max(pl.n_ghost, 1) * 3, np.nan) for pl in plans]         # one ring slot per rank
            for p, pl in zip(parts, plans):
                assert len(pl.peers) <= 8
                nghost, off, slot = pl.peer_memory_tables(p.rank, layouts)
                x = np.stack([_field(pl.owned_gid) + c for c in range(bs)], axis=1).reshape(-1)
                for i, r in enumerate(pl.peers):
                    idx = pl.send_idx[pl.send_ptr[i]:pl.send_ptr[i + 1]]
                    assert off[i] + len(idx) <= nghost[i] == plans[r].n_ghost and plans[r].peers[slot[i]] == p.rank
                    vals = x.reshape(-1, bs)[idx].reshape(-1)
                    rings[r][off[i] * bs: off[i] * bs + len(vals)] = vals
            for pl, ring in zip(plans, rings):
                want = np.stack([_field(pl.ghost_gid) + c for c in range(bs)], axis=1).reshape(-1)
                assert np.array_equal(ring[:pl.n_ghost * bs], want)
    # a broken pairing is refused
    bad = [dict(l) for l in [pl.peer_memory_layout() for pl in (p.nodes for p in parts)]]
    bad[int(parts[0].nodes.peers[0])]["recv_ptr"] = [v + (1 if k else 0) for k, v in enumerate(bad[int(parts[0].nodes.peers[0])]["recv_ptr"])]
    with pytest.raises(ValueError):
        parts[0].nodes.peer_memory_tables(0, bad)


def test_partition_rejects_too_many_ranks():
    from nsb200 import partition as P
    with pytest.raises(ValueError):
        P.build_partition(_mesh(2, (4, 2)), 0, 3)


def _gloo_worker(rank, world, port, dim, n, out):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    import nsb200
    from nsb200 import partition as P
    mesh = _mesh(dim, n)
    p = P.connect_distributed(P.build_partition(mesh, rank, world))
    ok = True
    for name, bs in (("nodes", dim), ("verts", 1)):
        pl = getattr(p, name)
        x = np.concatenate([_field(pl.owned_gid), np.full(pl.n_ghost, np.nan)])
        x = np.repeat(x, bs).reshape(-1, bs) + np.arange(bs)[None, :]
        # the exchange of comm.cu (pack -> send / recv in place), spelled with gloo point-to-point calls
        reqs, bufs = [], []
        for i, q in enumerate(pl.peers):
            s = torch.from_numpy(np.ascontiguousarray(x[pl.send_idx[pl.send_ptr[i]:pl.send_ptr[i + 1]]]))
            r = torch.empty((int(pl.recv_ptr[i + 1] - pl.recv_ptr[i]), bs), dtype=torch.float64)
            if s.numel():
                reqs.append(dist.isend(s, int(q)))
            if r.numel():
                reqs.append(dist.irecv(r, int(q)))
            bufs.append(r)
        for r_ in reqs:
            r_.wait()
        for i in range(len(pl.peers)):
            x[pl.n_owned + pl.recv_ptr[i]:pl.n_owned + pl.recv_ptr[i + 1]] = bufs[i].numpy()
        gid = np.concatenate([pl.owned_gid, pl.ghost_gid])
        ok &= bool(np.array_equal(x, _field(gid)[:, None] + np.arange(bs)[None, :]))
        # inner products: owned entries + all-reduce == the global inner product
        t = torch.tensor([float(np.dot(x[:pl.n_owned, 0], x[:pl.n_owned, 0]))], dtype=torch.float64)
        dist.all_reduce(t)
        g = _field(np.arange(pl.n_global))
        ok &= bool(abs(t.item() - np.dot(g, g)) <= 1e-12 * np.dot(g, g))
    flag = torch.tensor([1.0 if ok else 0.0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        with open(out, "w") as f:
            json.dump({"ok": bool(flag.item() == 1.0)}, f)
    dist.destroy_process_group()


@pytest.mark.parametrize("dim,n", [(2, (5, 6)), (3, (3, 3, 4))])
def test_halo_plans_over_gloo_world_size_2(tmp_path, dim, n):
    import torch.multiprocessing as mp
    out = str(tmp_path / "result.json")
    port = 29600 + (os.getpid() + dim) % 300
    mp.spawn(_gloo_worker, args=(2, port, dim, n, out), nprocs=2, join=True)
    with open(out) as f:
        assert json.load(f)["ok"]


@pytest.mark.gpu
@pytest.mark.parametrize("transport", ["p2p", "nccl"])
@pytest.mark.parametrize("cfg", ["--dim 3 --cells 6 --flow cavity", "--dim 3 --cells 5 --flow channel",
                                 "--dim 2 --cells 12 --flow channel"])
def test_partitioned_ipcs_matches_single_gpu(cfg, transport):
    """2 ranks / 2 GPUs through torchrun: fields of the partitioned run equal the single-GPU run to
    1e-8 (observed 1e-13) with identical Newton and Krylov iteration counts, with the halo exchange
    over NVLink peer memory (default) and over NCCL send/recv (NSB_HALO=nccl)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tools", "run_partitioned.py")] + cfg.split()
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT,
                         env=dict(os.environ, NSB_HALO=transport))
    lines = [l for l in res.stdout.splitlines() if l.startswith("{")]
    assert res.returncode == 0 and lines, res.stderr[-3000:]
    rep = json.loads(lines[-1])
    assert rep["halo_transport"] == transport
    assert rep["ok"] and rep["rel_err_velocity"] < 1e-8 and rep["rel_err_pressure"] < 1e-8
    assert rep["newton"]["partitioned"] == rep["newton"]["single"]


# ---- replicated pressure solve (ipcs.py::_solve_pressure_replicated): host logic over gloo --------
def _p1_stiffness_apply(coords, cells, p):
    """y = K p for the P1 Laplacian, cell by cell (plain numpy; rows of vertices whose cells are all present are complete)"""
    d = coords.shape[1]
    X = coords[cells]
    J = X[:, 1:, :] - X[:, :1, :]                                  # rows = edge vectors
    det = np.abs(np.linalg.det(J))
    fact = {2: 2.0, 3: 6.0}[d]
    Gref = np.concatenate([-np.ones((1, d)), np.eye(d)], axis=0)   # reference gradients of the barycentric functions
    G = np.einsum("ak,cki->cai", Gref, np.linalg.inv(J).transpose(0, 2, 1))
    Ke = np.einsum("c,cai,cbi->cab", det / fact, G, G)
    y = np.zeros(len(coords))
    np.add.at(y, cells, np.einsum("cab,cb->ca", Ke, p[cells]))
    return y


def _replicated_worker(rank, world, port, dim, n, out):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    import nsb200
    from nsb200 import partition as P
    mesh = _mesh(dim, n)
    part = P.build_partition(mesh, rank, world)
    nQo, gid = part.verts.n_owned, part.vert_gid
    p_glob = _field(np.arange(mesh.num_vertices) * 1.0)
    # owned rows of the LOCAL operator applied to the local [owned | ghost] view of a global field ...
    y_loc = _p1_stiffness_apply(part.mesh.coords, part.mesh.cells, p_glob[gid])
    # ... scattered into a zeroed global vector and summed over ranks (one contributor per entry) ...
    b_g = np.zeros(mesh.num_vertices)
    b_g[gid[:nQo]] = y_loc[:nQo]
    t = torch.from_numpy(b_g)
    dist.all_reduce(t)
    # ... equal the global operator applied to the global field; every rank holds the same bits afterwards
    y_ref = _p1_stiffness_apply(part.global_mesh.coords, part.global_mesh.cells, p_glob)
    ok = bool(np.allclose(t.numpy(), y_ref, rtol=0, atol=1e-12 * np.abs(y_ref).max()))
    same = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(same, t)
    ok &= all(bool(torch.equal(same[0], s)) for s in same)
    # gather back: local [owned | ghost] entries of the global solution
    ok &= bool(np.array_equal(p_glob[gid], _field(gid * 1.0)))
    ok &= len(np.unique(gid)) == len(gid)
    cnt = torch.zeros(mesh.num_vertices, dtype=torch.float64)
    cnt[torch.from_numpy(gid[:nQo].astype(np.int64))] = 1.0
    dist.all_reduce(cnt)
    ok &= bool((cnt == 1.0).all())                                  # each global vertex is owned exactly once
    flag = torch.tensor([1.0 if ok else 0.0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        with open(out, "w") as f:
            json.dump({"ok": bool(flag.item() == 1.0)}, f)
    dist.destroy_process_group()


@pytest.mark.parametrize("dim,n", [(2, (5, 6)), (3, (3, 3, 4))])
def test_replicated_pressure_rhs_over_gloo_world_size_2(tmp_path, dim, n):
    """scatter of the owned right-hand-side rows + one all-reduce reproduces the global right-hand side bit-identically
    on every rank, and the global solution gathers back to the local numbering"""
    import torch.multiprocessing as mp
    out = str(tmp_path / "result.json")
    port = 29900 + (os.getpid() + dim) % 90
    mp.spawn(_replicated_worker, args=(2, port, dim, n, out), nprocs=2, join=True)
    with open(out) as f:
        assert json.load(f)["ok"]
