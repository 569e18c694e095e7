"""Mesh partitioning for the multi-GPU path: one rank per GPU, cells split into
slabs, every P2 node / vertex owned by the lowest rank whose cells touch it.

A rank's local mesh holds every cell that touches one of its owned nodes (its own cells plus a
one-cell ghost layer), so each owned matrix row and residual entry is assembled completely from
local cells: no matrix entry ever crosses ranks, only vector halos (owner -> ghost copies before
an assembly or SpMV) and the all-reduced Krylov inner products do.

Local numbering is owned-first: nodes ``[owned, ascending global id | ghost, by (owner, global
id)]`` and the same rule for vertices, so that
  * the owned rows of every pattern are a prefix (the single-GPU kernels run unchanged on it),
  * the ghosts received from one peer are one contiguous range (NCCL receives in place).

The reference has no counterpart (it would be dolfin's MPI mesh distribution, never exercised by
its solvers); the single-GPU numbering contract of ``spaces.py`` is what a 1-rank partition
reduces to up to the documented permutation.  Setup only: nothing here runs per step.
"""
import numpy as np

from .grid import SimplexMesh


class HaloPlan:
    """Who sends what to whom for one index space (P2 nodes or vertices).

    ``recv_ptr[i]:recv_ptr[i+1]`` is the range of ghost slots (offsets into the ghost block, i.e.
    local index - n_owned) filled by ``peers[i]``; ``send_idx[send_ptr[i]:send_ptr[i+1]]`` are
    the local owned indices ``peers[i]`` needs, in the order of its ghost slots."""

    def __init__(self, n_owned, n_ghost, n_global, owned_gid, ghost_gid, ghost_owner):
        self.n_owned, self.n_ghost, self.n_global = int(n_owned), int(n_ghost), int(n_global)
        self.owned_gid, self.ghost_gid, self.ghost_owner = owned_gid, ghost_gid, ghost_owner
        self.peers = self.send_ptr = self.send_idx = self.recv_ptr = None

    def requests(self):
        """{owner rank: global ids of my ghosts it owns, in ghost-slot order}"""
        out = {}
        for q in np.unique(self.ghost_owner):
            out[int(q)] = self.ghost_gid[self.ghost_owner == q]
        return out

    def finalize(self, rank, requests_to_me):
        """``requests_to_me``: {peer rank: global ids that peer wants from me}."""
        mine = self.requests()
        peers = sorted(set(mine) | set(int(q) for q, g in requests_to_me.items() if len(g)))
        assert rank not in peers
        self.peers = np.asarray(peers, dtype=np.int32)
        recv_ptr, send_ptr, send_idx = [0], [0], []
        for q in peers:
            recv_ptr.append(recv_ptr[-1] + (len(mine[q]) if q in mine else 0))
            want = np.asarray(requests_to_me.get(q, np.zeros(0, np.int64)), dtype=np.int64)
            loc = np.searchsorted(self.owned_gid, want)
            if len(want) and not (loc.max() < self.n_owned and np.array_equal(self.owned_gid[loc], want)):
                raise ValueError("rank %d is asked by rank %d for nodes it does not own" % (rank, q))
            send_idx.append(loc.astype(np.int32))
            send_ptr.append(send_ptr[-1] + len(want))
        assert recv_ptr[-1] == self.n_ghost
        self.recv_ptr = np.asarray(recv_ptr, dtype=np.int32)
        self.send_ptr = np.asarray(send_ptr, dtype=np.int32)
        self.send_idx = np.concatenate(send_idx).astype(np.int32) if send_idx else np.zeros(0, np.int32)
        return self


    # ---- NVLink peer-memory transport (csrc/comm.cu nsb_p2p_*): what a rank publishes / derives ------------
    def peer_memory_layout(self):
        """what the neighbours need to know to PUSH into this rank's ring: (peers, recv_ptr, n_ghost)"""
        return dict(peers=[int(r) for r in self.peers], recv_ptr=[int(v) for v in self.recv_ptr], n_ghost=int(self.n_ghost))

    def peer_memory_tables(self, rank, layouts):
        """Per neighbour of this rank, in plan order: (its ghost count, the offset of this rank's part inside
        its ghost block, this rank's index in its peer list) -- the arguments of ``nsb_p2p_connect``.
        ``layouts[r]`` is rank r's ``peer_memory_layout()``.  Checks the pairing the push kernel relies on: the
        neighbour expects exactly as many entries from this rank as this rank sends (both ends size the
        exchange's grid from that count)."""
        nghost, off, slot = [], [], []
        for i, r in enumerate(self.peers):
            theirs = layouts[int(r)]
            if rank not in theirs["peers"]:
                raise ValueError("rank %d lists rank %d as a neighbour but not the other way round" % (rank, int(r)))
            s = theirs["peers"].index(rank)
            expects = theirs["recv_ptr"][s + 1] - theirs["recv_ptr"][s]
            sends = int(self.send_ptr[i + 1] - self.send_ptr[i])
            if expects != sends:
                raise ValueError("rank %d sends %d entries to rank %d, which expects %d" % (rank, sends, int(r), expects))
            # the flow control of the peer-memory exchange (csrc/comm.cu: acknowledgements per block of the pair) leans on
            # traffic in BOTH directions of every pair; a one-sided pair would let a pusher run ahead of a reader
            mine = int(self.recv_ptr[i + 1] - self.recv_ptr[i])
            if (sends == 0) != (mine == 0):
                raise ValueError("ranks %d and %d exchange in one direction only (%d sent, %d received): not supported by "
                                 "the peer-memory transport, use NSB_HALO=nccl" % (rank, int(r), sends, mine))
            nghost.append(theirs["n_ghost"])
            off.append(theirs["recv_ptr"][s])
            slot.append(s)
        return (np.asarray(nghost, dtype=np.int32), np.asarray(off, dtype=np.int32), np.asarray(slot, dtype=np.int32))


class Partition:
    """Local view of rank ``rank`` of ``nranks``: ``mesh`` (SimplexMesh with the global boundary
    markers carried over), ``node_perm`` (natural local P2 numbering -> owned-first numbering),
    halo plans ``nodes`` / ``verts`` and the local -> global maps."""

    def __init__(self, rank, nranks, mesh, node_perm, nodes, verts, vert_gid, node_gid, cell_gid):
        self.rank, self.nranks, self.mesh, self.node_perm = rank, nranks, mesh, node_perm
        self.nodes, self.verts = nodes, verts
        self.vert_gid, self.node_gid, self.cell_gid = vert_gid, node_gid, cell_gid


def cell_ranks(mesh, nranks, axis=None):
    """Slab decomposition: cut planes on grid planes of the vertex coordinates along ``axis``
    (default: the last axis), as many planes per slab as evenly possible."""
    axis = mesh.dim - 1 if axis is None else int(axis)
    x = mesh.coords[:, axis]
    planes = np.unique(np.round(x, 12))
    nlayers = len(planes) - 1
    if nlayers < nranks:
        raise ValueError("cannot cut %d layers into %d slabs" % (nlayers, nranks))
    cuts = planes[(np.arange(1, nranks) * nlayers) // nranks]
    centroid = x[mesh.cells].mean(axis=1)
    return np.searchsorted(cuts, centroid, side="right").astype(np.int32)


def global_node_owner(mesh, crank):
    """owner of every global P2 node (vertices then edges) = lowest rank of an incident cell"""
    nv = mesh.num_vertices
    gn = np.concatenate([mesh.cells.astype(np.int64), nv + mesh.cell_edges.astype(np.int64)], axis=1)
    owner = np.full(nv + mesh.num_edges, np.iinfo(np.int32).max, dtype=np.int32)
    np.minimum.at(owner, gn.ravel(), np.repeat(crank, gn.shape[1]))
    return gn, owner


def build_partition(mesh, rank, nranks, axis=None, crank=None):
    """Local mesh, numbering and (un-finalized) halo plans of one rank.  Deterministic and
    communication-free: every rank derives it from the same global mesh."""
    assert isinstance(mesh, SimplexMesh) and 0 <= rank < nranks
    nvg = mesh.num_vertices
    crank = cell_ranks(mesh, nranks, axis) if crank is None else np.asarray(crank, dtype=np.int32)
    gn, owner = global_node_owner(mesh, crank)
    cmask = (owner[gn] == rank).any(axis=1)
    cell_gid = np.nonzero(cmask)[0]
    if len(cell_gid) == 0:
        raise ValueError("rank %d owns nothing" % rank)

    def owned_first(gids):
        own = owner[gids]
        key = np.where(own == rank, -1, own)
        return gids[np.lexsort((gids, key))]

    lverts = owned_first(np.unique(mesh.cells[cell_gid]).astype(np.int64))
    g2l = np.full(nvg, -1, dtype=np.int64)
    g2l[lverts] = np.arange(len(lverts))
    local = SimplexMesh(mesh.coords[lverts], g2l[mesh.cells[cell_gid]])
    # boundary markers: exterior facets of the global mesh keep theirs, cut faces get 0
    def facet_keys(f):
        f = np.sort(f.astype(np.int64), axis=1)
        k = f[:, 0]
        for j in range(1, f.shape[1]):
            k = k * nvg + f[:, j]
        return k
    gk = facet_keys(mesh.bfacets)
    gorder = np.argsort(gk)
    lk = facet_keys(lverts[local.bfacets])
    pos = np.clip(np.searchsorted(gk[gorder], lk), 0, len(gk) - 1)
    hit = gk[gorder][pos] == lk
    local.bfacet_marker[:] = np.where(hit, mesh.bfacet_marker[gorder][pos], 0)
    local.global_boundary_ids = mesh.boundary_ids()      # a rank may not touch every boundary
    # P2 nodes: natural local numbering (vertices, then local edges) -> global ids -> owned first
    gedge = mesh.edge_index(lverts[local.edges[:, 0]], lverts[local.edges[:, 1]])
    nat_gid = np.concatenate([lverts, nvg + gedge.astype(np.int64)])
    lnodes = owned_first(nat_gid)
    node_perm = np.empty(len(nat_gid), dtype=np.int32)
    pos = np.argsort(lnodes, kind="stable")
    node_perm[:] = pos[np.searchsorted(lnodes[pos], nat_gid)]

    def plan(gids, n_global):
        own = owner[gids]
        n_owned = int((own == rank).sum())
        assert (own[:n_owned] == rank).all()
        return HaloPlan(n_owned, len(gids) - n_owned, n_global, gids[:n_owned], gids[n_owned:], own[n_owned:])

    nodes = plan(lnodes, nvg + mesh.num_edges)
    verts = plan(lverts, nvg)
    part = Partition(rank, nranks, local, node_perm, nodes, verts, lverts, lnodes, cell_gid)
    part.global_mesh = mesh          # host reference only: the replicated pressure solve works on the whole mesh
    return part


def connect(parts):
    """finalize the halo plans of all ranks held in one process (tests, single-process tools)"""
    for name in ("nodes", "verts"):
        req = [getattr(p, name).requests() for p in parts]
        for p in parts:
            getattr(p, name).finalize(p.rank, {q.rank: req[q.rank].get(p.rank, np.zeros(0, np.int64))
                                               for q in parts if q.rank != p.rank})
    return parts


def connect_distributed(part, group=None):
    """finalize the halo plans through torch.distributed (all ranks call it): every rank
    publishes, per owner, the global ids of its ghosts."""
    import torch.distributed as dist
    for name in ("nodes", "verts"):
        pl = getattr(part, name)
        gathered = [None] * part.nranks
        dist.all_gather_object(gathered, pl.requests(), group=group)
        pl.finalize(part.rank, {q: gathered[q].get(part.rank, np.zeros(0, np.int64))
                                for q in range(part.nranks) if q != part.rank})
    return part
