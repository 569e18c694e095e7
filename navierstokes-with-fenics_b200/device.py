"""Device-side state of one mesh: PyTorch tensors as device buffers (plumbing only), the
nsb_ctx of libnsb200.so, patterns and the thin typed wrappers around the C-ABI calls."""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib as L
from .spaces import TaylorHoodSpaces


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class Pattern:
    """CSR pattern + row->(cell, local row, slots) incidence, resident on the device."""

    def __init__(self, dev, nrows, cell_rows, cell_cols, nrows_owned=None):
        """``nrows_owned`` (partitioned meshes): keep only the owned prefix of the rows; the
        columns still span every local index (owned + ghost)."""
        lib = dev.lib
        cell_rows = np.ascontiguousarray(cell_rows, dtype=np.int32)
        cell_cols = np.ascontiguousarray(cell_cols, dtype=np.int32)
        ncells, nlr = cell_rows.shape
        nlc = cell_cols.shape[1]
        h = C.c_void_p()
        L.check(lib.nsb_pattern_build(nrows, ncells, nlr, cell_rows.ctypes.data, nlc, cell_cols.ctypes.data,
                                      C.byref(h)))
        try:
            nnz, npairs, maxrow = C.c_int64(), C.c_int64(), C.c_int32()
            L.check(lib.nsb_pattern_sizes(h, C.byref(nnz), C.byref(npairs), C.byref(maxrow)))
            self.nrows, self.nnz, self.npairs, self.max_row = nrows, nnz.value, npairs.value, maxrow.value
            self.nloc_col = nlc
            rowptr = np.empty(nrows + 1, dtype=np.int32)
            col = np.empty(self.nnz, dtype=np.int32)
            pair_ptr = np.empty(nrows + 1, dtype=np.int32)
            pair_cell = np.empty(self.npairs, dtype=np.int32)
            pair_iloc = np.empty(self.npairs, dtype=np.uint8)
            pair_slots = np.empty(self.npairs * nlc, dtype=np.uint8)
            L.check(lib.nsb_pattern_fill(h, rowptr.ctypes.data, col.ctypes.data, pair_ptr.ctypes.data,
                                         pair_cell.ctypes.data, pair_iloc.ctypes.data, pair_slots.ctypes.data))
        finally:
            lib.nsb_pattern_free(h)
        if nrows_owned is not None and nrows_owned < nrows:
            nrows = int(nrows_owned)
            rowptr, pair_ptr = rowptr[:nrows + 1].copy(), pair_ptr[:nrows + 1].copy()
            col = col[:rowptr[nrows]].copy()
            pair_cell, pair_iloc = pair_cell[:pair_ptr[nrows]].copy(), pair_iloc[:pair_ptr[nrows]].copy()
            pair_slots = pair_slots[:int(pair_ptr[nrows]) * nlc].copy()
            self.nrows, self.nnz, self.npairs = nrows, int(rowptr[nrows]), int(pair_ptr[nrows])
        self.rowptr_host, self.col_host = rowptr, col
        self._dev, self._tiles = dev, {}
        self.sched = None
        if nlc <= 15 and nlr == nlc:
            # conflict-free scatter schedule of the warp-cooperative matrix assembly
            cnt = (C.c_int64 * 4)()
            a_ = (nrows, rowptr.ctypes.data, pair_ptr.ctypes.data, pair_slots.ctypes.data, nlc, L.SCHED_MAX_BLOCKS)
            if lib.nsb_sched_build(*a_, cnt, None, None, None, None, None, None) == 0:
                g_row, c_first = np.empty(cnt[0] + 1, np.int32), np.empty(cnt[0] + 1, np.int32)
                r_ptr, sched = np.empty(cnt[1] + 1, np.int32), np.empty(32 * cnt[2], np.uint16)
                f_ptr, fold = np.empty(cnt[0] + 1, np.int32), np.empty(max(int(cnt[3]), 1), np.uint32)
                L.check(lib.nsb_sched_build(*a_, cnt, g_row.ctypes.data, c_first.ctypes.data, r_ptr.ctypes.data,
                                            sched.ctypes.data, f_ptr.ctypes.data, fold.ctypes.data))
                # unsigned words travel as signed bit patterns (torch has no unsigned arithmetic, none is needed)
                self.sched = (int(cnt[0]), dev.upload(g_row), dev.upload(c_first), dev.upload(r_ptr),
                              dev.upload(sched.view(np.int16)), dev.upload(f_ptr), dev.upload(fold.view(np.int32)))
        up = dev.upload
        self.rowptr, self.col = up(rowptr), up(col)
        self.pair_ptr, self.pair_cell = up(pair_ptr), up(pair_cell)
        self.pair_iloc, self.pair_slots = up(pair_iloc), up(pair_slots)


    def tiles(self, max_entries):
        """TMA row tiles (nsb_tiles_build) for tiles of <= max_entries entries, cached;
        (device tensor, ntiles) or (None, 0) when the pattern has no aligned partition."""
        if max_entries not in self._tiles:
            lib = self._dev.lib
            nt = C.c_int32()
            if lib.nsb_tiles_build(self.nrows, self.rowptr_host.ctypes.data, max_entries, C.byref(nt), None) != 0:
                self._tiles[max_entries] = (None, 0)
            else:
                rec = np.empty(nt.value * 64, dtype=np.uint32)
                L.check(lib.nsb_tiles_build(self.nrows, self.rowptr_host.ctypes.data, max_entries, C.byref(nt),
                                            rec.ctypes.data))
                self._tiles[max_entries] = (self._dev.upload(rec), nt.value)
        return self._tiles[max_entries]


class Device:
    """One GPU, one mesh.  Raises if CUDA or libnsb200.so is unavailable (no CPU fallback)."""

    def __init__(self, mesh, device_index=0, mixed=False, partition=None, node_perm=None, patterns=None):
        """``partition`` (partition.Partition with finalized halo plans): this rank's share of a
        mesh-partitioned run; ``mesh`` is then ignored in favour of the partition's local mesh and
        torch.distributed must be initialised (it carries the NCCL id to the other ranks).
        ``node_perm`` (ingest mode): natural P2 node number (vertex v / N0 + e) -> the node
        number to use on the device, e.g. the one a real dolfin's re-ordered dofmap assigns; every map,
        pattern and kernel then works in that numbering.
        ``patterns``: names of the patterns to build up front (default: "WW" if mixed else "P2P2", "P1P1"); a
        device that only serves the pressure space (the replicated pressure solve of a partitioned run) passes
        ("P1P1",) and gets no cell scratch."""
        if not torch.cuda.is_available():
            raise RuntimeError("nsb200 needs a CUDA device: there is no CPU fallback")
        self.lib = L.load()
        self.torch_device = torch.device("cuda", device_index)
        torch.cuda.set_device(self.torch_device)
        self.ctx = C.c_void_p()
        L.check(self.lib.nsb_create(C.byref(self.ctx), device_index))
        self.partition = partition
        if partition is not None:
            assert not mixed, "the monolithic solvers run on one GPU (SURVEY.md 8e: replicas only)"
            mesh = partition.mesh
        self.mesh = mesh
        assert partition is None or node_perm is None
        self.spaces = sp = TaylorHoodSpaces(mesh, node_perm if partition is None else partition.node_perm)
        self.dim = sp.dim
        self.n_owned_nodes = sp.n_nodes if partition is None else partition.nodes.n_owned
        self.n_owned_verts = sp.n_vertices if partition is None else partition.verts.n_owned
        if partition is not None:
            self._setup_comm(partition)
        self.coords = self.upload(mesh.coords)
        self.cell_verts = self.upload(mesh.cells)
        self.cell_nodes = self.upload(sp.cell_nodes)
        self.check(self.lib.nsb_set_mesh(self.ctx, sp.dim, mesh.num_vertices, _ptr(self.coords), mesh.num_cells,
                                         _ptr(self.cell_verts), sp.n_nodes, _ptr(self.cell_nodes)))
        if patterns is None:
            patterns = ("WW",) if mixed else ("P2P2", "P1P1")
        if "P2P2" in patterns:
            self.scratch = self.empty(mesh.num_cells * self.lib.nsb_scratch_doubles_per_cell(sp.dim))
            self.check(self.lib.nsb_set_scratch(self.ctx, _ptr(self.scratch), self.scratch.numel()))
        self.patterns = {}
        self.coef = None
        self._launch_hint = 0
        for name in patterns:
            self.pattern(name)

    def __del__(self):
        try:
            if getattr(self, "ctx", None):
                self.lib.nsb_destroy(self.ctx)
                self.ctx = None
        except Exception:
            pass

    # ---- multi-GPU ---------------------------------------------------------------------------
    def _setup_comm(self, part):
        """NCCL communicator (id broadcast through torch.distributed) and the two halo plans."""
        import torch.distributed as dist
        ident = torch.zeros(128, dtype=torch.uint8)
        if part.rank == 0:
            buf = (C.c_uint8 * 128)()
            L.check(self.lib.nsb_comm_unique_id(buf))
            ident = torch.tensor(list(buf), dtype=torch.uint8)
        if part.nranks > 1:
            if dist.get_backend() == "nccl":          # pure-NCCL process group: device tensors only
                ident = ident.to(self.torch_device)
            dist.broadcast(ident, src=0)
        buf = (C.c_uint8 * 128)(*ident.cpu().tolist())
        self.check(self.lib.nsb_comm_init(self.ctx, part.rank, part.nranks, buf))
        self._halo_keep = []
        for which, pl in ((L.HALO_NODES, part.nodes), (L.HALO_VERTS, part.verts)):
            assert pl.peers is not None, "halo plan not finalized (partition.connect*)"
            send_idx = self.upload(pl.send_idx) if len(pl.send_idx) else None
            sendbuf = self.empty(3 * max(len(pl.send_idx), 1))
            ghostbuf = self.empty(3 * max(pl.n_ghost, 1))
            peers, sp_, rp_ = (np.ascontiguousarray(a, dtype=np.int32) for a in (pl.peers, pl.send_ptr, pl.recv_ptr))
            self.check(self.lib.nsb_set_halo(self.ctx, which, pl.n_owned, pl.n_ghost, pl.n_global, len(peers),
                                             peers.ctypes.data, sp_.ctypes.data, _ptr(send_idx), rp_.ctypes.data,
                                             _ptr(sendbuf), _ptr(ghostbuf)))
            self._halo_keep.append((send_idx, sendbuf, ghostbuf))
        self.halo_transport = "nccl"
        if part.nranks > 1 and os.environ.get("NSB_HALO", "p2p") != "nccl":
            self._connect_peer_memory(part)

    def _connect_peer_memory(self, part):
        """Halo exchange over NVLink peer memory (nsb_p2p_*): every rank exports its IPC block, the
        handles and ghost layouts travel through torch.distributed, every rank maps its neighbours.
        All-or-nothing: if any rank cannot map a neighbour, every rank stays on NCCL send/recv."""
        import torch.distributed as dist
        mine, ok = {}, True
        try:
            for which, pl in ((L.HALO_NODES, part.nodes), (L.HALO_VERTS, part.verts)):
                buf = (C.c_uint8 * 64)()
                self.check(self.lib.nsb_p2p_export(self.ctx, which, buf))
                mine[which] = dict(handle=bytes(buf), **pl.peer_memory_layout())
            if part.nranks <= 8:                          # all-reduce block of the Krylov scalars
                buf = (C.c_uint8 * 64)()
                self.check(self.lib.nsb_p2p_export(self.ctx, L.HALO_NONE, buf))
                mine[L.HALO_NONE] = dict(handle=bytes(buf))
        except RuntimeError:
            ok = False
        everyone = [None] * part.nranks
        dist.all_gather_object(everyone, (ok, mine))
        ok = all(e[0] for e in everyone)
        if ok:
            try:
                for which, pl in ((L.HALO_NODES, part.nodes), (L.HALO_VERTS, part.verts)):
                    ng, of, sl = pl.peer_memory_tables(part.rank, [e[1].get(which) for e in everyone])
                    handles = b"".join(everyone[int(r)][1][which]["handle"] for r in pl.peers)
                    hb = (C.c_uint8 * max(len(handles), 1)).from_buffer_copy(handles or b"\0")
                    self.check(self.lib.nsb_p2p_connect(self.ctx, which, len(pl.peers), hb, ng.ctypes.data,
                                                        of.ctypes.data, sl.ctypes.data))
                if part.nranks <= 8:
                    handles = b"".join(e[1][L.HALO_NONE]["handle"] for e in everyone)
                    hb = (C.c_uint8 * len(handles)).from_buffer_copy(handles)
                    self.check(self.lib.nsb_p2p_connect(self.ctx, L.HALO_NONE, part.nranks, hb, None, None, None))
            except (RuntimeError, ValueError, KeyError):
                ok = False
        verdict = [None] * part.nranks
        dist.all_gather_object(verdict, ok)              # also the barrier between connect and first use
        if all(verdict):
            self.halo_transport = "p2p"
        else:
            self.check(self.lib.nsb_p2p_disconnect(self.ctx))
            dist.barrier()

    def halo_status(self):
        """raise if a bounded wait of the peer-memory halo exchange ever timed out"""
        if self.partition is None or self.halo_transport != "p2p":
            return
        err = C.c_int32(0)
        self.check(self.lib.nsb_p2p_status(self.ctx, C.byref(err)))
        if err.value:
            raise RuntimeError("peer-memory halo exchange timed out (code %d: %s)" % (
                err.value, {1: "a neighbour stopped consuming", 2: "a neighbour stopped sending"}.get(err.value, "?")))

    def close(self):
        """collective on a partitioned run (every rank calls it): tears the NCCL communicator
        down, then frees the context"""
        if getattr(self, "ctx", None):
            if self.partition is not None:
                if self.partition.nranks > 1:
                    import torch.distributed as dist
                    torch.cuda.synchronize(self.torch_device)
                    try:
                        self.halo_status()
                        failure = None
                    except RuntimeError as e:
                        failure = e
                    dist.barrier()                       # nobody pushes into a block that is being freed
                self.check(self.lib.nsb_comm_finalize(self.ctx))
            self.lib.nsb_destroy(self.ctx)
            self.ctx = None
            if self.partition is not None and self.partition.nranks > 1 and failure is not None:
                raise failure

    def halo_exchange(self, which, bs, x):
        """refresh the ghost block of the [owned | ghost] vector ``x`` from the owners (no-op on
        one GPU)"""
        if self.partition is None or self.partition.nranks == 1:
            return
        n_owned = self.n_owned_nodes if which == L.HALO_NODES else self.n_owned_verts
        ghost = C.c_void_p(x.data_ptr() + 8 * bs * n_owned)
        self.check(self.lib.nsb_halo_exchange(self.ctx, which, bs, _ptr(x), ghost, self.stream))

    def allreduce(self, x):
        """x += the same vector of every other rank, in place (no-op on one GPU)"""
        if self.partition is None or self.partition.nranks == 1:
            return
        self.check(self.lib.nsb_allreduce(self.ctx, _ptr(x), x.numel(), self.stream))

    # ---- plumbing ---------------------------------------------------------------------------
    def check(self, status):
        L.check(status, self.ctx)

    def upload(self, array):
        return torch.from_numpy(np.ascontiguousarray(array)).to(self.torch_device)

    def host_buffer(self, n):
        """pinned host staging buffer (plain host memory when the device is not a CUDA device: CPU test doubles)"""
        t = torch.empty(int(n), dtype=torch.float64)
        return t.pin_memory() if self.torch_device.type == "cuda" else t

    def zeros(self, n, dtype=torch.float64):
        return torch.zeros(int(n), dtype=dtype, device=self.torch_device)

    def empty(self, n, dtype=torch.float64):
        return torch.empty(int(n), dtype=dtype, device=self.torch_device)

    @property
    def stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.torch_device).cuda_stream)

    def pattern(self, name):
        if name not in self.patterns:
            sp = self.spaces
            if name == "P2P2":
                pat = Pattern(self, sp.n_nodes, sp.cell_nodes, sp.cell_nodes, self.n_owned_nodes)
                pat.halo = L.HALO_NODES
                which = L.PAT_P2P2
            elif name == "P1P1":
                pat = Pattern(self, sp.n_vertices, self.mesh.cells, self.mesh.cells, self.n_owned_verts)
                pat.halo = L.HALO_VERTS
                which = L.PAT_P1P1
            elif name == "WW":
                cd = sp.cell_dofs_W
                pat = Pattern(self, sp.ndofs_W, cd, cd)
                which = L.PAT_WW
                self.cell_dofs_W = self.upload(cd)
                self.check(self.lib.nsb_set_mixed_dofmap(self.ctx, sp.ndofs_W, _ptr(self.cell_dofs_W)))
            else:
                raise KeyError(name)
            self.check(self.lib.nsb_set_pattern(self.ctx, which, pat.nrows, _ptr(pat.rowptr), _ptr(pat.col),
                                                _ptr(pat.pair_ptr), _ptr(pat.pair_cell), _ptr(pat.pair_iloc),
                                                _ptr(pat.pair_slots)))
            if pat.sched is not None:
                ng, g_row, c_first, r_ptr, sched, f_ptr, fold = pat.sched
                self.check(self.lib.nsb_set_schedule(self.ctx, which, ng, L.SCHED_MAX_BLOCKS, _ptr(g_row),
                                                     _ptr(c_first), _ptr(r_ptr), _ptr(sched), _ptr(f_ptr), _ptr(fold)))
            self.patterns[name] = pat
        return self.patterns[name]

    def matrix(self, pattern, vals, kind, bs, bc_mask=None, vals32=None, diag32=None, compress=False):
        """nsb_matrix descriptor (keeps the tensors alive through attributes).  ``vals32``: float32
        copy of the BSR values for the SpMVs inside the Chebyshev preconditioner (``to_f32``);
        ``diag32``: one float per block, its block-scalar approximation (``block_scalar_f32``)."""
        pat = self.patterns[pattern]
        tiles, ntiles = pat.tiles(L.TILE_ENTRIES[(kind, bs)]) if (kind, bs) in L.TILE_ENTRIES else (None, 0)
        m = L.Matrix(kind, bs, pat.nrows, pat.nnz, pat.rowptr.data_ptr(), pat.col.data_ptr(), vals.data_ptr(),
                     bc_mask.data_ptr() if bc_mask is not None else None,
                     tiles.data_ptr() if tiles is not None else None, ntiles,
                     getattr(pat, "halo", 0) if self.partition is not None else 0,
                     vals32.data_ptr() if vals32 is not None else None,
                     diag32.data_ptr() if diag32 is not None else None, None)
        m._keep = (pat, vals, bc_mask, tiles, vals32, diag32)
        if compress and kind == L.CSR and bs == 1:
            # SELL-32 copy for the persistent CG kernel (+ float values / 16-bit column deltas for its polynomial)
            rp = pat.rowptr_host.astype(np.int64)
            nsl = (pat.nrows + 31) // 32
            lens = np.zeros(nsl * 32, dtype=np.int64)
            lens[:pat.nrows] = rp[1:] - rp[:-1]
            width = lens.reshape(nsl, 32).max(axis=1)
            sp_host = np.zeros(nsl + 1, dtype=np.int64)
            np.cumsum(32 * width, out=sp_host[1:])
            if sp_host[-1] < 2 ** 31:
                tot = int(sp_host[-1])
                slice_ptr = self.upload(sp_host.astype(np.int32))
                scol, sval = self.empty(tot, dtype=torch.int32), self.empty(tot)
                v32, c16 = self.empty(tot, dtype=torch.float32), self.empty(tot, dtype=torch.int16)
                over = C.c_int32()
                self.check(self.lib.nsb_sell_fill(self.ctx, C.byref(m), nsl, _ptr(slice_ptr), _ptr(scol), _ptr(sval), _ptr(v32),
                                                  _ptr(c16), C.byref(over), self.stream))
                sell = L.Sell(nsl, 0, slice_ptr.data_ptr(), scol.data_ptr(), sval.data_ptr(), v32.data_ptr(),
                              c16.data_ptr() if over.value == 0 else None)
                m.sell = C.pointer(sell)
                m._keep = m._keep + (slice_ptr, scol, sval, v32, c16, sell)
        return m

    # ---- coefficients -----------------------------------------------------------------------
    def set_coefficients(self, k, alpha, c_conv=1.0, c_p=1.0, c_visc=1.0, c_body=1.0):
        arr = (C.c_double * 8)(k, alpha[0], alpha[1], alpha[2], c_conv, c_p, c_visc, c_body)
        self.check(self.lib.nsb_set_coefficients(self.ctx, arr))
        self.coef = tuple(arr)

    def set_forms(self, conv=0, visc=0, coriolis=None, generic=False):
        """weak-form variants (nsb_set_forms): conv 0 standard | 1 rotational | 2 divergence | 3 skew_symmetric,
        visc 0 reduced | 1 traction, coriolis = 2 c_Omega Omega (3 components; 2D: only the last)"""
        arr = None
        if coriolis is not None:
            arr = (C.c_double * 3)(*[float(v) for v in coriolis])
        self.check(self.lib.nsb_set_forms(self.ctx, int(conv), int(visc), arr, 1 if generic else 0))

    # ---- assembly ---------------------------------------------------------------------------
    def assemble_ipcs_momentum(self, flags, ustar, un=None, unm1=None, pn=None, b=None, A=None):
        self.check(self.lib.nsb_assemble_ipcs_momentum(self.ctx, flags, _ptr(ustar), _ptr(un), _ptr(unm1),
                                                       _ptr(pn), _ptr(b), _ptr(A), self.stream))

    def assemble_poisson(self, flags, pn=None, ustar=None, b=None, A=None):
        self.check(self.lib.nsb_assemble_poisson(self.ctx, flags, _ptr(pn), _ptr(ustar), _ptr(b), _ptr(A),
                                                 self.stream))

    def assemble_correction(self, flags, ustar=None, pnew=None, pold=None, b=None, A=None):
        self.check(self.lib.nsb_assemble_correction(self.ctx, flags, _ptr(ustar), _ptr(pnew), _ptr(pold),
                                                    _ptr(b), _ptr(A), self.stream))

    def assemble_p1_mass(self, flags, w=None, A=None):
        self.check(self.lib.nsb_assemble_p1_mass(self.ctx, flags, _ptr(w), _ptr(A), self.stream))

    def assemble_mixed(self, flags, w0, w1=None, w2=None, b=None, A=None):
        self.check(self.lib.nsb_assemble_mixed(self.ctx, flags, _ptr(w0), _ptr(w1), _ptr(w2), _ptr(b), _ptr(A),
                                               self.stream))

    def set_mixed_bc(self, mask, g):
        self._mixed_bc = (mask, g)
        self.check(self.lib.nsb_set_mixed_bc(self.ctx, _ptr(mask), _ptr(g)))

    # ---- Dirichlet rows ---------------------------------------------------------------------
    def bc_apply_matrix(self, pattern, vals, kind, bs, bc_dofs):
        if bc_dofs.numel() == 0:
            return
        pat = self.patterns[pattern]
        self.check(self.lib.nsb_bc_apply_matrix(self.ctx, kind, bs, _ptr(pat.rowptr), _ptr(pat.col), _ptr(vals),
                                                bc_dofs.numel(), _ptr(bc_dofs), self.stream))

    def bc_apply_vector(self, mode, b, bc_dofs, bc_vals, x=None):
        if bc_dofs.numel() == 0:
            return
        self.check(self.lib.nsb_bc_apply_vector(self.ctx, mode, _ptr(b), _ptr(x), bc_dofs.numel(), _ptr(bc_dofs),
                                                _ptr(bc_vals), self.stream))

    # ---- linear algebra ---------------------------------------------------------------------
    def spmv(self, A, x, y):
        self.check(self.lib.nsb_spmv(self.ctx, C.byref(A), _ptr(x), _ptr(y), self.stream))

    def inv_diagonal(self, A, out):
        self.check(self.lib.nsb_inv_diagonal(self.ctx, C.byref(A), _ptr(out), self.stream))

    def precond(self, inv_diag, kind=L.PC_JACOBI, degree=1, lambda_min=0.0, lambda_max=0.0, lowp=False):
        """nsb_precond descriptor; a plain tensor passed to cg/gmres means Jacobi.  ``lowp``: the Chebyshev
        polynomial runs in single precision inside a flexible GMRES (needs the matrix's vals32 copy)."""
        m = L.Precond(kind, degree, inv_diag.data_ptr() if inv_diag is not None else None, lambda_min, lambda_max,
                      1 if lowp else 0, 0, None)
        m._keep = inv_diag
        return m

    def _pc(self, M):
        return M if isinstance(M, L.Precond) else self.precond(M)

    def lambda_max(self, A, inv_diag, iters=12, work=None, warm_start=False, lowp=False):
        """``warm_start``: ``work`` (kept by the caller between calls) still holds the previous iterate;
        ``lowp``: power iteration on the single-precision block-scalar copy (nsb_lambda_max32)"""
        n = A.nrows * A.bs
        if work is None or work.numel() < 2 * n:
            assert not warm_start, "a warm start needs the work vector of the previous call"
            work = self.empty(2 * n)
        out = C.c_double()
        fn = self.lib.nsb_lambda_max32 if lowp else self.lib.nsb_lambda_max
        self.check(fn(self.ctx, C.byref(A), _ptr(inv_diag), iters, 1 if warm_start else 0, _ptr(work), C.byref(out),
                      self.stream))
        return out.value

    def vorticity(self, u):
        """``dolfin.project(curl(u), DG_1)`` (source/ns_problem.py:55-81): [ncells, d + 1] in 2D, [ncells, 4, 3] in 3D --
        the DG-1 dofs of a cell are its vertex values, in the cell's local vertex order"""
        d, nc = self.dim, self.mesh.num_cells
        out = self.empty(nc * (d + 1) * (1 if d == 2 else 3))
        self.check(self.lib.nsb_cell_vorticity(self.ctx, _ptr(u), _ptr(out), self.stream))
        return out.view(nc, d + 1) if d == 2 else out.view(nc, d + 1, 3)

    def pressure_gradient(self, p):
        """``dolfin.project(grad(p), [DG_0]^d)`` (source/ns_problem.py:83-101): [ncells, d]"""
        out = self.empty(self.mesh.num_cells * self.dim)
        self.check(self.lib.nsb_cell_pressure_gradient(self.ctx, _ptr(p), _ptr(out), self.stream))
        return out.view(self.mesh.num_cells, self.dim)

    def rhs_basis(self, n, lmax=10):
        """state of the successive-right-hand-side projection (nsb_rhs_guess / nsb_rhs_update) for a system of size n"""
        return dict(n=n, lmax=lmax, l=C.c_int32(0), X=self.empty(lmax * n), AX=self.empty(lmax * n), work=self.empty(2 * n))

    def rhs_guess(self, basis, b, x):
        self.check(self.lib.nsb_rhs_guess(self.ctx, basis["n"], basis["l"].value, _ptr(basis["X"]), _ptr(b), _ptr(x),
                                          self.stream))
        return basis["l"].value

    def rhs_update(self, basis, A, x):
        self.check(self.lib.nsb_rhs_update(self.ctx, C.byref(A), C.byref(basis["l"]), basis["lmax"], _ptr(basis["X"]),
                                           _ptr(basis["AX"]), _ptr(x), _ptr(basis["work"]), self.stream))
        return basis["l"].value

    def saddle_lambda_max(self, A, inv_diag, n_u, iters=12, work=None):
        """largest eigenvalue of D^-1 F, F = velocity block of the mixed matrix ``A``"""
        n = A.nrows
        if work is None or work.numel() < 2 * n:
            work = self.empty(2 * n)
        out = C.c_double()
        self.check(self.lib.nsb_saddle_lambda_max(self.ctx, C.byref(A), _ptr(inv_diag), n_u, iters, _ptr(work), C.byref(out),
                                                  self.stream))
        return out.value

    def cg(self, A, M, b, x, rtol=1e-12, atol=0.0, maxit=10000, nullspace=False, work=None):
        n = A.nrows * A.bs
        need = self.lib.nsb_cg_workspace_doubles(n)
        if work is None or work.numel() < need:
            work = self.empty(need)
        it, res = C.c_int32(), C.c_double()
        M = self._pc(M)
        st = self.lib.nsb_cg(self.ctx, C.byref(A), C.byref(M), _ptr(b), _ptr(x), rtol, atol, maxit,
                             1 if nullspace else 0, _ptr(work), C.byref(it), C.byref(res), self.stream)
        self.check(st)
        return it.value, res.value

    def gmres(self, A, M, b, x, rtol=1e-12, atol=0.0, maxit=2000, restart=40, work=None, x0_is_zero=False):
        n = A.nrows * A.bs
        M = self._pc(M)
        need = self.lib.nsb_gmres_workspace_doubles(n, restart, 1 if (M.lowp or M.kind == L.PC_SADDLE) else 0)
        if work is None or work.numel() < need:
            work = self.empty(need)
        it, res = C.c_int32(), C.c_double()
        st = self.lib.nsb_gmres(self.ctx, C.byref(A), C.byref(M), _ptr(b), _ptr(x), rtol, atol, maxit, restart,
                                1 if x0_is_zero else 0, _ptr(work), C.byref(it), C.byref(res), self.stream)
        self.check(st)
        return it.value, res.value

    def facet_table(self, boundary_id):
        """(owning cell, local number of the opposite vertex) of every exterior facet marked
        ``boundary_id``, on the device (cached)"""
        key = int(boundary_id)
        if not hasattr(self, "_facet_tables"):
            self._facet_tables = {}
        if key not in self._facet_tables:
            m = self.mesh
            sel = np.nonzero(m.bfacet_marker == key)[0]
            cells = m.bfacet_cell[sel].astype(np.int32)
            cv, fv = m.cells[cells], m.bfacets[sel]
            inside = (cv[:, :, None] == fv[:, None, :]).any(axis=2)
            opp = np.argmin(inside, axis=1).astype(np.uint8)
            assert len(sel) == 0 or (inside.sum(axis=1) == m.dim).all()
            self._facet_tables[key] = (self.upload(cells), self.upload(opp), len(sel))
        return self._facet_tables[key]

    def boundary_force(self, table, u, p, c_visc):
        cells, opp, n = table
        out = (C.c_double * 3)()
        if n == 0 and self.partition is None:
            return np.zeros(self.dim)
        self.check(self.lib.nsb_boundary_force(self.ctx, n, _ptr(cells), _ptr(opp), _ptr(u), _ptr(p), float(c_visc), out,
                                               self.stream))
        return np.array(out[:self.dim])

    def direct_solve(self, pattern, vals, b, x, reorder=1):
        """K16 (library leg): x = A^-1 b by cuSOLVER sparse QR; returns the seconds it took."""
        pat = self.patterns[pattern]
        sec = C.c_double()
        self.check(self.lib.nsb_direct_solve(self.ctx, pat.nrows, pat.nnz, _ptr(pat.rowptr), _ptr(pat.col), _ptr(vals),
                                             _ptr(b), _ptr(x), reorder, C.byref(sec), self.stream))
        return sec.value

    def direct_seconds(self, reset=False):
        sec, n = C.c_double(), C.c_int64()
        self.check(self.lib.nsb_direct_seconds(self.ctx, C.byref(sec), C.byref(n), 1 if reset else 0))
        return sec.value, n.value

    # ---- L2 projection, error norms, CFL (csrc/project.cu) ----------------------
    def _cell_lattice_values(self, fn, k, ncomp):
        """values of ``fn(x[n,d]) -> [n,ncomp]`` at the P_k lattice nodes of every cell (dolfin's per-cell
        interpolation of an Expression of degree k), device tensor [ncells, nk, ncomp]"""
        from . import lagrange
        m, d = self.mesh, self.dim
        lam = lagrange._bary(lagrange.lattice_points(d, k))                # [nk, d+1]
        X = np.einsum("km,cmi->cki", lam, m.coords[m.cells])               # [ncells, nk, d]
        F = np.asarray(fn(X.reshape(-1, d)), dtype=np.float64).reshape(m.num_cells, lam.shape[0], ncomp)
        return self.upload(F), lam.shape[0]

    def mass_operator(self, space):
        """(nsb_matrix, inverse diagonal) of the P2 (x I_d) / P1 mass matrix without Dirichlet rows, cached"""
        if not hasattr(self, "_mass_ops"):
            self._mass_ops = {}
        if space not in self._mass_ops:
            if space == "V":
                pat = self.pattern("P2P2")
                A = self.empty(pat.nnz)
                self.assemble_correction(L.MAT, A=A)
                M = self.matrix("P2P2", A, L.CSR, self.dim)
            else:
                pat = self.pattern("P1P1")
                A = self.empty(pat.nnz)
                self.assemble_p1_mass(L.MAT, A=A)
                M = self.matrix("P1P1", A, L.CSR, 1)
            dinv = self.empty(M.nrows * M.bs)
            self.inv_diagonal(M, dinv)
            self._mass_ops[space] = (M, dinv)
        return self._mass_ops[space]

    def cell_rhs(self, fn, space, degree):
        """b_I = int (I_k fn) phi_I dx over the owned rows of V (interleaved) or Q: right-hand side of a projection,
        body-force load vector (source/ns_solver_base.py:158-171)"""
        from . import lagrange
        d = self.dim
        ncomp = d if space == "V" else 1
        which, pname = (L.PAT_P2P2, "P2P2") if space == "V" else (L.PAT_P1P1, "P1P1")
        pat = self.pattern(pname)
        F, nk = self._cell_lattice_values(fn, degree, ncomp)
        R = self.upload(lagrange.cross_mass(d, "P2" if space == "V" else "P1", degree))
        b = self.empty(pat.nrows * ncomp)
        self.check(self.lib.nsb_assemble_cell_rhs(self.ctx, which, nk, ncomp, _ptr(R), _ptr(F), _ptr(b), self.stream))
        return b

    def project(self, fn, space, degree, out=None, rtol=1e-14):
        """``dolfin.project(Expression(degree=degree), V|Q)`` (source/ns_solver_base.py:1151,1168): cell-wise P_k
        interpolant of ``fn``, exact right-hand side on the device, mass solve by Jacobi-CG (LU in the
        reference).  Returns the [owned | ghost] dof vector (V interleaved / Q)."""
        from . import lagrange
        assert space in ("V", "Q") and degree >= 1
        d = self.dim
        ncomp = d if space == "V" else 1
        which, pname, halo = (L.PAT_P2P2, "P2P2", L.HALO_NODES) if space == "V" else (L.PAT_P1P1, "P1P1", L.HALO_VERTS)
        b = self.cell_rhs(fn, space, degree)
        M, dinv = self.mass_operator(space)
        n_own = M.nrows * M.bs
        n_all = (self.spaces.n_nodes * d) if space == "V" else self.spaces.n_vertices
        x = self.zeros(n_all) if out is None else out
        x[:n_own].zero_()
        self.cg(M, dinv, b, x, rtol=rtol)
        self.halo_exchange(halo, ncomp, x)
        return x

    def errornorm(self, exact_fn, uh, space, degree_rise=3):
        """``dolfin.errornorm(u_exact, u_h)`` (L2, default degree_rise = 3; App. A.7): both interpolated into
        P_{k+3} on every cell, || difference ||_L2"""
        from . import lagrange
        assert self.partition is None, "errornorm runs on one GPU"
        d = self.dim
        ncomp, k0, which = (d, 2, L.PAT_P2P2) if space == "V" else (1, 1, L.PAT_P1P1)
        k = k0 + degree_rise
        F, nk = self._cell_lattice_values(exact_fn, k, ncomp)
        pts = lagrange.lattice_points(d, k)
        E = lagrange.eval_p2(d, pts) if space == "V" else lagrange.eval_p1(d, pts)
        Xq, Wq = lagrange.simplex_rule(d, 2 * k)
        Psi = lagrange.eval_lagrange(d, k, Xq)
        work = self.empty(F.numel())
        out = C.c_double()
        tabs = (self.upload(E), self.upload(Psi), self.upload(Wq))       # named: they must outlive the launch
        self.check(self.lib.nsb_cell_error_norm2(self.ctx, which, ncomp, _ptr(uh), nk, _ptr(tabs[0]), len(Wq),
                                                 _ptr(tabs[1]), _ptr(tabs[2]), _ptr(F), _ptr(work),
                                                 C.byref(out), self.stream))
        return float(np.sqrt(max(out.value, 0.0)))

    def cfl(self, u, step_size, degree=2):
        """max local CFL number of source/ns_problem.py:554-587 for the P2 velocity ``u`` (V numbering)"""
        from . import lagrange
        if not hasattr(self, "_cfl_tabs"):
            B, P = lagrange.cfl_tables(self.dim)
            self._cfl_tabs = (self.upload(B), self.upload(P), B.shape[0])
        B, P, nq = self._cfl_tabs
        out = C.c_double()
        self.check(self.lib.nsb_cfl_max(self.ctx, _ptr(u), float(degree) * float(step_size), nq, _ptr(B), _ptr(P),
                                        C.byref(out), self.stream))
        return out.value

    def shift(self, x, s):
        """x -= s"""
        self.check(self.lib.nsb_vec_shift(self.ctx, x.numel(), float(s), _ptr(x), self.stream))

    def p1_integral(self, p):
        """(int p dx, int 1 dx) for a P1 function in Q numbering (source/ns_solver_base.py:1194)"""
        n_own = self.n_owned_verts
        if not hasattr(self, "_p1_weights"):
            self.pattern("P1P1")
            self._p1_weights = self.empty(n_own)
            self.assemble_p1_mass(L.VEC, w=self._p1_weights)
            self._volume = self.sum(self._p1_weights)
        return self.dot(self._p1_weights, p[:n_own]), self._volume

    def set_block_scalar_out(self, out):
        """register (None: unregister) the float tensor the momentum-Jacobian assembly fills with the block-scalar copy"""
        self._block_scalar_out = out
        self.check(self.lib.nsb_set_block_scalar_out(self.ctx, _ptr(out), 0 if out is None else out.numel()))

    def block_scalar_rows_f32(self, vals, bs, block_rows, out):
        """refresh the block-scalar copy on the listed block rows (after bc_apply_matrix)"""
        pat = self.patterns["P2P2"]
        self.check(self.lib.nsb_bsr_block_scalar_rows_f32(self.ctx, block_rows.numel(), _ptr(block_rows), _ptr(pat.rowptr), bs,
                                                          _ptr(vals), _ptr(out), self.stream))

    def block_scalar_f32(self, vals, bs, out):
        """out[k] = float(mean of the diagonal of the bs x bs block k of ``vals``)"""
        self.check(self.lib.nsb_bsr_block_scalar_f32(self.ctx, out.numel(), bs, _ptr(vals), _ptr(out), self.stream))

    def to_f32(self, x, y):
        self.check(self.lib.nsb_vec_to_f32(self.ctx, x.numel(), _ptr(x), _ptr(y), self.stream))

    def copy(self, x, y):
        self.check(self.lib.nsb_vec_copy(self.ctx, x.numel(), _ptr(x), _ptr(y), self.stream))

    def axpy(self, a, x, y):
        self.check(self.lib.nsb_vec_axpy(self.ctx, x.numel(), float(a), _ptr(x), _ptr(y), self.stream))

    def scale(self, a, x):
        self.check(self.lib.nsb_vec_scale(self.ctx, x.numel(), float(a), _ptr(x), self.stream))

    def dot(self, x, y):
        out = C.c_double()
        self.check(self.lib.nsb_vec_dot(self.ctx, x.numel(), _ptr(x), _ptr(y), C.byref(out), self.stream))
        return out.value

    def norm(self, x):
        return float(np.sqrt(self.dot(x, x)))

    def sum(self, x):
        out = C.c_double()
        self.check(self.lib.nsb_vec_sum(self.ctx, x.numel(), _ptr(x), C.byref(out), self.stream))
        return out.value

    def scatter(self, idx, x, y):
        self.check(self.lib.nsb_vec_scatter(self.ctx, idx.numel(), _ptr(idx), _ptr(x), _ptr(y), self.stream))

    def gather(self, idx, x, y):
        self.check(self.lib.nsb_vec_gather(self.ctx, idx.numel(), _ptr(idx), _ptr(x), _ptr(y), self.stream))

    def ipcs_advance(self, u0, u1, u2, p, p_old):
        self.check(self.lib.nsb_ipcs_advance(self.ctx, u0.numel(), _ptr(u0), _ptr(u1), _ptr(u2), p.numel(), _ptr(p),
                                             _ptr(p_old), self.stream))

    @property
    def launches(self):
        return self.launch_count()

    # ---- instrumentation --------------------------------------------------------------------
    def timer_enable(self, on=True, classes=None):
        """``classes``: names out of ``_lib.TIMER_CLASSES`` to time (default: all)"""
        code = 1 if on else 0
        if on and classes is not None:
            code = sum(1 << L.TIMER_CLASSES.index(c) for c in classes) << 1
        self.check(self.lib.nsb_timer_enable(self.ctx, code))

    def timer_report(self):
        """{class: (milliseconds, timed regions)} since the last report (synchronises)."""
        n = len(L.TIMER_CLASSES)
        ms, cnt = (C.c_double * n)(), (C.c_int64 * n)()
        self.check(self.lib.nsb_timer_report(self.ctx, ms, cnt))
        return {name: (ms[i], cnt[i]) for i, name in enumerate(L.TIMER_CLASSES)}

    def launch_count(self):
        """kernels launched by libnsb200 for this context (counted inside the library)."""
        out = C.c_int64()
        self.check(self.lib.nsb_launch_count(self.ctx, C.byref(out)))
        return out.value

    def fp64_peak(self):
        out = C.c_double()
        self.check(self.lib.nsb_fp64_peak(self.ctx, C.byref(out)))
        return out.value
