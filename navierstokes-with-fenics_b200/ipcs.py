"""IPCSSolver: mirror of ``source/ns_ipcs_solver.py`` whose per-step work (assembly of the
three sub-steps, Dirichlet rows, the linear solves, the history shift) runs in libnsb200.so.

Same constructor arguments, ``set_boundary_conditions`` / ``set_equation_coefficients`` /
``set_initial_conditions`` / ``solve`` / ``advance_time`` / ``solution`` as the reference class.
Differences forced by replacing LU: Krylov tolerances ``linear_rtol`` and the
zero-mean representative for the all-Neumann pressure problem (H4).
"""
import numpy as np
import torch

from . import _lib as L
from .device import Device
from .solver_base import SolverBase, Expression, NewtonNotConverged, newton_converged


def initial_fields(dev, initial_conditions):
    """(velocity in V numbering, pressure in Q numbering or None) of an initial-condition dictionary, on the
    device; shared by the IPCS and the monolithic mirrors (source/ns_solver_base.py:1123-1171)"""
    sp = dev.spaces
    v = initial_conditions["velocity"]
    if isinstance(v, Expression):
        u0 = dev.project(v, "V", v.degree)
    else:
        x = sp.node_coordinates()
        if callable(v):
            u0 = np.asarray(v(x), dtype=np.float64).reshape(len(x), sp.dim)
        else:
            assert len(v) == sp.dim
            u0 = np.broadcast_to(np.asarray(v, dtype=np.float64), (len(x), sp.dim))
        u0 = dev.upload(np.ascontiguousarray(u0).reshape(-1))
    p0 = None
    if "pressure" in initial_conditions:
        p = initial_conditions["pressure"]
        if isinstance(p, Expression):
            p0 = dev.project(p, "Q", p.degree)
        else:
            xv = sp.mesh.coords
            p0 = np.asarray(p(xv), dtype=np.float64).reshape(-1) if callable(p) else np.full(len(xv), float(p))
            p0 = dev.upload(p0)
    return u0, p0


class IPCSSolver(SolverBase):
    def __init__(self, mesh, boundary_markers, form_convective_term, time_stepping, tol=1e-10, max_iter=50,
                 device_index=0, linear_rtol=1e-12, gmres_restart=40, cache_constant_matrices=True,
                 momentum_preconditioner="chebyshev", chebyshev_degree=10, chebyshev_ratio=30.0, partition=None,
                 preconditioner_precision="f32-scalar", pressure_preconditioner="chebyshev", pressure_chebyshev_degree=8,
                 pressure_chebyshev_ratio=300.0, pressure_solve="auto", form_viscous_term="standard",
                 linear_atol_factor=1.0, correction_rtol=None, extrapolate_pressure_guess=True, device=None,
                 newton_forcing=0.001, pressure_rhs_projection=0, extrapolate_correction_guess=1,
                 mass_chebyshev_degree=2, mass_chebyshev_ratio=20.0, extrapolate_newton_update=1):
        """``partition`` (partition.Partition, halo plans finalized): this rank's share of a
        mesh-partitioned multi-GPU run; ``mesh`` / ``boundary_markers`` are then
        the partition's local mesh and its carried-over markers.  Field vectors hold
        [owned | ghost] entries, every linear system lives on the owned entries."""
        if partition is not None:
            mesh, boundary_markers = partition.mesh, partition.mesh.bfacet_marker
        super().__init__(mesh, boundary_markers, form_convective_term, form_viscous_term)
        assert isinstance(max_iter, int) and max_iter > 0 and isinstance(tol, float) and tol > 0.0
        assert hasattr(time_stepping, "n_levels") and hasattr(time_stepping, "coefficients")
        self._time_stepping = time_stepping
        self._tol, self._maxiter = tol, max_iter
        # Krylov tolerances (relative residual) replacing the reference's LU: 1e-12 for the momentum and pressure solves,
        # three digits more for the mass matrix of the correction step.  Measured on the 16^3 cavity against the oracle at
        # 1e-14 (tools/probe_parity16.py, profiles/r2_parity16_tolerances.jsonl): with every solve at 1e-12 the fields
        # after two steps differ by 1.2e-8 / 2.2e-8 (velocity / pressure); tightening the pressure solve or the last
        # Newton update changes nothing (Newton residual 5e-15 either way); with the mass solve at 1e-13 they agree to
        # 6e-10 / 3e-12.  The unscaled P2 mass matrix is the ill-conditioned one (vertex functions of a P2 tetrahedron
        # have tiny diagonal entries), and its solution IS the new velocity.
        self._linear_rtol, self._restart = linear_rtol, gmres_restart
        # At C5 (h 4.5x smaller, 11 steps) the same experiment with the GPU path alone (profiles/r2_mass_tolerance_c5.jsonl):
        # ||u|| moves by 4e-8 relative between mass tolerances 1e-13 and 1e-14 and by 4.5e-9 between 1e-14 and 1e-15 --
        # one digit of the mass tolerance is one digit of the fields, with a factor ~4e5 between them.  Default:
        # three digits below linear_rtol (1e-15: fields reproducible to ~5e-10 whatever the initial guess, the number of
        # ranks or the summation order); the extrapolated correction guess (below) pays for the extra iterations.
        self._correction_rtol = 1e-3 * linear_rtol if correction_rtol is None else correction_rtol
        # Newton updates are solved to ||r|| <= max(linear_rtol ||b||, linear_atol_factor * linear_rtol * r0), r0 = the
        # step's first Newton residual.  1.0 (default): every update to 1e-12 r0 -- the converged u* then differs from an
        # LU-exact Newton iterate by O(1e-12 r0), inside dolfin's own stopping tolerance (1e-10).  Measured at 16^3,
        # dt = 1e-2 against the oracle at 1e-15 (profiles/r2_forcing_parity_16cubed.jsonl): 1.1e-10 / 1.4e-10 (velocity /
        # pressure) with 1.0, and the same with 0.1 and 0.01 (one or two more GMRES iterations): the factor is a knob for
        # harder problems, not needed here.  (The 1.9e-8 once attributed to it was the mass-solve tolerance, above.)
        self._linear_atol_factor = linear_atol_factor
        # Forcing term of the FIRST Newton update of a step (inexact Newton, Eisenstat-Walker): Newton contracts the
        # residual by rho = r1 / r0 per step (observed at the previous time step, 1e-4 .. 1e-6 here), so a linear
        # residual of newton_forcing * rho * r0 perturbs r1 by at most newton_forcing (0.1 %) -- the iteration counts and
        # the converged u* (whose accuracy is set by the LAST update, solved to linear_rtol * r0 as before) do not
        # change (tools/probe_newton.py, C5 and 16^3: identical counts over 14 steps, fields equal to 7e-13), the first
        # GMRES solve needs 4-5 iterations instead of 10.  Used only when the predicted r1 = rho * r0 is
        # ten times above Newton's own tolerance, i.e. when a second, exact update is certain to follow.  0 switches it
        # off (every update to linear_rtol * r0); the first step of a run has no history and is solved exactly.
        self._newton_forcing, self._newton_rho, self._newton_r1 = float(newton_forcing), None, None
        # Pressure projection: the same SPD operator every step -> initial guess = energy-norm best approximation of the
        # new solution in the span of the last `pressure_rhs_projection` solutions (Fischer 1998; nsb_rhs_guess /
        # nsb_rhs_update).  For the all-Neumann problem on one GPU and in the replicated solve; 0 = off (default: the
        # guess is then the linear extrapolation of the last two pressures).  The CG tolerance is unchanged.  Measured
        # on the C5 start-up transient (profiles/r2_pressure_projection_probe.jsonl): 54 iterations instead of 73 once
        # ten solutions are in the basis, but 70-79 again after the basis is thinned out -- on average no better than
        # the extrapolation (72), hence not the default.
        self._p_basis_size, self._p_basis = int(pressure_rhs_projection), None
        self._extrapolate_corr, self._corr_prev, self._corr_k = bool(extrapolate_correction_guess), None, None
        # Mass solve of the correction step: Jacobi-CG (degree 0) or CG preconditioned by a single-precision Chebyshev
        # polynomial in D^-1 M of this degree on [lambda_max / ratio, lambda_max] -- the float copy of the Kronecker mass
        # matrix runs through the block-scalar kernel of the momentum preconditioner (8 instead of 12 bytes per entry,
        # one aligned 8-byte bf16 gather per entry instead of three unaligned doubles, the step fused in its epilogue)
        # Measured at C5 (profiles/r2_mass_chebyshev_sweep.jsonl): Jacobi 22 iterations, 93.2 ms/step; degree 2: 12 iterations,
        # 89.7 ms/step (ratios 8 .. 30 alike); degree 3: 10 iterations, 91.0; degree 4: 9 iterations, 92.9.
        self._mass_cheb = (int(mass_chebyshev_degree), float(mass_chebyshev_ratio))
        # Initial guess of the FIRST Newton update of a step: the total update of the previous step, scaled by the
        # step-size ratio (u* moves smoothly in time, so does its change per step).  The linear solve keeps its tolerance;
        # GMRES then starts from the true residual of that guess (one more operator SpMV).
        # Measured at C5 (steps 6-15 from rest): first solve 5 -> 4 iterations, 87.4 -> 84.5 ms per step, checksums equal to 5e-14.
        self._extrapolate_dx, self._dx_acc, self._dx_prev, self._dx_k = bool(extrapolate_newton_update), None, None, None
        self._cache = cache_constant_matrices
        assert momentum_preconditioner in ("jacobi", "chebyshev")
        assert pressure_preconditioner in ("jacobi", "chebyshev")
        # pressure Poisson: CG preconditioned by a Chebyshev polynomial in D^-1 L (SPD: always safe).  About as
        # many SpMVs as Jacobi-CG but 7x fewer CG iterations, i.e. 7x fewer inner products -- the all-reduces
        # that bound the partitioned runs -- and fewer launches on one GPU
        self._p_pc = (pressure_preconditioner, pressure_chebyshev_degree, pressure_chebyshev_ratio)
        self._mom_pc, self._cheb_degree, self._cheb_ratio = momentum_preconditioner, chebyshev_degree, chebyshev_ratio
        # ``device``: an existing Device of this mesh (the dolfin-named compatibility layer shares one per mesh)
        self.dev = dev = device if device is not None else Device(mesh, device_index, partition=partition)
        sp = dev.spaces
        nV, nQ = sp.ndofs_V, sp.ndofs_Q                      # local: owned + ghost
        self._nV_own, self._nQ_own = sp.dim * dev.n_owned_nodes, dev.n_owned_verts
        # fields (source/ns_ipcs_solver.py:66-82)
        self._velocities = [dev.zeros(nV) for _ in range(time_stepping.n_levels() + 1)]
        while len(self._velocities) < 3:
            self._velocities.append(dev.zeros(nV))
        self._intermediate_velocity = dev.zeros(nV)
        self._pressure, self._old_pressure = dev.zeros(nQ), dev.zeros(nQ)
        # initial guess of the pressure CG: p^n + omega (p^n - p^{n-1}) instead of p^n (the reference's LU needs none; the
        # converged solution is the same to the solver tolerance, the iteration count is not)
        self._extrapolate_p = extrapolate_pressure_guess
        self._p_hist, self._p_hist_steps, self._prev_k = None, 0, None
        # work space
        p22, p11 = dev.pattern("P2P2"), dev.pattern("P1P1")
        d = sp.dim
        self._A_mom = dev.empty(p22.nnz * d * d)
        # single-precision copy of the Jacobian values, read only by the SpMVs INSIDE the Chebyshev
        # polynomial; GMRES itself applies the FP64 matrix and tests the FP64 residual.
        #   "f32-scalar" (default) the polynomial in single precision (padded vectors, arithmetic) inside a
        #                flexible GMRES, which keeps z_j = M^-1 v_j and so tolerates the rounding; its matrix
        #                is the block-scalar part S (x) I of J (mean block diagonals, 8 B per block entry)
        #   "f32"        the same with the full d x d blocks in single precision (40 B per block entry)
        #   "f32-matrix" single-precision values only, FP64 vectors: still a fixed linear preconditioner
        #   "f64"        everything FP64
        assert preconditioner_precision in ("f32-scalar", "f32", "f32-matrix", "f64")
        cheb = momentum_preconditioner == "chebyshev"
        self._pc_lowp = cheb and preconditioner_precision in ("f32", "f32-scalar")
        self._A_mom32 = dev.empty(p22.nnz * d * d, dtype=torch.float32) \
            if (cheb and preconditioner_precision in ("f32", "f32-matrix")) else None
        self._S_mom32 = dev.empty(p22.nnz, dtype=torch.float32) \
            if (cheb and preconditioner_precision == "f32-scalar") else None
        self._A_mass = dev.empty(p22.nnz)
        self._A_lap = dev.empty(p11.nnz)
        nVo, nQo = self._nV_own, self._nQ_own
        self._bV, self._dxV, self._dinvV = dev.empty(nVo), dev.empty(nVo), dev.empty(nVo)
        self._dinv_mass = dev.empty(nVo)
        self._bQ, self._dinvQ = dev.empty(nQo), dev.empty(nQo)
        self._work = dev.empty(max(dev.lib.nsb_gmres_workspace_doubles(nVo, gmres_restart, 1 if self._pc_lowp else 0),
                                   dev.lib.nsb_cg_workspace_doubles(nVo)))
        # Pressure problem on a partitioned mesh (405 k rows at C5, an operator that fits one GPU's
        # L2): "replicated" = every rank all-reduces the right-hand side once and solves the WHOLE system itself with
        # the persistent CG kernel -- no halo exchange, no all-reduce inside the solve, bit-identical on all ranks;
        # "partitioned" = owned rows per rank, one halo exchange per SpMV and one all-reduce per inner product.
        assert pressure_solve in ("auto", "replicated", "partitioned")
        if pressure_solve == "auto":
            pressure_solve = "replicated" if (partition is not None and partition.nranks > 1
                                              and partition.verts.n_global <= 4_000_000) else "partitioned"
        self._p_replicated = partition is not None and pressure_solve == "replicated"
        if self._p_replicated:
            self._pdev = pdev = Device(partition.global_mesh, device_index, patterns=("P1P1",))
            nQg = pdev.spaces.ndofs_Q
            self._A_lap_g = pdev.empty(pdev.patterns["P1P1"].nnz)
            self._bQ_g, self._p_g, self._dinvQ_g = pdev.zeros(nQg), pdev.zeros(nQg), pdev.empty(nQg)
            self._work_g = pdev.empty(pdev.lib.nsb_cg_workspace_doubles(nQg))
            self._own_gid = dev.upload(partition.vert_gid[:nQo].astype(np.int32))     # owned local vertex -> global
            self._loc_gid = dev.upload(partition.vert_gid.astype(np.int32))           # every local vertex -> global
            self._p_g_stale = True
        self._pi_work = None
        self._have_constant = False
        self._bc_ready = False
        self.newton_iterations, self.krylov_iterations = [], []
        self.chebyshev_fallbacks = 0
        self._alpha, self._next_step_size = None, None

    # ---- BC tables on the device --------------------------------------------------------------
    def _setup_boundary_conditions(self):
        """source/ns_ipcs_solver.py:45-64: velocity BCs on V, pressure BCs on Q."""
        dev, sp = self.dev, self.dev.spaces
        vd, vv = sp.dirichlet_table(self._velocity_bcs, "V")
        pd, pv = sp.dirichlet_table(self._pressure_bcs, "Q")
        # Dirichlet rows exist for owned dofs only; ghosts receive their values from the owner
        self._vbc_keep, self._pbc_keep = vd < self._nV_own, pd < self._nQ_own
        vd, vv, pd, pv = vd[self._vbc_keep], vv[self._vbc_keep], pd[self._pbc_keep], pv[self._pbc_keep]
        assert dev.partition is not None or len(vd) + len(pd) > 0
        self._vbc_dofs, self._vbc_vals = dev.upload(vd), dev.upload(vv)
        self._vbc_nodes = dev.upload(np.unique(np.asarray(vd) // dev.dim).astype(np.int32))     # block rows with a Dirichlet dof
        self._pbc_dofs, self._pbc_vals = dev.upload(pd), dev.upload(pv)
        # Dirichlet or all-Neumann pressure problem: one decision for all ranks
        self._pressure_dirichlet = len(pd) > 0
        if dev.partition is not None:
            self._pressure_dirichlet = dev.sum(dev.upload(np.array([float(len(pd))]))) > 0.0
        mV = np.zeros(sp.ndofs_V, dtype=np.uint8)
        mV[vd] = 1
        mQ = np.zeros(sp.ndofs_Q, dtype=np.uint8)
        mQ[pd] = 1
        self._maskV, self._maskQ = dev.upload(mV), dev.upload(mQ)
        if self._p_replicated:
            gsp = self._pdev.spaces
            gd, gv = gsp.dirichlet_table(self._pressure_bcs, "Q")
            mg = np.zeros(gsp.ndofs_Q, dtype=np.uint8)
            mg[gd] = 1
            self._pbc_dofs_g, self._pbc_vals_g, self._maskQ_g = dev.upload(gd), dev.upload(gv), dev.upload(mg)
        self._time_dependent_bcs = any(isinstance(v, Expression) for _, _, v in
                                       self._velocity_bcs + self._pressure_bcs)
        self._bc_ready = True
        self._have_constant = False

    def _update_bc_values(self, force=False):
        """Dirichlet data: time-dependent values are re-evaluated on the host (the Python
        ``Expression`` callbacks of the reference, H11) and uploaded from pinned host memory.
        ``force`` uploads the tables even when they are constant (dolfin re-evaluates every
        DirichletBC on each apply; bench.py's end-to-end leg counts these bytes)."""
        if not (self._time_dependent_bcs or force):
            return 0
        sp = self.dev.spaces
        if not hasattr(self, "_vbc_host"):
            self._vbc_host = self.dev.host_buffer(self._vbc_vals.numel())
            self._pbc_host = self.dev.host_buffer(self._pbc_vals.numel())
            self._vbc_host.copy_(self._vbc_vals)
            self._pbc_host.copy_(self._pbc_vals)
        if self._time_dependent_bcs:
            _, vv = sp.dirichlet_table(self._velocity_bcs, "V")
            self._vbc_host.copy_(torch.from_numpy(vv[self._vbc_keep]))
            if self._pbc_dofs.numel():
                _, pv = sp.dirichlet_table(self._pressure_bcs, "Q")
                self._pbc_host.copy_(torch.from_numpy(pv[self._pbc_keep]))
            if self._p_replicated and self._pbc_dofs_g.numel():
                _, gv = self._pdev.spaces.dirichlet_table(self._pressure_bcs, "Q")
                self._pbc_vals_g.copy_(self.dev.upload(gv))
        self._vbc_vals.copy_(self._vbc_host, non_blocking=True)
        nbytes = self._vbc_host.numel() * 8
        if self._pbc_dofs.numel():
            self._pbc_vals.copy_(self._pbc_host, non_blocking=True)
            nbytes += self._pbc_host.numel() * 8
        return nbytes

    # ---- checkpoint / resume of the time loop (source/ns_problem.py:601-653 writes the solution and the
    # time-stepping state; here both stay on the device) -------------------------------------------
    def get_state(self):
        """device copies of the BDF history, pressures and the time-stepping object's state"""
        import copy
        return dict(velocities=[v.clone() for v in self._velocities], ustar=self._intermediate_velocity.clone(),
                    pressure=self._pressure.clone(), old_pressure=self._old_pressure.clone(),
                    p_hist=None if self._p_hist is None else self._p_hist.clone(), p_hist_steps=self._p_hist_steps,
                    prev_k=self._prev_k,
                    time_stepping=copy.deepcopy(self._time_stepping.__dict__))

    def set_state(self, state):
        import copy
        for v, s in zip(self._velocities, state["velocities"]):
            v.copy_(s)
        self._intermediate_velocity.copy_(state["ustar"])
        self._pressure.copy_(state["pressure"])
        self._old_pressure.copy_(state["old_pressure"])
        self._p_hist = None if state.get("p_hist") is None else state["p_hist"].clone()
        self._p_hist_steps, self._prev_k = state.get("p_hist_steps", 0), state.get("prev_k")
        self._prev_k2 = None                      # the p^{n-2} buffer of the quadratic pressure guess is not part of the state
        # predictors built from the steps before the restore say nothing about the restored state
        self._newton_rho = self._newton_r1 = None
        self._corr_prev = self._corr_k = None
        self._dx_prev = self._dx_k = None
        if self._p_replicated:
            self._p_g_stale = True
        self._time_stepping.__dict__.update(copy.deepcopy(state["time_stepping"]))

    def fetch_solution(self, out=None, wait=True):
        """Device -> pinned host copy of the mixed solution vector (what the reference hands to
        its XDMF writer, source/ns_problem.py:733-735).  Returns (host tensor, bytes).
        ``wait=False``: the copy runs on a side stream behind the kernels issued so far and overlaps the next time
        step; ``wait_solution()`` (or the next ``fetch_solution`` / ``solution``) waits for it -- the host tensor
        must not be read before."""
        self.wait_solution()                      # the previous copy still owns the host buffer and self._w
        w = self.solution
        if out is None:
            if not hasattr(self, "_w_host"):
                self._w_host = self.dev.host_buffer(w.numel())
            out = self._w_host
        cur = torch.cuda.current_stream(self.dev.torch_device)
        if wait or self.dev.torch_device.type != "cuda":
            out.copy_(w, non_blocking=True)
            cur.synchronize()
            return out, w.numel() * 8
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(self.dev.torch_device)
        ready = torch.cuda.Event()
        ready.record(cur)
        self._copy_stream.wait_event(ready)
        with torch.cuda.stream(self._copy_stream):
            out.copy_(w, non_blocking=True)
        self._copy_done = torch.cuda.Event()
        self._copy_done.record(self._copy_stream)
        return out, w.numel() * 8

    def wait_solution(self):
        """block until the copy started by ``fetch_solution(wait=False)`` has arrived"""
        ev = getattr(self, "_copy_done", None)
        if ev is not None:
            ev.synchronize()
            self._copy_done = None

    # ---- constant operators (pressure stiffness, velocity mass), quirk Q8 ----------------------
    def _assemble_constant(self):
        dev, d = self.dev, self.dev.dim
        dev.assemble_poisson(L.MAT, A=self._A_lap)
        dev.assemble_correction(L.MAT, A=self._A_mass)
        self._lap = dev.matrix("P1P1", self._A_lap, L.CSR, 1, self._maskQ if self._pressure_dirichlet else None,
                               compress=dev.partition is None or dev.partition.nranks == 1)
        self._A_mass32 = None
        if self._mass_cheb[0] > 1 and d in (2, 3):
            self._A_mass32 = dev.empty(self._A_mass.numel(), dtype=torch.float32)
            dev.to_f32(self._A_mass, self._A_mass32)
        self._mass = dev.matrix("P2P2", self._A_mass, L.CSR, d, self._maskV, diag32=self._A_mass32)
        self._jac = dev.matrix("P2P2", self._A_mom, L.BSR, d, None, vals32=self._A_mom32, diag32=self._S_mom32)
        if self._p_replicated:          # the whole pressure operator on every rank (its own single-rank context)
            pdev = self._pdev
            pdev.assemble_poisson(L.MAT, A=self._A_lap_g)
            self._lap_g = pdev.matrix("P1P1", self._A_lap_g, L.CSR, 1, self._maskQ_g if self._pressure_dirichlet else None,
                                      compress=True)
            lap, dinvQ, ldev = self._lap_g, self._dinvQ_g, pdev
        else:
            lap, dinvQ, ldev = self._lap, self._dinvQ, dev
        ldev.inv_diagonal(lap, dinvQ)
        if self._p_pc[0] == "chebyshev":
            lmax = 1.1 * ldev.lambda_max(lap, dinvQ, iters=30, work=self._work_g if self._p_replicated else self._work)
            self._pcQ = ldev.precond(dinvQ, L.PC_CHEBYSHEV, self._p_pc[1], lmax / self._p_pc[2], lmax)
        else:
            self._pcQ = ldev.precond(dinvQ)
        dev.inv_diagonal(self._mass, self._dinv_mass)
        self._pc_mass = self._dinv_mass
        if self._A_mass32 is not None:
            lmax = 1.1 * dev.lambda_max(self._mass, self._dinv_mass, iters=30, work=self._work, lowp=True)
            self._pc_mass = dev.precond(self._dinv_mass, L.PC_CHEBYSHEV, self._mass_cheb[0], lmax / self._mass_cheb[1], lmax,
                                        lowp=True)
        self._have_constant = True

    def _update_time_stepping_coefficients(self):
        """source/ns_ipcs_solver.py:210-227"""
        ts = self._time_stepping
        alpha = list(ts.coefficients(1))
        while len(alpha) < 3:
            alpha.append(0.0)
        self._alpha, self._next_step_size = tuple(alpha), ts.get_next_step_size()
        co = self._equation_coefficients
        self.dev.set_coefficients(self._next_step_size, self._alpha, co["convective_term"], co["pressure_term"],
                                  co["viscous_term"])

    def set_initial_conditions(self, initial_conditions):
        """source/ns_solver_base.py:1123-1171 + source/ns_ipcs_solver.py:229-238: ``dolfin.project`` of the initial
        data into the velocity / pressure space, assigned to velocities[0], [1], pressure, old_pressure.
        An ``Expression(fn, degree=k)`` is projected on the device (cell-wise P_k interpolant, exact right-hand
        side, mass solve: ``Device.project``); constants are in the space, their projection is the constant; a bare
        callable is taken as nodal data of degree <= 2 (velocity) / 1 (pressure), for which the projection is the
        identity."""
        assert isinstance(initial_conditions, dict) and "velocity" in initial_conditions
        sp, dev = self.dev.spaces, self.dev
        u0, p0 = initial_fields(dev, initial_conditions)
        self._velocities[0].copy_(u0)
        self._velocities[1].copy_(u0)
        if p0 is not None:
            self._pressure.copy_(p0)
            self._old_pressure.copy_(p0)
        if self._p_replicated:
            self._p_g_stale = True
        self._newton_rho = self._newton_r1 = None
        self._corr_prev = self._corr_k = None
        self._dx_prev = self._dx_k = None
        del sp

    # ---- one time step -------------------------------------------------------------------------
    def solve(self, upload_bcs=False):
        """``InstationarySolverBase.solve`` (source/ns_solver_base.py:1174-1203)"""
        if not self._bc_ready:
            self._setup_boundary_conditions()
        self._set_time(self._time_stepping.next_time)
        self._apply_forms(self.dev)
        self._load = self._external_load(self.dev)
        self.h2d_bytes = self._update_bc_values(force=upload_bcs)
        self._update_time_stepping_coefficients()
        if not (self._have_constant and self._cache):
            self._assemble_constant()
        self._solve_time_step()

    def _residual(self):
        dev = self.dev
        dev.assemble_ipcs_momentum(L.VEC, self._intermediate_velocity, self._velocities[1], self._velocities[2],
                                   self._old_pressure, b=self._bV)
        if getattr(self, "_load", None) is not None:      # F += -c_b int f.w dx + int_Gamma t.w ds
            dev.axpy(1.0, self._load[:self._nV_own], self._bV)
        dev.bc_apply_vector(L.BC_NEWTON, self._bV, self._vbc_dofs, self._vbc_vals, x=self._intermediate_velocity)
        return dev.norm(self._bV)

    def _solve_time_step(self):
        """source/ns_ipcs_solver.py:198-208: the three sub-step solvers in sequence"""
        self._solve_diffusion_step()
        self._solve_projection_step()
        self._solve_correction_step()

    def _solve_diffusion_step(self):
        """``self._diffusion_solver.solve()`` (source/ns_ipcs_solver.py:202; forms :106-147)"""
        dev, d = self.dev, self.dev.dim
        ustar = self._intermediate_velocity
        # (1) tentative velocity: dolfin NewtonSolver semantics (atol = tol, rtol = 10 tol)
        atol, rtol = self._tol, 10.0 * self._tol
        r0 = r = self._residual()
        it, kry = 0, []
        self.newton_residuals = [r0]           # residual norms of this step's Newton iteration (diagnostic)
        converged = newton_converged(r, r0, atol, rtol)
        while not converged and it < self._maxiter:
            # the block-scalar f32 copy of the preconditioner is written by the assembly kernel itself (registered only
            # for the duration of the call: the Device may be shared), then refreshed on the Dirichlet block rows
            dev.set_block_scalar_out(self._S_mom32)
            dev.assemble_ipcs_momentum(L.MAT, ustar, A=self._A_mom)
            dev.set_block_scalar_out(None)
            dev.bc_apply_matrix("P2P2", self._A_mom, L.BSR, d, self._vbc_dofs)
            dev.inv_diagonal(self._jac, self._dinvV)
            if self._A_mom32 is not None:
                dev.to_f32(self._A_mom, self._A_mom32)
            if self._S_mom32 is not None:
                dev.block_scalar_rows_f32(self._A_mom, d, self._vbc_nodes, self._S_mom32)
            if self._mom_pc == "chebyshev":
                if it == 0:      # spectrum of D^-1 J changes little within a step: estimate once per step,
                    # and little from step to step: the power iteration continues from the previous
                    # step's iterate (1 iteration = 2 SpMVs) instead of starting over (8)
                    warm = self._pi_work is not None
                    if not warm:
                        self._pi_work = dev.empty(2 * self._nV_own)
                    # on the block-scalar f32 operator when that is what the polynomial is built from
                    self._lmax = 1.1 * dev.lambda_max(self._jac, self._dinvV, iters=1 if warm else 8,
                                                      work=self._pi_work, warm_start=warm,
                                                      lowp=self._S_mom32 is not None)
                pc = dev.precond(self._dinvV, L.PC_CHEBYSHEV, self._cheb_degree, self._lmax / self._cheb_ratio,
                                 self._lmax, lowp=self._pc_lowp)
            else:
                pc = dev.precond(self._dinvV)
            x0_zero = True
            if self._extrapolate_dx and it == 0 and self._dx_prev is not None and self._dx_k:
                dev.copy(self._dx_prev, self._dxV)
                dev.scale(self._next_step_size / self._dx_k, self._dxV)
                x0_zero = False
            else:
                self._dxV.zero_()
            # every Newton update is solved to linear_rtol relative to the step's INITIAL residual r0
            # (the accuracy LU-exact Newton reaches anyway): later iterations need fewer Krylov steps
            eta = self._linear_rtol
            if it == 0 and self._newton_forcing > 0.0 and self._newton_rho is not None:
                # predicted residual after this update: the SMALLER of (previous contraction) x r0 and the previous
                # step's r1 itself -- when r0 jumps (restart, a step change of the boundary data) the contraction of the
                # last step says nothing, and dolfin's own exit rule (r2 < 1e-9 r0) leaves u* sensitive to r1 at that level
                r1_pred = min(self._newton_rho * r0, self._newton_r1)
                # only when a second (exact) update is certain to follow: the last update sets the accuracy of u*
                if r1_pred > 10.0 * max(atol, rtol * r0):
                    eta = min(1e-3, max(eta, self._newton_forcing * r1_pred / r0))
            kw = dict(rtol=eta, atol=self._linear_atol_factor * eta * r0, restart=self._restart, work=self._work)
            if self._mom_pc == "chebyshev":
                # the polynomial is only safe while the spectrum of D^-1 J stays near the real interval:
                # strongly non-normal Jacobians (coarse meshes, sharp inflow transients) stall it.  Give it
                # three restart cycles, then finish from the current iterate with plain Jacobi.
                try:
                    n_it, _ = dev.gmres(self._jac, pc, self._bV, self._dxV, maxit=3 * self._restart, x0_is_zero=x0_zero, **kw)
                except RuntimeError as err:
                    if "did not converge" not in str(err):
                        raise
                    self.chebyshev_fallbacks += 1
                    n_j, _ = dev.gmres(self._jac, dev.precond(self._dinvV), self._bV, self._dxV, **kw)
                    n_it = 3 * self._restart + n_j
            else:
                n_it, _ = dev.gmres(self._jac, pc, self._bV, self._dxV, x0_is_zero=x0_zero, **kw)
            kry.append(n_it)
            if self._extrapolate_dx:
                if self._dx_acc is None:
                    self._dx_acc = dev.empty(self._dxV.numel())
                if it == 0:
                    dev.copy(self._dxV, self._dx_acc)
                else:
                    dev.axpy(1.0, self._dxV, self._dx_acc)
            dev.axpy(-1.0, self._dxV, ustar)
            dev.halo_exchange(L.HALO_NODES, d, ustar)
            r = self._residual()
            self.newton_residuals.append(r)
            it += 1
            if it == 1:
                # contraction of the first update; an inexact first update that was followed by a poor contraction
                # is not trusted as a predictor for the next step
                ok = r0 > 0.0 and (eta == self._linear_rtol or r / r0 < 1e-2)
                self._newton_rho, self._newton_r1 = (r / r0, r) if ok else (None, None)
            converged = newton_converged(r, r0, atol, rtol)
        if self._extrapolate_dx and it > 0:
            self._dx_acc, self._dx_prev = self._dx_prev, self._dx_acc
            self._dx_k = self._next_step_size
        self.newton_iterations.append(it)
        self.newton_residual = r
        if not converged:
            raise NewtonNotConverged("Newton solver did not converge (%d iterations, residual %g)" % (it, r))
        self._kry_momentum = kry

    def _solve_projection_step(self):
        """``self._projection_solver.solve()`` (source/ns_ipcs_solver.py:205; forms :149-171)"""
        dev = self.dev
        ustar = self._intermediate_velocity
        dev.assemble_poisson(L.VEC, pn=self._old_pressure, ustar=ustar, b=self._bQ)
        self._extrapolate_pressure_guess()
        if self._p_replicated:
            n_p = self._solve_pressure_replicated()
        elif self._pressure_dirichlet:
            dev.bc_apply_vector(L.BC_LINEAR, self._bQ, self._pbc_dofs, self._pbc_vals)
            dev.bc_apply_vector(L.BC_LINEAR, self._pressure, self._pbc_dofs, self._pbc_vals)
            n_p, _ = dev.cg(self._lap, self._pcQ, self._bQ, self._pressure, rtol=self._linear_rtol, work=self._work)
            dev.halo_exchange(L.HALO_VERTS, 1, self._pressure)
        else:
            basis = self._pressure_basis(dev, self._pressure.numel()) if dev.partition is None else None
            if basis is not None:
                dev.rhs_guess(basis, self._bQ, self._pressure)
            n_p, _ = dev.cg(self._lap, self._pcQ, self._bQ, self._pressure, rtol=self._linear_rtol,
                            nullspace=True, work=self._work)
            if basis is not None:
                dev.rhs_update(basis, self._lap, self._pressure)
            dev.halo_exchange(L.HALO_VERTS, 1, self._pressure)
        self._kry_pressure = n_p

    def _pressure_basis(self, dev, n):
        """the successive-right-hand-side basis of the all-Neumann pressure problem (None: switched off)"""
        if self._p_basis_size <= 0 or self._pressure_dirichlet:
            return None
        if self._p_basis is None or self._p_basis["n"] != n:
            self._p_basis = dev.rhs_basis(n, min(self._p_basis_size, 16))
        return self._p_basis

    def _solve_correction_step(self):
        """``self._velocity_correction_solver.solve()`` (source/ns_ipcs_solver.py:208; forms :173-196)"""
        dev, d = self.dev, self.dev.dim
        ustar = self._intermediate_velocity
        dev.assemble_correction(L.VEC, ustar=ustar, pnew=self._pressure, pold=self._old_pressure, b=self._bV)
        dev.bc_apply_vector(L.BC_LINEAR, self._bV, self._vbc_dofs, self._vbc_vals)
        dev.copy(ustar, self._velocities[0])
        # initial guess u* + omega (u^n - u*^n): the correction of the previous step, scaled by the step-size ratio
        # (the correction is -(k / alpha0) times the projected gradient of the pressure increment, smooth in time)
        if self._extrapolate_corr and self._corr_prev is not None and self._corr_k:
            dev.axpy(self._next_step_size / self._corr_k, self._corr_prev, self._velocities[0])
        dev.bc_apply_vector(L.BC_LINEAR, self._velocities[0], self._vbc_dofs, self._vbc_vals)
        n_c, _ = dev.cg(self._mass, self._pc_mass, self._bV, self._velocities[0], rtol=self._correction_rtol,
                        work=self._work)
        if self._extrapolate_corr:
            if self._corr_prev is None:
                self._corr_prev = dev.empty(self._velocities[0].numel())
            dev.copy(self._velocities[0], self._corr_prev)
            dev.axpy(-1.0, ustar, self._corr_prev)
            self._corr_k = self._next_step_size
        dev.halo_exchange(L.HALO_NODES, d, self._velocities[0])
        dev.halo_status()
        self.krylov_iterations.append(dict(momentum=self._kry_momentum, pressure=self._kry_pressure, correction=n_c))

    def _extrapolate_pressure_guess(self):
        """x0 of the pressure CG: ``_pressure`` holds p^n on entry; with the history vector p^{n-1} it becomes
        p^n + omega (p^n - p^{n-1}), omega = k_{n+1} / k_n.  The history vector is refreshed to p^n first-hand.
        ``extrapolate_pressure_guess=2``: on uniform steps the quadratic 3 p^n - 3 p^{n-1} + p^{n-2} (measured at C5,
        steps 6-15 from rest: 71 -> 64 CG iterations, 0.4 ms per step; not the default -- the linear guess is the safer
        one when the pressure is not smooth in time, and the gain is small)."""
        if not self._extrapolate_p:
            return
        dev = self.dev
        x = self._p_g if self._p_replicated and not self._p_g_stale else self._pressure
        if self._p_replicated and self._p_g_stale:
            self._p_hist_steps = 0            # the global copy is rebuilt from the local field: restart the history
        if self._p_hist is None or self._p_hist.numel() != x.numel():
            self._p_hist, self._p_hist_steps = dev.zeros(x.numel()), 0
        k = self._next_step_size
        h2 = getattr(self, "_p_hist2", None)
        if int(self._extrapolate_p) >= 2 and (h2 is None or h2.numel() != x.numel()):
            h2 = self._p_hist2 = dev.zeros(x.numel())
        uniform = bool(self._prev_k) and abs(k / self._prev_k - 1.0) < 1e-12 and \
            bool(getattr(self, "_prev_k2", None)) and abs(self._prev_k / self._prev_k2 - 1.0) < 1e-12
        if int(self._extrapolate_p) >= 2 and self._p_hist_steps >= 2 and uniform:
            # quadratic extrapolation on a uniform step: 3 p^n - 3 p^{n-1} + p^{n-2}, built in the p^{n-2} buffer
            dev.axpy(-3.0, self._p_hist, h2)
            dev.axpy(3.0, x, h2)
            guess, self._p_hist2 = h2, self._p_hist          # p^{n-1} becomes p^{n-2}
            self._p_hist = guess                                # _swap_guess leaves p^n in this buffer
            self._swap_guess(x, guess)
        elif self._p_hist_steps >= 1 and self._prev_k:
            omega = k / self._prev_k
            if h2 is not None:
                dev.copy(self._p_hist, h2)
            # x <- (1 + omega) x - omega hist ; hist <- old x      (two passes over a 3 MB vector)
            tmp = self._p_hist
            dev.scale(-omega, tmp)
            dev.axpy(1.0 + omega, x, tmp)          # tmp = (1 + omega) p^n - omega p^{n-1}
            self._swap_guess(x, tmp)
        else:
            dev.copy(x, self._p_hist)
        self._p_hist_steps += 1
        self._prev_k2, self._prev_k = self._prev_k, k

    def _swap_guess(self, x, guess):
        """x holds p^n, guess the extrapolated start: afterwards x = guess and the history vector = p^n"""
        dev = self.dev
        if not hasattr(self, "_p_tmp") or self._p_tmp.numel() != x.numel():
            self._p_tmp = dev.empty(x.numel())
        dev.copy(x, self._p_tmp)
        dev.copy(guess, x)
        dev.copy(self._p_tmp, self._p_hist)

    def _solve_pressure_replicated(self):
        """every rank solves the whole pressure system: owned right-hand-side entries -> zeroed global vector ->
        one all-reduce (each entry has a single non-zero contributor: exact) -> persistent CG on the global operator
        (no exchange inside) -> local [owned | ghost] entries gathered back, ghosts included"""
        dev, pdev = self.dev, self._pdev
        if self._p_g_stale:              # initial guess = the current pressure field (initial conditions)
            self._p_g.zero_()
            dev.scatter(self._own_gid, self._pressure[:self._nQ_own], self._p_g)
            dev.allreduce(self._p_g)
            self._p_g_stale = False
        self._bQ_g.zero_()
        dev.scatter(self._own_gid, self._bQ, self._bQ_g)
        dev.allreduce(self._bQ_g)
        if self._pressure_dirichlet:
            pdev.bc_apply_vector(L.BC_LINEAR, self._bQ_g, self._pbc_dofs_g, self._pbc_vals_g)
            pdev.bc_apply_vector(L.BC_LINEAR, self._p_g, self._pbc_dofs_g, self._pbc_vals_g)
        basis = self._pressure_basis(pdev, self._p_g.numel())
        if basis is not None:
            pdev.rhs_guess(basis, self._bQ_g, self._p_g)
        n_p, _ = pdev.cg(self._lap_g, self._pcQ, self._bQ_g, self._p_g, rtol=self._linear_rtol,
                         nullspace=not self._pressure_dirichlet, work=self._work_g)
        if basis is not None:
            pdev.rhs_update(basis, self._lap_g, self._p_g)
        dev.gather(self._loc_gid, self._p_g, self._pressure)
        return n_p

    # ---- instrumentation over both contexts (the replicated pressure solve has its own) -------------
    def _devices(self):
        return [self.dev] + ([self._pdev] if self._p_replicated else [])

    def timer_enable(self, on=True, classes=None):
        for d in self._devices():
            d.timer_enable(on, classes)

    def timer_report(self):
        out = {}
        for d in self._devices():
            for k, (ms, cnt) in d.timer_report().items():
                a = out.get(k, (0.0, 0))
                out[k] = (a[0] + ms, a[1] + cnt)
        return out

    def launch_count(self):
        return sum(d.launch_count() for d in self._devices())

    def advance_time(self):
        """``_advance_solution`` source/ns_ipcs_solver.py:35-43"""
        self.dev.ipcs_advance(self._velocities[0], self._velocities[1], self._velocities[2], self._pressure,
                              self._old_pressure)

    # ---- output ---------------------------------------------------------------------------------
    @property
    def velocity(self):
        return self._velocities[0]

    @property
    def pressure(self):
        return self._pressure

    @property
    def solution(self):
        """mixed-space vector (source/ns_ipcs_solver.py:240-247), with the mean-pressure shift
        of source/ns_solver_base.py:1190-1203 applied to the copy when requested."""
        dev, sp = self.dev, self.dev.spaces
        self.wait_solution()                      # an overlapped copy of the previous solution still reads self._w
        if not hasattr(self, "_w"):
            self._w = dev.zeros(sp.ndofs_W)
            self._v2w, self._q2w = dev.upload(sp.V_to_W()), dev.upload(sp.Q_to_W())
        dev.scatter(self._v2w, self._velocities[0], self._w)
        dev.scatter(self._q2w, self._pressure, self._w)
        if hasattr(self, "_mean_pressure_value"):
            # source/ns_solver_base.py:1190-1203, on the mixed copy like the reference: int p / int 1 from the P1
            # mass weights, then p - shift (a constant shift of a P1 function IS its L2 projection)
            shift = self.mean_pressure() - self._mean_pressure_value
            dev.shift(self._w[sp.dim * sp.n_nodes:], shift)
        return self._w

    def checksums(self, lagged=False):
        """{||u||_2, ||p - mean p||_2} over the owned dofs of all ranks (all-reduced): what bench.py prints so
        that runs on 1, 2, 4 and 8 GPUs can be compared.  ``lagged``: the fields after ``advance_time``
        (velocities[1], old_pressure) instead of the current ones."""
        dev = self.dev
        u = (self._velocities[1] if lagged else self._velocities[0])[:self._nV_own]
        p = (self._old_pressure if lagged else self._pressure)[:self._nQ_own]
        nq = float(dev.partition.verts.n_global) if dev.partition is not None else float(self._nQ_own)
        pp, ps = dev.dot(p, p), dev.sum(p)
        return {"u_l2": float(np.sqrt(dev.dot(u, u))), "p_minus_mean_l2": float(np.sqrt(max(pp - ps * ps / nq, 0.0)))}

    def mean_pressure(self):
        """int p dx / int 1 dx (source/ns_solver_base.py:1194)"""
        integral, volume = self.dev.p1_integral(self._pressure)
        return integral / volume

    def cfl_number(self, step_size=None):
        """``InstationaryProblem._compute_cfl_number`` (source/ns_problem.py:554-587) for the current velocity"""
        k = self._time_stepping.get_next_step_size() if step_size is None else step_size
        return self.dev.cfl(self._velocities[0], k)
