"""Shared machinery of the two monolithic solver mirrors (``bdf.py``, ``stationary.py``): the mixed
Taylor-Hood space W on the device, its Dirichlet table, dolfin's NewtonSolver loop and the GPU
direct solve of the saddle-point Newton system (K16, library leg timed apart from the kernels).

Reference: ``NonlinearVariationalSolver`` / ``NewtonSolver`` as driven by
source/ns_bdf_solver.py:91-106 and source/ns_solver_base.py:938-988 (semantics: SURVEY.md App. A.5).
"""
import numpy as np

from . import _lib as L
from .device import Device
from .solver_base import SolverBase, Expression, NewtonNotConverged


class MixedSolverBase(SolverBase):
    def __init__(self, mesh, boundary_markers, form_convective_term="standard", device_index=0, reorder=1,
                 linear_solver="auto", linear_rtol=1e-13, gmres_restart=60, form_viscous_term="standard", device=None):
        """``linear_solver`` (K16, the Newton update the reference hands to sparse LU): "gmres" = hand-written
        flexible GMRES with the block-triangular Cahouet-Chabard preconditioner (csrc/krylov.cu saddle_apply);
        "direct" = cuSOLVER sparse QR (library leg); "auto" = direct below 20 000 dofs and for stationary problems
        (no mass term: the Schur complement approximation degrades with the Reynolds number), gmres otherwise."""
        super().__init__(mesh, boundary_markers, form_convective_term, form_viscous_term)
        assert linear_solver in ("auto", "gmres", "direct")
        self._linear_solver, self._linear_rtol, self._gmres_restart = linear_solver, linear_rtol, gmres_restart
        # knobs of the block-triangular preconditioner (nsb_saddle): inner CG tolerance, velocity-block Chebyshev
        # degree and the ratio lambda_max / lambda_min of its interval
        self.saddle_options = dict(inner_rtol=1e-2, vel_degree=8, vel_ratio=8.0, lp_degree=8, lp_ratio=300.0)
        self.krylov_iterations, self.inner_iterations = [], 0
        # inexact first Newton update of a time step (instationary solver, hand-written K16 only); 0 = every update exact
        self.newton_forcing, self._newton_rho, self._newton_r1 = 0.0, None, None
        # ``device``: an existing Device of this mesh (the dolfin-named compat layer shares one per mesh)
        self.dev = dev = Device(mesh, device_index, mixed=True) if device is None else device
        self._nW = dev.spaces.ndofs_W
        self._pat = dev.pattern("WW")
        self._A = dev.empty(self._pat.nnz)
        self._b, self._dx = dev.empty(self._nW), dev.empty(self._nW)
        self._reorder = reorder
        self.pin_pressure = None          # None: pin iff the flow is enclosed (see _dirichlet_table); True / False force it
        self._bc_ready = False
        self.newton_iterations, self.newton_history = [], []

    # ---- Dirichlet data on W.sub(0) / W.sub(1) (source/ns_solver_base.py:546-660) ----------------
    def _setup_boundary_conditions(self):
        dev, sp = self.dev, self.dev.spaces
        dofs, vals = self._dirichlet_table()
        assert len(dofs) > 0
        self._bc_dofs, self._bc_vals = dev.upload(dofs), dev.upload(vals)
        mask = np.zeros(self._nW, dtype=np.uint8)
        mask[dofs] = 1
        g = np.zeros(self._nW)
        g[dofs] = vals
        self._bc_mask, self._bc_g = dev.upload(mask), dev.upload(g)
        dev.set_mixed_bc(self._bc_mask, self._bc_g)
        self._time_dependent_bcs = any(isinstance(v, Expression) for _, _, v in
                                       self._velocity_bcs + self._pressure_bcs)
        self._bc_ready = True

    def _dirichlet_table(self):
        """Dirichlet dofs / values on W.  Without any pressure condition (cavity flows) the mixed
        Jacobian is singular in p and the reference leans on LU pivoting: here the
        first pressure dof is pinned to 0, which selects one member of the same solution family
        (the velocity is unaffected; pressures compare up to a constant)."""
        sp = self.dev.spaces
        dofs, vals = sp.dirichlet_table(self._velocity_bcs + self._pressure_bcs, "W")
        pin = self.pin_pressure
        if pin is None:
            # enclosed flow = every exterior facet carries a Dirichlet condition on the whole velocity vector; a
            # natural (do-nothing / traction) boundary fixes the pressure level by itself
            d = sp.dim
            covered = {}
            for bid, comp, _ in self._velocity_bcs:
                covered.setdefault(int(bid), set()).update(range(d) if comp is None else [int(comp)])
            markers = set(int(m) for m in np.unique(self._mesh.bfacet_marker))
            pin = all(len(covered.get(m, ())) == d for m in markers)
        if not self._pressure_bcs and pin:
            pin = int(sp.Q_to_W()[0])
            dofs, vals = np.append(dofs, pin), np.append(vals, 0.0)
            order = np.argsort(dofs, kind="stable")
            dofs, vals = dofs[order].astype(np.int32), vals[order]
        return dofs, vals

    def _update_bc_values(self):
        if not self._time_dependent_bcs:
            return
        dev = self.dev
        dofs, vals = self._dirichlet_table()
        self._bc_vals.copy_(dev.upload(vals))
        dev.scatter(self._bc_dofs, self._bc_vals, self._bc_g)

    # ---- dolfin NewtonSolver (convergence_criterion "residual", relaxation 1) -------------------
    def _newton(self, residual, jacobian, x, atol, rtol, maxit, error_on_nonconvergence=True):
        """``residual()`` fills self._b from x and returns its l2 norm, ``jacobian()`` fills self._A.
        The update dx = J^-1 b is the K16 direct solve; x -= dx."""
        dev = self.dev
        r0 = r = residual()
        hist, it = [r], 0
        converged = (r / r0 < rtol if r0 > 0 else False) or r < atol
        while not converged and it < maxit:
            jacobian()
            if self._use_gmres():
                # forcing term for the FIRST update of a time step (same rule as ipcs.py: only when the predicted r1 --
                # the smaller of previous contraction x r0 and the previous step's r1 -- is ten times above the Newton
                # tolerance, so that a second, exact update is certain to follow and sets the accuracy)
                eta = None
                if it == 0 and self.newton_forcing > 0.0 and self._newton_rho is not None:
                    r1_pred = min(self._newton_rho * r0, self._newton_r1)
                    if r1_pred > 10.0 * max(atol, rtol * r0):
                        eta = min(1e-3, max(self._linear_rtol, self.newton_forcing * r1_pred / r0))
                self._gmres_update(r0, eta)
            else:
                eta = None
                dev.direct_solve("WW", self._A, self._b, self._dx, self._reorder)
            dev.axpy(-1.0, self._dx, x)
            r = residual()
            hist.append(r)
            it += 1
            if it == 1 and self.newton_forcing > 0.0:
                ok = r0 > 0.0 and (eta is None or r / r0 < 1e-2)
                self._newton_rho, self._newton_r1 = (r / r0, r) if ok else (None, None)
            converged = (r / r0 < rtol) or (r < atol)
        self.newton_history = hist
        if not converged and error_on_nonconvergence:
            raise NewtonNotConverged("Newton solver did not converge (%d iterations, residual %g)" % (it, r))
        return it

    def _prepare_forms_and_load(self):
        """weak-form switches to the device; external load (body force, tractions) scattered into W numbering"""
        dev = self.dev
        self._apply_forms(dev)
        load = self._external_load(dev)
        self._loadW = None
        if load is not None:
            if not hasattr(self, "_loadW_buf"):
                self._loadW_buf = dev.zeros(self._nW)
            dev.scatter(self._maps()[0], load, self._loadW_buf)
            self._loadW = self._loadW_buf

    # ---- K16: hand-written Krylov solve of the Newton update ---------------------------------------
    _mass_coefficient = 0.0      # alpha0 / k of the Jacobian's mass term; set by the instationary mirror

    def _use_gmres(self):
        if self._linear_solver == "auto":
            return self._nW >= 20000 and self._mass_coefficient > 0.0
        return self._linear_solver == "gmres"

    def _setup_saddle(self):
        """P1 mass / stiffness matrices, masks and work space of the block-triangular preconditioner (once)"""
        import ctypes as C
        dev, sp = self.dev, self.dev.spaces
        nu, npr = sp.dim * sp.n_nodes, sp.n_vertices
        pat = dev.pattern("P1P1")
        self._Mp_vals, self._Lp_vals = dev.empty(pat.nnz), dev.empty(pat.nnz)
        dev.assemble_p1_mass(L.MAT, A=self._Mp_vals)
        dev.assemble_poisson(L.MAT, A=self._Lp_vals)
        dofs = self._bc_dofs.cpu().numpy()
        ident = np.zeros(npr, dtype=np.uint8)
        ident[dofs[dofs >= nu] - nu] = 1                       # pressure Dirichlet dofs (and the pin): identity rows of A
        # natural (do-nothing) velocity boundaries: the Schur complement acts like a Dirichlet Laplacian there
        m = self._mesh
        with_bc = set(int(b) for b, _, _ in self._velocity_bcs)
        lmask = ident.copy()
        for marker in np.unique(m.bfacet_marker):
            if int(marker) not in with_bc:
                lmask[np.unique(m.bfacets[m.bfacet_marker == marker])] = 1
        self._p_ident = dev.upload(ident) if ident.any() else None
        self._lp_mask = dev.upload(lmask) if lmask.any() else None
        self._Mp = dev.matrix("P1P1", self._Mp_vals, L.CSR, 1, self._p_ident)
        self._Lp = dev.matrix("P1P1", self._Lp_vals, L.CSR, 1, self._lp_mask)
        self._Mp_dinv, self._Lp_dinv = dev.empty(npr), dev.empty(npr)
        dev.inv_diagonal(self._Mp, self._Mp_dinv)
        dev.inv_diagonal(self._Lp, self._Lp_dinv)
        lmax = 1.1 * dev.lambda_max(self._Lp, self._Lp_dinv, iters=30)
        so = self.saddle_options
        self._saddle_work = dev.empty(nu + 11 * npr)
        self._Amat = dev.matrix("WW", self._A, L.CSR, 1)
        self._A_dinv = dev.empty(self._nW)
        self._gmres_work = dev.empty(dev.lib.nsb_gmres_workspace_doubles(self._nW, self._gmres_restart, 1))
        co = self._equation_coefficients
        cp2 = co["pressure_term"] ** 2
        self._saddle = L.Saddle(nu, npr, self._Mp, self._Lp, self._Mp_dinv.data_ptr(), self._Lp_dinv.data_ptr(),
                                self._p_ident.data_ptr() if self._p_ident is not None else None,
                                co["viscous_term"] / cp2, 0.0, 0 if lmask.any() else 1, int(so["vel_degree"]), 0.0, 0.0,
                                int(so["lp_degree"]), lmax / so["lp_ratio"], lmax, float(so["inner_rtol"]), 5000, 0,
                                self._saddle_work.data_ptr(), 0)
        self._vel_lmax = None
        del C

    def _gmres_update(self, r0, eta=None):
        """dx = J^-1 b by flexible GMRES, right-preconditioned by the block-triangular Cahouet-Chabard operator;
        solved to ``linear_rtol`` relative to the step's initial Newton residual, like the IPCS momentum solve"""
        import ctypes as C
        dev = self.dev
        if not hasattr(self, "_saddle"):
            self._setup_saddle()
        S = self._saddle
        S.c_lap = self._mass_coefficient / self._equation_coefficients["pressure_term"] ** 2
        dev.inv_diagonal(self._Amat, self._A_dinv)
        if self._vel_lmax is None:
            self._vel_lmax = 1.1 * dev.saddle_lambda_max(self._Amat, self._A_dinv, S.n_u, iters=12, work=self._gmres_work)
        S.vel_lambda_max, S.vel_lambda_min = self._vel_lmax, self._vel_lmax / self.saddle_options["vel_ratio"]
        pc = L.Precond(L.PC_SADDLE, 1, self._A_dinv.data_ptr(), 0.0, 0.0, 0, 0, C.pointer(S))
        self._dx.zero_()
        t0 = S.inner_iterations
        # ||r|| <= max(rtol ||b||, 0.01 rtol r0): the first update to rtol, the later (small) ones two digits below the
        # step's initial residual -- the residual norm weighs velocity rows ~ h^2 alpha0/k against pressure rows ~ h,
        # and the pressure only reaches the 1e-8 parity bar with the last update solved that far
        rtol_ = self._linear_rtol if eta is None else eta
        n_it, _ = dev.gmres(self._Amat, pc, self._b, self._dx, rtol=rtol_, atol=(0.01 if eta is None else 1.0) * rtol_ * r0,
                            maxit=20 * self._gmres_restart, restart=self._gmres_restart, work=self._gmres_work, x0_is_zero=True)
        self.krylov_iterations.append(n_it)
        self.inner_iterations += S.inner_iterations - t0

    # ---- sub-space views of a mixed vector (FunctionAssigner, source/ns_solver_base.py:424-476) --
    def _maps(self):
        if not hasattr(self, "_v2w"):
            sp = self.dev.spaces
            self._v2w, self._q2w = self.dev.upload(sp.V_to_W()), self.dev.upload(sp.Q_to_W())
        return self._v2w, self._q2w

    def split(self, w):
        """(velocity in V numbering, pressure in Q numbering) of the mixed vector ``w``"""
        dev, sp = self.dev, self.dev.spaces
        v2w, q2w = self._maps()
        u, p = dev.empty(sp.ndofs_V), dev.empty(sp.ndofs_Q)
        dev.gather(v2w, w, u)
        dev.gather(q2w, w, p)
        return u, p

    def direct_solve_seconds(self, reset=False):
        """(seconds, solves) spent in the K16 library leg since the last reset"""
        return self.dev.direct_seconds(reset)

    # ---- drag / lift (a13) ---------------------------------------------------------------------
    def boundary_force(self, w, boundary_id, viscous_factor=1.0):
        """F = -int_Gamma ( -p n + c_visc * sym(grad u) n ) ds over the facets marked ``boundary_id``
        (demo/dfg_benchmark.py:44-66; drag = 2 F_x, lift = 2 F_y for the DFG 2D-2 scaling).  The reference's
        functional carries HALF the viscous stress of a Newtonian fluid (sigma = -p I + 2 c_visc sym(grad u));
        ``viscous_factor=2`` gives the physical traction, the one the published DFG coefficients refer to."""
        u, p = self.split(w)
        return boundary_force(self.dev, u, p, boundary_id, viscous_factor * self._equation_coefficients["viscous_term"])


def boundary_force(dev, u, p, boundary_id, c_visc):
    """Exterior-facet functional on the device (nsb_boundary_force): returns the d force components."""
    tab = dev.facet_table(boundary_id)
    return dev.boundary_force(tab, u, p, c_visc)
