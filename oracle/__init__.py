"""CPU oracle for the nsb200 hot path -- TEST INFRASTRUCTURE, NOT PRODUCT.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this package.  The product
(``navierstokes-with-fenics_b200``) never imports it and has no CPU fallback.

What it is
----------
A numpy/scipy restatement of what the reference
(LKM-code-base/NavierStokes-with-Fenics, ``/root/reference``) asks dolfin
2019.1.0 / FFC / PETSc-LU to compute for the P2-P1 Navier-Stokes solvers:

* meshes + facet markers of ``source/grid_generator.py:111-208``
* Taylor-Hood / collapsed sub-space dof maps and sparsity patterns
  (``source/ns_solver_base.py:501-524``, ``:401-422``)
* the weak forms of ``source/ns_ipcs_solver.py:106-196``,
  ``source/ns_bdf_solver.py:19-100`` and ``source/ns_solver_base.py:894-949``
* dolfin's DirichletBC / Newton / LU semantics
* the step drivers ``source/ns_ipcs_solver.py:198-208``,
  ``source/ns_bdf_solver.py:102-106``, ``source/ns_solver_base.py:951-988``

PARITY UNPINNED (against FEniCS itself)
---------------------------------------
The arithmetic of the path lives in third-party packages that are neither in
``/root/reference`` nor installable here (fenics-dolfin/ffc/fiat/ufl 2019.1.0,
petsc 3.14.5, suitesparse 5.6.0; ``environment.yml:19-25,75,99``) and the
reference's own tests pin no number at this boundary
(``tests/test_ipcs_solver.py:52-54`` only runs to completion).  The oracle is
therefore pinned to what *can* be pinned here:

* the reference's pure-python time stepping (imported unchanged from
  ``/root/reference/source`` when present; golden tables of
  ``tests/test_bdf_time_stepping.py:67-114`` committed under ``tests/golden``)
* exact element integrals (sympy), patch tests, the Poiseuille steady state
  (``tests/test_stationary_solvers.py:173-175``) and the Taylor-Green exact
  solution (``convergence_test/taylor_green_vortex.py:111-117``)
* the size formulas of SURVEY.md App. C.

Design rule: the oracle is deliberately *generic and slow* (collapsed
Gauss-Jacobi quadrature of generous degree, einsum over cells, topology by
sorting), so that it shares no closed-form tables or code with the CUDA path it
checks.
"""

from . import mesh, fem, dofmap, forms, bc, solvers, native, functionals, projection  # noqa: F401
