"""Drop-in boundary as BASELINE.json's north star words it: the reference's own solver classes on top of libnsb200.

``install()`` registers the dolfin-named module of this package (and the few ``ufl`` type names the reference checks
with ``isinstance``) in ``sys.modules``; afterwards ``source/ns_ipcs_solver.py`` / ``ns_solver_base.py`` /
``auxiliary_*.py`` / ``grid_generator.py`` / ``bdf_time_stepping.py`` of the reference import and run unchanged, with
every per-step floating-point operation in the CUDA library.  Scope: the instationary IPCS path with the default weak
forms; everything else the reference can do is served by the nsb200 mirror classes (same names and semantics), and a
form outside the recognised vocabulary raises ``NotImplementedError`` instead of computing something else.
"""
import sys
import types


def install(device_factory=None):
    """register ``dolfin`` / ``ufl``; ``device_factory(simplex_mesh) -> Device``: replace the CUDA device (tests)"""
    from . import dolfin as D
    from . import expr as X
    D._BACKEND["device_factory"] = device_factory
    sys.modules["dolfin"] = D
    fn = types.ModuleType("dolfin.function")
    arg = types.ModuleType("dolfin.function.argument")
    arg.Argument = D.Argument
    fun = types.ModuleType("dolfin.function.function")
    fun.Function = D.Function
    fn.argument, fn.function = arg, fun
    D.function = fn
    cpp = types.ModuleType("dolfin.cpp")
    cmesh = types.ModuleType("dolfin.cpp.mesh")
    cmesh.MeshFunctionSizet, cmesh.MeshFunctionInt = D._MeshFunctionSizet, D._MeshFunctionInt
    cpp.mesh = cmesh
    D.cpp = cpp
    for name, mod in (("dolfin.function", fn), ("dolfin.function.argument", arg), ("dolfin.function.function", fun),
                      ("dolfin.cpp", cpp), ("dolfin.cpp.mesh", cmesh)):
        sys.modules[name] = mod
    ufl = types.ModuleType("ufl")
    ufl.form = types.ModuleType("ufl.form")
    ufl.form.Form = X.Form
    ufl.tensors = types.ModuleType("ufl.tensors")
    ufl.tensors.ListTensor = X.ListTensor
    ufl.indexed = types.ModuleType("ufl.indexed")
    ufl.indexed.Indexed = X.Indexed
    for name, mod in (("ufl", ufl), ("ufl.form", ufl.form), ("ufl.tensors", ufl.tensors), ("ufl.indexed", ufl.indexed)):
        sys.modules[name] = mod
    # grid_generator.py imports mshr at module level (annulus meshes): present as a stub, unusable
    if "mshr" not in sys.modules:
        mshr = types.ModuleType("mshr")

        def _absent(*a, **k):
            raise NotImplementedError("mshr is not available: only the structured generators of grid_generator.py "
                                      "(hyper_cube, hyper_rectangle, open_hyper_cube) work on the compat layer")
        mshr.Sphere = mshr.Circle = mshr.generate_mesh = _absent
        mshr._nsb_stub = True
        sys.modules["mshr"] = mshr
    return D


def uninstall():
    if getattr(sys.modules.get("mshr"), "_nsb_stub", False):
        del sys.modules["mshr"]
    for name in [n for n in sys.modules if n == "dolfin" or n.startswith("dolfin.") or n == "ufl" or n.startswith("ufl.")]:
        del sys.modules[name]
