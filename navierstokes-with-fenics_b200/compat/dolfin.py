"""``dolfin``-named compatibility module: the subset of the dolfin 2019.1 Python API that the reference's solver
classes call, backed by libnsb200.so.  ``compat.install()`` registers it as ``sys.modules["dolfin"]``
so that ``source/ns_ipcs_solver.py``, ``ns_solver_base.py``, ``auxiliary_classes.py``, ``auxiliary_methods.py`` and
``grid_generator.py`` of the reference import and run UNCHANGED on top of it.

What it is: meshes / facet markers (structured generators of source/grid_generator.py:111-208), Taylor-Hood function
spaces and their sub-space algebra, ``Function`` / ``Constant`` / ``Expression`` / ``DirichletBC`` /
``FunctionAssigner`` objects, term-recording forms (``compat/expr.py``) and the variational problem / solver classes.
``{Non,}LinearVariationalSolver.solve()`` recognises the three IPCS sub-problems (``compat/recognise.py``) and runs them
through the same device code as the nsb200 mirror class.  What it is not: a finite-element library -- a form outside
the recognised vocabulary raises ``NotImplementedError``; XDMF output, mshr, DG post-processing projections are stubs.
"""
import math
import re

import numpy as np

from .. import grid as _grid
from ..solver_base import Expression as _NsbExpression
from ..spaces import TaylorHoodSpaces
from . import expr as _x
from .expr import (dot, inner, grad, div, curl, cross, as_vector, sqrt, derivative, Measure, Form)   # noqa: F401

DOLFIN_EPS = _grid.DOLFIN_EPS
_BACKEND = {"device_factory": None}


def set_log_level(level):
    pass


def info(msg):
    pass


class _MPI:
    comm_world = "comm_world"


MPI = _MPI()


# =====================================================================================================
# meshes and markers
# =====================================================================================================
class Point:
    def __init__(self, *xs):
        self._x = np.array([float(v) for v in xs])

    def __getitem__(self, i):
        return float(self._x[i])

    def dot(self, o):
        return float(self._x @ o._x)

    def __sub__(self, o):
        return Point(*(self._x - o._x))

    def array(self):
        return self._x


class _Geometry:
    def __init__(self, m):
        self._m = m

    def dim(self):
        return self._m.dim


class Mesh:
    """wraps an ordered simplicial mesh (nsb200.grid.SimplexMesh) and owns the per-mesh Device"""

    def __init__(self, simplex_mesh=None):
        if simplex_mesh is None:
            raise NotImplementedError("reading meshes from files is outside the compat layer (no gmsh / meshio here)")
        self._m = simplex_mesh
        self._device = None
        self._engine = None
        self._spaces = None

    def geometry(self):
        return _Geometry(self._m)

    def topology(self):
        return _Geometry(self._m)

    def num_cells(self):
        return self._m.num_cells

    def num_vertices(self):
        return self._m.num_vertices

    def coordinates(self):
        return self._m.coords

    def ufl_cell(self):
        return "triangle" if self._m.dim == 2 else "tetrahedron"

    # ---- backend -----------------------------------------------------------------------------
    def spaces(self):
        if self._spaces is None:
            self._spaces = self._dev().spaces if self._device is not None else TaylorHoodSpaces(self._m)
        return self._spaces

    def _dev(self):
        """the Device of this mesh (created on first use; tests install a recording fake through set_backend)"""
        if self._device is None:
            factory = _BACKEND["device_factory"]
            if factory is None:
                from ..device import Device
                self._device = Device(self._m)
            else:
                self._device = factory(self._m)
            self._spaces = self._device.spaces
        return self._device


def RectangleMesh(p0, p1, nx, ny, diagonal="right"):
    assert diagonal == "right"
    return Mesh(_grid._structured(tuple(p0.array()), tuple(p1.array()), (int(nx), int(ny)))._unmarked())


def BoxMesh(p0, p1, nx, ny, nz):
    return Mesh(_grid._structured(tuple(p0.array()), tuple(p1.array()), (int(nx), int(ny), int(nz)))._unmarked())


class _Facet:
    def __init__(self, mesh, k):
        self._mesh, self._k = mesh, k

    def exterior(self):
        return True

    def index(self):
        return self._k

    def midpoint(self):
        m = self._mesh._m
        return Point(*m.coords[m.bfacets[self._k]].mean(axis=0))

    def normal(self):
        m = self._mesh._m
        fv = m.bfacets[self._k]
        X = m.coords[fv]
        if m.dim == 2:
            t = X[1] - X[0]
            n = np.array([t[1], -t[0]])
        else:
            n = np.cross(X[1] - X[0], X[2] - X[0])
        n = n / np.linalg.norm(n)
        cell = m.cells[m.bfacet_cell[self._k]]
        opp = [v for v in cell if v not in fv][0]
        if n @ (m.coords[opp] - X[0]) > 0.0:
            n = -n
        return Point(*n)


def facets(mesh):
    """exterior facets only: interior facets carry no marker in this layer and every reference loop over
    ``dlfn.facets`` acts on exterior ones (source/auxiliary_methods.py:24-66)"""
    for k in range(len(mesh._m.bfacets)):
        yield _Facet(mesh, k)


def vertices(mesh):
    raise NotImplementedError("vertex iterators are only used by the gmsh import path of grid_generator.py")


class _MeshFunctionSizet:
    """``MeshFunction("size_t", mesh, dim - 1)``: the facet markers; stored on the boundary facets of the mesh"""

    def __init__(self, mesh, dim):
        assert dim == mesh._m.dim - 1, "only facet markers are supported"
        self._mesh, self._dim = mesh, dim

    def dim(self):
        return self._dim

    def set_all(self, v):
        self._mesh._m.bfacet_marker[:] = int(v)

    def __getitem__(self, f):
        k = f.index() if isinstance(f, _Facet) else int(f)
        return int(self._mesh._m.bfacet_marker[k])

    def array(self):
        return self._mesh._m.bfacet_marker


class _MeshFunctionInt(_MeshFunctionSizet):
    pass


def MeshFunction(kind, mesh, dim, value=None):
    mf = (_MeshFunctionSizet if kind == "size_t" else _MeshFunctionInt)(mesh, dim)
    if value is not None:
        mf.set_all(value)
    return mf


class SubDomain:
    def inside(self, x, on_boundary):
        raise NotImplementedError

    def mark(self, mesh_function, marker):
        """dolfin: a facet is marked when all its vertices and its midpoint are inside"""
        m = mesh_function._mesh._m
        X = m.coords[m.bfacets]                                     # [nf, d, dim]
        pts = np.concatenate([X, X.mean(axis=1, keepdims=True)], axis=1)
        for k in range(len(X)):
            if all(self.inside(p, True) for p in pts[k]):
                m.bfacet_marker[k] = int(marker)


def _cpp_to_python(code):
    code = code.replace("&&", " and ").replace("||", " or ").replace("!", " not ").replace(" not =", "!=")
    code = re.sub(r"\bstd::", "", code)
    return code.replace("fabs", "abs")


_MATH = {"sin": np.sin, "cos": np.cos, "tan": np.tan, "exp": np.exp, "sqrt": np.sqrt, "pow": np.power, "log": np.log,
         "abs": np.abs, "pi": math.pi, "DOLFIN_EPS": DOLFIN_EPS, "DOLFIN_PI": math.pi, "M_PI": math.pi, "atan2": np.arctan2,
         "tanh": np.tanh, "cosh": np.cosh, "sinh": np.sinh, "near": lambda a, b, eps=DOLFIN_EPS: abs(a - b) < eps}


class CompiledSubDomain(SubDomain):
    def __init__(self, code, **params):
        self._code = compile(_cpp_to_python(code), "<CompiledSubDomain>", "eval")
        self._params = params

    def inside(self, x, on_boundary):
        env = dict(_MATH, x=x, on_boundary=on_boundary, **self._params)
        return bool(eval(self._code, {"__builtins__": {}}, env))


class MeshValueCollection:
    def __init__(self, *a, **k):
        raise NotImplementedError("gmsh / XDMF mesh import is outside the compat layer")


class XDMFFile:
    def __init__(self, *a, **k):
        raise NotImplementedError("XDMF output is outside the hot path (SURVEY.md 2: out of scope)")


File = XDMFFile


# =====================================================================================================
# elements and function spaces
# =====================================================================================================
class FiniteElement:
    def __init__(self, family, cell, degree):
        self.family_name, self.cell, self.deg, self.kind = family, cell, degree, "scalar"

    def degree(self):
        return self.deg

    def family(self):
        return "Lagrange" if self.family_name in ("CG", "Lagrange", "P") else self.family_name


class VectorElement(FiniteElement):
    def __init__(self, family, cell, degree, dim=None):
        super().__init__(family, cell, degree)
        self.kind = "vector"


class MixedElement:
    def __init__(self, elements):
        self.elements = list(elements)
        self.kind = "mixed"


class FunctionSpace:
    """The Taylor-Hood family of this path: ``kind`` "W" (mixed [P2]^d x P1), "V" (collapsed [P2]^d, interleaved),
    "Q" (collapsed P1); ``path`` the chain of ``.sub(i)`` indices of a sub-space view."""

    def __init__(self, mesh, element, constrained_domain=None, _kind=None, _path=()):
        if constrained_domain is not None:
            raise NotImplementedError("periodic constraints are not on the GPU path (SURVEY.md 8f-4)")
        assert isinstance(mesh, Mesh)
        self._mesh, self._element, self._path = mesh, element, tuple(_path)
        if _kind is None:
            if isinstance(element, MixedElement):
                ev, ep = element.elements
                if not (ev.kind == "vector" and ev.deg == 2 and ep.kind == "scalar" and ep.deg == 1
                        and ev.family() == ep.family() == "Lagrange"):
                    raise NotImplementedError("only the P2-P1 Taylor-Hood element is on the GPU path")
                _kind = "W"
            elif element.kind == "vector" and element.deg == 2 and element.family() == "Lagrange":
                _kind = "V"
            elif element.kind == "scalar" and element.deg == 1 and element.family() == "Lagrange":
                _kind = "Q"
            else:
                raise NotImplementedError("function space outside the Taylor-Hood family (%s, degree %s): DG / higher "
                                          "order spaces are only used by post-processing" % (element.kind, element.deg))
        self.kind = _kind

    def mesh(self):
        return self._mesh

    def ufl_element(self):
        return self._element

    def dim(self):
        sp = self._mesh.spaces()
        if self._path:
            raise NotImplementedError("dimension of a sub-space view")
        return {"W": sp.ndofs_W, "V": sp.ndofs_V, "Q": sp.ndofs_Q}[self.kind]

    def sub(self, i):
        return FunctionSpace(self._mesh, self._element, _kind=self.kind, _path=self._path + (int(i),))

    def num_sub_spaces(self):
        return 2 if self.kind == "W" and not self._path else self._mesh._m.dim

    def collapse(self):
        assert self.kind == "W" and len(self._path) == 1
        el = self._element.elements[self._path[0]]
        return FunctionSpace(self._mesh, el, _kind="V" if self._path[0] == 0 else "Q")

    def key(self):
        return (id(self._mesh), self.kind, self._path)

    def __eq__(self, o):
        return isinstance(o, FunctionSpace) and self.key() == o.key()

    def __hash__(self):
        return hash(self.key())

    def __contains__(self, f):
        return isinstance(f, Function) and f.function_space() == self


# =====================================================================================================
# coefficients
# =====================================================================================================
class Constant(_x.Expr):
    def __init__(self, value, name=None, cell=None):
        self._name = name
        self._set(value)

    def _set(self, value):
        if isinstance(value, Constant):
            value = value._v
        if isinstance(value, (tuple, list, np.ndarray)):
            self._v = tuple(float(v) for v in value)
            self.rank = 1
        else:
            self._v = float(value)
            self.rank = 0

    def assign(self, value):
        self._set(value)

    def values(self):
        return np.atleast_1d(np.array(self._v, dtype=np.float64))

    def __float__(self):
        assert self.rank == 0
        return self._v

    def value_rank(self):
        return self.rank


class Expression(_NsbExpression, _x.Expr):
    """``dlfn.Expression(cpp_code | (cpp_code, ...), degree=k, **parameters)``: the C++ strings are evaluated with numpy
    (``x[i]`` -> coordinate column, math functions of <cmath>); parameters stay assignable attributes
    (``expression.time = t``, source/ns_solver_base.py:1045-1051)."""

    def __init__(self, code, degree=2, element=None, **params):
        object.__setattr__(self, "_params", dict(params))
        codes = (code,) if isinstance(code, str) else tuple(code)
        object.__setattr__(self, "_codes", [compile(_cpp_to_python(c), "<Expression>", "eval") for c in codes])
        object.__setattr__(self, "rank", 0 if isinstance(code, str) else 1)
        object.__setattr__(self, "degree", int(degree))
        object.__setattr__(self, "fn", None)

    def __getattr__(self, name):
        params = object.__getattribute__(self, "_params")
        if name in params:
            return params[name]
        raise AttributeError(name)

    def __setattr__(self, name, value):
        if name in self._params:
            self._params[name] = value
        else:
            object.__setattr__(self, name, value)

    def value_rank(self):
        return self.rank

    def __call__(self, X):
        X = np.asarray(X, dtype=np.float64)
        env = dict(_MATH, x=[X[:, i] for i in range(X.shape[1])], **self._params)
        cols = [np.broadcast_to(np.asarray(eval(c, {"__builtins__": {}}, env), dtype=np.float64), (len(X),)) for c in self._codes]
        return cols[0].copy() if self.rank == 0 else np.stack(cols, axis=1)


class _Vector:
    """``Function.vector()`` / ``dlfn.Vector``: thin view of the device tensor"""

    def __init__(self, function):
        self._f = function

    def norm(self, kind="l2"):
        dev = self._f._dev()
        t = self._f._tensor
        if kind == "l2":
            return dev.norm(t)
        if kind == "linf":
            return float(t.abs().max())
        raise NotImplementedError(kind)

    def get_local(self):
        return self._f._tensor.cpu().numpy().copy()

    def set_local(self, a):
        self._f._tensor.copy_(self._f._dev().upload(np.asarray(a, dtype=np.float64)))

    def size(self):
        return self._f._tensor.numel()

    def zero(self):
        self._f._tensor.zero_()

    def __getitem__(self, i):
        return self.get_local()[i]


class _RawVectorOwner:
    """owner of a bare device tensor that lives on the space of ``like`` (the residual vector of a NewtonSolver)"""

    def __init__(self, tensor, like):
        self._tensor, self._like = tensor, like

    def _dev(self):
        return self._like._dev()

    def function_space(self):
        return self._like.function_space()


def Vector(v):
    assert isinstance(v, _Vector)
    g = Function(v._f.function_space())
    g.assign(v._f)
    return g.vector()


class Function(_x.Expr):
    """a dof vector on the device in the numbering of nsb200.spaces (V interleaved, Q, W component-blocked)"""

    def __init__(self, space, name=None, _tensor=None, _parent=None):
        assert isinstance(space, FunctionSpace)
        self._space, self._name, self._parent = space, name, _parent
        self.rank = 1 if (space.kind == "V" or (space.kind == "W" and space._path == (0,))) else 0
        if _parent is None:
            assert not space._path, "functions live on whole (or collapsed) spaces"
            self._tensor = _tensor if _tensor is not None else space.mesh()._dev().zeros(space.dim())

    def _dev(self):
        return self._space.mesh()._dev()

    def function_space(self):
        return self._space

    def ufl_element(self):
        el = self._space._element
        return el.elements[self._space._path[0]] if self._space._path and isinstance(el, MixedElement) else el

    def name(self):
        return self._name

    def rename(self, name, label=""):
        self._name = name

    def vector(self):
        assert self._parent is None
        return _Vector(self)

    def assign(self, other):
        assert isinstance(other, Function) and other.function_space() == self._space and self._parent is None
        self._dev().copy(other._tensor, self._tensor)

    def split(self, deepcopy=False):
        assert self._space.kind == "W" and not self._space._path and not deepcopy
        return tuple(Function(self._space.sub(i), _parent=self) for i in range(2))

    def sub(self, i):
        return self.split()[i]


def split(f):
    """``dlfn.split(w)`` for use inside forms: parts that stand for the velocity / pressure of the mixed function"""
    assert isinstance(f, (Function, Argument)) and f.function_space().kind == "W"
    parts = []
    for i in range(2):
        lt = _x.ListTensor([])
        lt.split_of, lt.index, lt.rank = f, i, 1 - i
        parts.append(lt)
    return tuple(parts)


class Argument(_x.Expr):
    def __init__(self, space, number):
        self._space, self.number = space, number
        self.rank = 1 if space.kind in ("V",) else 0

    def function_space(self):
        return self._space


def TestFunction(space):
    return Argument(space, 0)


def TrialFunction(space):
    return Argument(space, 1)


def TestFunctions(space):
    return split(Argument(space, 0))


def TrialFunctions(space):
    return split(Argument(space, 1))


def SpatialCoordinate(mesh):
    raise NotImplementedError("SpatialCoordinate only appears in the Euler-acceleration term (time-dependent rotation)")


def FacetNormal(mesh):
    raise NotImplementedError("facet normals in forms: post-processing only (drag / lift is nsb_boundary_force)")


def CellDiameter(mesh):
    raise NotImplementedError("use IPCSSolver.cfl_number (nsb_cfl_max) for the CFL diagnostic")


# =====================================================================================================
# boundary conditions, assigners, projection, scalar assembly
# =====================================================================================================
class DirichletBC:
    def __init__(self, space, value, markers, marker_id, method="topological"):
        assert isinstance(space, FunctionSpace) and isinstance(markers, _MeshFunctionSizet)
        self.space, self.value, self.markers, self.marker_id = space, value, markers, int(marker_id)

    def as_tuple(self):
        """(boundary id, component | None | "p", value) in the format of nsb200.spaces.dirichlet_table"""
        sp = self.space
        comp = None
        if sp.kind == "Q" or (sp.kind == "W" and sp._path[:1] == (1,)):
            comp = "p"
        elif sp.kind == "V" and len(sp._path) == 1:
            comp = sp._path[0]
        elif sp.kind == "W" and len(sp._path) == 2:
            comp = sp._path[1]
        v = self.value
        if isinstance(v, Constant):
            v = v._v
        return (self.marker_id, comp, v)


class FunctionAssigner:
    """mixed <-> collapsed sub-space copies (source/ns_solver_base.py:424-476): gather / scatter with the V->W, Q->W maps"""

    def __init__(self, receiving, assigning):
        self._recv, self._assg = receiving, assigning

    @staticmethod
    def _map(dev, kind):
        if not hasattr(dev, "_compat_maps"):
            sp = dev.spaces
            dev._compat_maps = {"V": dev.upload(sp.V_to_W()), "Q": dev.upload(sp.Q_to_W())}
        return dev._compat_maps[kind]

    def assign(self, receiving, assigning):
        recv = receiving if isinstance(receiving, (list, tuple)) else [receiving]
        assg = assigning if isinstance(assigning, (list, tuple)) else [assigning]
        if len(recv) == 1 and len(assg) > 1:          # [V, Q] -> W
            w = recv[0]
            for f in assg:
                dev = f._dev()
                dev.scatter(self._map(dev, f.function_space().kind), f._tensor, w._tensor)
        elif len(assg) == 1 and len(recv) > 1:        # W -> [V, Q]
            w = assg[0]
            for f in recv:
                dev = f._dev()
                dev.gather(self._map(dev, f.function_space().kind), w._tensor, f._tensor)
        else:
            r, a = recv[0], assg[0]
            dev = r._dev()
            if a._parent is not None:                 # W.sub(i) view -> collapsed function
                dev.gather(self._map(dev, r.function_space().kind), a._parent._tensor, r._tensor)
            elif r._parent is not None:               # collapsed function -> W.sub(i) view
                dev.scatter(self._map(dev, a.function_space().kind), a._tensor, r._parent._tensor)
            else:
                r.assign(a)


def project(expression, space, **kwargs):
    """``dlfn.project`` onto the collapsed velocity / pressure space (source/ns_solver_base.py:1151,1168,1201)"""
    assert space.kind in ("V", "Q") and not space._path
    dev = space.mesh()._dev()
    out = Function(space)
    d = dev.dim
    if isinstance(expression, Constant):
        vals = expression.values()
        n = out._tensor.numel()
        out._tensor.copy_(dev.upload(np.tile(vals, n // len(vals))))
    elif isinstance(expression, Expression):
        dev.project(expression, space.kind, expression.degree, out=out._tensor)
    else:
        raise NotImplementedError("project() of a general expression: only Constants and Expressions are on the hot path")
    del d
    return out


def assemble(form):
    """scalar functionals of the mean-pressure fix (source/ns_solver_base.py:1194): ``p * dx`` and ``Constant * dx``"""
    assert isinstance(form, Form) and len(form.integrals) == 1 and form.integrals[0].measure.kind == "dx"
    mono = _x.expand(form.integrals[0].integrand)
    assert len(mono) == 1
    m = mono[0]
    mesh = form.integrals[0].measure.domain
    if m.atom is None:
        zero = Function(FunctionSpace(mesh, FiniteElement("CG", mesh.ufl_cell(), 1)))
        return m.value() * mesh._dev().p1_integral(zero._tensor)[1]
    leaves = _x.atom_leaves(m.atom)
    if m.atom[0] == "leaf" and isinstance(leaves[0], Function):
        f = leaves[0]
        if f._parent is not None and f.function_space()._path == (1,):
            sp = f._dev().spaces
            p = f._parent._tensor[sp.dim * sp.n_nodes:]
            return m.value() * f._dev().p1_integral(p)[0]
        if f.function_space().kind == "Q":
            return m.value() * f._dev().p1_integral(f._tensor)[0]
    raise NotImplementedError("assemble() of this scalar form is outside the compat vocabulary")


def norm(v, kind="l2"):
    return v.norm(kind)


def solve(*a, **k):
    raise NotImplementedError("dlfn.solve: use the variational solver classes")


class LocalSolver:
    def __init__(self, *a, **k):
        raise NotImplementedError("LocalSolver (CFL diagnostic): use IPCSSolver.cfl_number / nsb_cfl_max")


# =====================================================================================================
# variational problems and solvers
# =====================================================================================================
class _Parameters(dict):
    def __missing__(self, key):
        self[key] = _Parameters()
        return self[key]


class NonlinearVariationalProblem:
    def __init__(self, F, u, bcs=None, J=None):
        assert isinstance(F, Form) and isinstance(u, Function)
        self.F, self.u, self.bcs, self.J = F, u, list(bcs or []), J


class NonlinearVariationalSolver:
    def __init__(self, problem):
        self.problem = problem
        self.parameters = _Parameters()
        ns = self.parameters["newton_solver"]
        ns.update(absolute_tolerance=1e-10, relative_tolerance=1e-9, maximum_iterations=50, error_on_nonconvergence=True)
        from .backend import bind_nonlinear
        self._run = bind_nonlinear(problem)

    def solve(self):
        return self._run(self.parameters["newton_solver"])


class LinearVariationalProblem:
    def __init__(self, a, L, u, bcs=None):
        assert isinstance(a, Form) and isinstance(L, Form) and isinstance(u, Function)
        self.a, self.L, self.u, self.bcs = a, L, u, list(bcs or [])


class LinearVariationalSolver:
    def __init__(self, problem):
        self.problem = problem
        self.parameters = _Parameters()
        from .backend import bind_linear
        self._run = bind_linear(problem)

    def solve(self):
        return self._run()


class NonlinearProblem:
    """base class of ``CustomNonlinearProblem`` (source/auxiliary_classes.py:120-164)"""

    def __init__(self):
        pass


class SystemAssembler:
    """``SystemAssembler(J, F, bcs)`` of source/auxiliary_classes.py:141: bound to the stationary mixed problem"""

    def __init__(self, J, F, bcs):
        from .backend import StationaryBinding
        self._binding = StationaryBinding(J, F, bcs)

    def assemble(self, *args):
        from .backend import _MatrixHandle
        if len(args) == 2 and all(isinstance(a, _Vector) for a in args):
            b, x = args
            self._binding.assemble_vector(b._f._tensor, x._f._tensor)
        elif len(args) == 1 and isinstance(args[0], _MatrixHandle):
            self._binding.assemble_matrix()
        else:
            raise NotImplementedError("SystemAssembler.assemble(b, x) and .assemble(A) are the calls of this path")


class NewtonSolver:
    """``dlfn.NewtonSolver(comm, linear_solver, factory)`` with a ``NonlinearProblem`` (source/ns_solver_base.py:938-983):
    dolfin's loop (residual criterion), the problem's ``F(b, x)`` / ``J(A, x)`` called back every iteration, the update
    solved on the device (K16)"""

    def __init__(self, *a, **k):
        self.parameters = _Parameters()
        self.parameters.update(absolute_tolerance=1e-10, relative_tolerance=1e-9, maximum_iterations=50,
                               error_on_nonconvergence=True)

    def solve(self, problem, x):
        binding = getattr(getattr(problem, "assembler", None), "_binding", None)
        if binding is None or not isinstance(x, _Vector):
            raise NotImplementedError("NewtonSolver.solve: the problem must assemble through a SystemAssembler")
        return binding.newton(problem, x, self.parameters)


class PETScLUSolver:
    def __init__(self, *a, **k):
        pass


class PETScFactory:
    @staticmethod
    def instance():
        return None
