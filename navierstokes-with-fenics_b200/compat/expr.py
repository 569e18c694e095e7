"""Term-recording expression objects of the dolfin-named compatibility layer (SURVEY.md H1 option (ii)).

The reference solver classes build their weak forms symbolically (``dot(dot(grad(u), u), v) * dV``,
``dlfn.derivative(F, u)``, source/ns_ipcs_solver.py:126-136).  Nothing is compiled here: the objects below only RECORD
the expression tree; ``expand`` normalises an integrand into a sum of monomials

    numeric factor  x  prod Constant^power  x  one multilinear atom of Functions / test / trial functions

and ``recognise.py`` matches the monomials against the closed vocabulary of forms the CUDA kernels implement.  An
integrand outside that vocabulary raises ``NotImplementedError`` -- the layer is a drop-in for this path only.
"""
import numbers


class Expr:
    """base of everything that can appear in an integrand"""
    rank = 0          # tensor rank as far as it is needed to tell scalars from vectors

    def __add__(self, o):
        return Sum([self, as_expr(o)])

    def __radd__(self, o):
        return Sum([as_expr(o), self]) if not (isinstance(o, numbers.Number) and o == 0) else self

    def __sub__(self, o):
        return Sum([self, Scaled(-1.0, as_expr(o))])

    def __rsub__(self, o):
        return Sum([as_expr(o), Scaled(-1.0, self)])

    def __neg__(self):
        return Scaled(-1.0, self)

    def __mul__(self, o):
        if isinstance(o, Measure):
            return Form([Integral(self, o)])
        return Product(self, as_expr(o))

    def __rmul__(self, o):
        return Product(as_expr(o), self)

    def __truediv__(self, o):
        return Quotient(self, as_expr(o))

    def __getitem__(self, i):
        return Indexed(self, i)

    @property
    def T(self):
        return Transposed(self)


class Number(Expr):
    def __init__(self, v):
        self.v = float(v)


def as_expr(o):
    if isinstance(o, Expr):
        return o
    if isinstance(o, numbers.Number):
        return Number(o)
    raise TypeError("cannot use %r in a form" % (o,))


class Sum(Expr):
    def __init__(self, terms):
        self.terms = terms
        self.rank = terms[0].rank


class Scaled(Expr):
    def __init__(self, c, a):
        self.c, self.a, self.rank = float(c), a, a.rank


class Product(Expr):
    def __init__(self, a, b):
        self.a, self.b = a, b
        self.rank = a.rank + b.rank


class Quotient(Expr):
    def __init__(self, a, b):
        self.a, self.b, self.rank = a, b, a.rank


class Indexed(Expr):
    def __init__(self, a, i):
        self.a, self.i, self.rank = a, i, max(a.rank - 1, 0)


class Transposed(Expr):
    def __init__(self, a):
        self.a, self.rank = a, a.rank


class Op(Expr):
    """multilinear operator node: dot, inner, grad, div, curl, cross"""
    RANK = {"dot": lambda a, b: max(a + b - 2, 0), "inner": lambda a, b: 0, "cross": lambda a, b: 1}

    def __init__(self, name, *args):
        self.name, self.args = name, args
        if name == "grad":
            self.rank = args[0].rank + 1
        elif name == "div":
            self.rank = max(args[0].rank - 1, 0)
        elif name == "curl":
            self.rank = args[0].rank       # 3D; scalar in 2D -- not needed for the supported vocabulary
        else:
            self.rank = Op.RANK[name](args[0].rank, args[1].rank)


def dot(a, b):
    return Op("dot", as_expr(a), as_expr(b))


def inner(a, b):
    return Op("inner", as_expr(a), as_expr(b))


def grad(a):
    return Op("grad", as_expr(a))


def div(a):
    return Op("div", as_expr(a))


def curl(a):
    return Op("curl", as_expr(a))


def cross(a, b):
    return Op("cross", as_expr(a), as_expr(b))


class ListTensor(Expr):
    """``as_vector([...])`` / the parts of ``split(function)``"""
    rank = 1

    def __init__(self, components):
        self.components = list(components)


def as_vector(components):
    return ListTensor([as_expr(c) for c in components])


def sqrt(a):
    raise NotImplementedError("sqrt() only appears in post-processing forms, which are outside the hot path")


# ---- measures and forms ---------------------------------------------------------------------------
class Measure:
    def __init__(self, kind, domain=None, subdomain_data=None, subdomain_id=None, metadata=None):
        assert kind in ("dx", "ds", "dS")
        self.kind, self.domain, self.subdomain_data, self.subdomain_id, self.metadata = kind, domain, subdomain_data, subdomain_id, metadata

    def __call__(self, subdomain_id):
        return Measure(self.kind, self.domain, self.subdomain_data, subdomain_id, self.metadata)

    def __rmul__(self, integrand):
        return Form([Integral(as_expr(integrand), self)])


class Integral:
    def __init__(self, integrand, measure):
        self.integrand, self.measure = integrand, measure


class Form:
    """sum of integrals (``ufl.form.Form``)"""

    def __init__(self, integrals):
        self.integrals = list(integrals)

    def __add__(self, o):
        return Form(self.integrals + o.integrals)

    def __sub__(self, o):
        return Form(self.integrals + [Integral(Scaled(-1.0, i.integrand), i.measure) for i in o.integrals])

    def __neg__(self):
        return Form([Integral(Scaled(-1.0, i.integrand), i.measure) for i in self.integrals])

    def __radd__(self, o):
        assert o == 0
        return self


class Derivative:
    """``dlfn.derivative(F, u)``: the Gateaux derivative is never expanded -- the kernels assemble the exact Jacobian of
    the recognised residual; only the pairing (F, u) is checked"""

    def __init__(self, form, u):
        self.form, self.u = form, u


def derivative(form, u, du=None):
    assert isinstance(form, Form)
    return Derivative(form, u)


# ---- normalisation --------------------------------------------------------------------------------------
class Monomial:
    """coef * prod(const ** power) * atom ; atom = nested tuple (op name, children...) with leaves"""

    def __init__(self, coef, consts, atom):
        self.coef, self.consts, self.atom = coef, consts, atom      # consts: tuple of (Constant, power)

    def value(self):
        v = self.coef
        for c, pw in self.consts:
            v *= float(c) ** pw
        return v


def _is_constant_scalar(e):
    from .dolfin import Constant
    return isinstance(e, Constant) and e.rank == 0


def expand(e):
    """list of Monomials of the expression ``e`` (scalar or tensor valued); atoms of tensor-valued sub-expressions are
    combined by the multilinear operators above them"""
    from .dolfin import Constant, Function, Argument
    if isinstance(e, Number):
        return [Monomial(e.v, (), None)]
    if isinstance(e, Constant) and e.rank == 0:
        return [Monomial(1.0, ((e, 1),), None)]
    if isinstance(e, (Function, Argument, Constant)):
        return [Monomial(1.0, (), ("leaf", e))]
    if isinstance(e, Sum):
        out = []
        for t in e.terms:
            out += expand(t)
        return out
    if isinstance(e, Scaled):
        return [Monomial(e.c * m.coef, m.consts, m.atom) for m in expand(e.a)]
    if isinstance(e, Product):
        out = []
        for ma in expand(e.a):
            for mb in expand(e.b):
                if ma.atom is None:
                    atom = mb.atom
                elif mb.atom is None:
                    atom = ma.atom
                else:
                    atom = ("mul", ma.atom, mb.atom)
                out.append(Monomial(ma.coef * mb.coef, ma.consts + mb.consts, atom))
        return out
    if isinstance(e, Quotient):
        den = expand(e.b)
        if len(den) != 1 or den[0].atom is not None:
            raise NotImplementedError("division by a non-constant expression is outside the form vocabulary")
        d = den[0]
        inv = tuple((c, -pw) for c, pw in d.consts)
        return [Monomial(m.coef / d.coef, m.consts + inv, m.atom) for m in expand(e.a)]
    if isinstance(e, Op):
        parts = [expand(a) for a in e.args]
        out = []
        if len(parts) == 1:
            for m in parts[0]:
                if m.atom is None:
                    continue               # derivative of a constant
                out.append(Monomial(m.coef, m.consts, (e.name, m.atom)))
            return out
        for ma in parts[0]:
            for mb in parts[1]:
                if ma.atom is None or mb.atom is None:
                    raise NotImplementedError("%s() of a scalar constant" % e.name)
                out.append(Monomial(ma.coef * mb.coef, ma.consts + mb.consts, (e.name, ma.atom, mb.atom)))
        return out
    if isinstance(e, Transposed):
        return [Monomial(m.coef, m.consts, ("T", m.atom)) for m in expand(e.a)]
    if isinstance(e, Indexed):
        return [Monomial(m.coef, m.consts, ("index", m.atom, e.i)) for m in expand(e.a)]
    if isinstance(e, ListTensor):
        # only the trivial use is in the vocabulary: the velocity / pressure parts of a split mixed function stand for
        # the sub-function itself
        if getattr(e, "split_of", None) is not None:
            return [Monomial(1.0, (), ("leaf", e))]
        raise NotImplementedError("as_vector() expressions (rotational form in 2D) are outside the compat vocabulary; "
                                  "use the nsb200 mirror classes, which implement every weak-form variant")
    raise NotImplementedError("unsupported expression node %s" % type(e).__name__)


def atom_signature(atom, names):
    """canonical string of an atom with the leaves replaced through ``names(leaf) -> str``"""
    if atom is None:
        return "1"
    if atom[0] == "leaf":
        return names(atom[1])
    if atom[0] == "index":
        return "%s[%s]" % (atom_signature(atom[1], names), atom[2])
    return "%s(%s)" % (atom[0], ",".join(atom_signature(a, names) for a in atom[1:]))


def atom_leaves(atom, out=None):
    out = [] if out is None else out
    if atom is None:
        return out
    if atom[0] == "leaf":
        out.append(atom[1])
    elif atom[0] == "index":
        atom_leaves(atom[1], out)
    else:
        for a in atom[1:]:
            atom_leaves(a, out)
    return out
