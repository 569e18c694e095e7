"""Matching recorded forms against the closed vocabulary of this hot path ("each recognised term maps to a
kernel flag + a device scalar slot").  A form is expanded into monomials (``expr.expand``); every monomial must be one
of the terms listed here, with its Constants bound to a coefficient slot -- anything else raises NotImplementedError
with the offending term, so an edited reference form never silently runs a different equation.

IPCS tentative velocity   source/ns_ipcs_solver.py:126-136     F = sum_i alpha_i / k (u_i . w) + c_c (grad(u*) u*) . w
                                                                    - c_p div(w) p^n + c_nu grad(u*) : grad(w)
pressure projection       source/ns_ipcs_solver.py:160-164     a = grad(p) . grad(q) ;  L = grad(p^n) . grad(q) - alpha_0 / k div(u*) q
velocity correction       source/ns_ipcs_solver.py:183-187     a = v . w ;  L = u* . w - k / alpha_0 grad(p - p^n) . w
"""
from . import expr as X


class Unrecognised(NotImplementedError):
    pass


def _monomials(form):
    out = []
    for integral in form.integrals:
        if integral.measure.kind != "dx" or integral.measure.subdomain_id is not None:
            raise Unrecognised("only volume integrals are in the compat vocabulary (boundary tractions: use the nsb200 "
                               "mirror classes, which implement them)")
        out += X.expand(integral.integrand)
    return out


def _sig(m, roles):
    def names(leaf):
        from .dolfin import Argument
        if isinstance(leaf, Argument):
            return "w" if leaf.number == 0 else "v"
        for name, obj in roles.items():
            if obj is leaf:
                return name
        return "?%s" % type(leaf).__name__
    return X.atom_signature(m.atom, names)


def _functions(monos):
    from .dolfin import Function
    seen = []
    for m in monos:
        for leaf in X.atom_leaves(m.atom):
            if isinstance(leaf, Function) and all(leaf is not s for s in seen):
                seen.append(leaf)
    return seen


def recognise_ipcs_diffusion(F, u, J):
    """returns dict(ustar, history [(Function, coefficient monomial)], p_old, c_conv, c_p, c_visc) with the coefficients
    as zero-argument callables (the Constants may be re-assigned between steps)"""
    if not (isinstance(J, X.Derivative) and J.form is F and J.u is u):
        raise Unrecognised("the Jacobian must be dlfn.derivative(F, u) of the same residual (Newton linearisation)")
    monos = _monomials(F)
    funcs = _functions(monos)
    p_old = [f for f in funcs if f.function_space().kind == "Q"]
    vel = [f for f in funcs if f.function_space().kind == "V"]
    if len(p_old) != 1 or u not in vel:
        raise Unrecognised("tentative-velocity residual: expected the unknown velocity and exactly one pressure function")
    roles = {"u": u, "p": p_old[0]}
    for i, f in enumerate(v for v in vel if v is not u):
        roles["h%d" % i] = f
    out = dict(ustar=u, p_old=p_old[0], acc={}, c_conv=None, c_p=None, c_visc=None)
    for m in monos:
        s = _sig(m, roles)
        if s == "dot(u,w)" or (s.startswith("dot(h") and s.endswith(",w)")):
            key = "u" if s == "dot(u,w)" else s[4:s.index(",")]
            out["acc"].setdefault(key, []).append(m)
        elif s == "dot(dot(grad(u),u),w)" and out["c_conv"] is None:
            out["c_conv"] = m
        elif s in ("mul(div(w),p)", "mul(p,div(w))") and out["c_p"] is None:
            out["c_p"] = m
        elif s == "inner(grad(u),grad(w))" and out["c_visc"] is None:
            out["c_visc"] = m
        else:
            raise Unrecognised("term %s of the tentative-velocity residual is not in the compat vocabulary (standard "
                               "convective form, reduced viscous form); the nsb200 mirror classes cover the other "
                               "weak-form variants" % s)
    if None in (out["c_conv"], out["c_p"], out["c_visc"]) or "u" not in out["acc"]:
        raise Unrecognised("tentative-velocity residual: a term of the momentum equation is missing")
    out["history"] = {k: roles[k] for k in out["acc"] if k != "u"}
    return out


def recognise_ipcs_projection(a, L, u):
    ma, mL = _monomials(a), _monomials(L)
    if len(ma) != 1 or _sig(ma[0], {}) != "dot(grad(v),grad(w))" or abs(ma[0].value() - 1.0) > 0:
        raise Unrecognised("pressure projection: the bilinear form must be dot(grad(p), grad(q)) * dx")
    funcs = _functions(mL)
    pq = [f for f in funcs if f.function_space().kind == "Q"]
    uv = [f for f in funcs if f.function_space().kind == "V"]
    if len(pq) != 1 or len(uv) != 1:
        raise Unrecognised("pressure projection: expected the old pressure and the tentative velocity on the right-hand side")
    roles = {"p": pq[0], "u": uv[0]}
    out = dict(p_new=u, p_old=pq[0], ustar=uv[0], c_grad=None, c_div=None)
    for m in mL:
        s = _sig(m, roles)
        if s == "dot(grad(p),grad(w))" and out["c_grad"] is None:
            out["c_grad"] = m
        elif s in ("dot(div(u),w)", "mul(div(u),w)") and out["c_div"] is None:
            out["c_div"] = m
        else:
            raise Unrecognised("term %s of the pressure-projection right-hand side is not in the compat vocabulary" % s)
    if None in (out["c_grad"], out["c_div"]):
        raise Unrecognised("pressure projection: a right-hand-side term is missing")
    return out


def recognise_ipcs_correction(a, L, u):
    ma, mL = _monomials(a), _monomials(L)
    if len(ma) != 1 or _sig(ma[0], {}) != "dot(v,w)" or abs(ma[0].value() - 1.0) > 0:
        raise Unrecognised("velocity correction: the bilinear form must be dot(v, w) * dx")
    funcs = _functions(mL)
    uv = [f for f in funcs if f.function_space().kind == "V"]
    pq = [f for f in funcs if f.function_space().kind == "Q"]
    if len(uv) != 1 or len(pq) != 2:
        raise Unrecognised("velocity correction: expected the tentative velocity, the new and the old pressure")
    out = dict(u_new=u, ustar=uv[0], c_u=None, grads=[])
    roles = {"u": uv[0], "p0": pq[0], "p1": pq[1]}
    for m in mL:
        s = _sig(m, roles)
        if s == "dot(u,w)" and out["c_u"] is None:
            out["c_u"] = m
        elif s in ("dot(grad(p0),w)", "dot(grad(p1),w)"):
            out["grads"].append((roles[s[9:11]], m))
        else:
            raise Unrecognised("term %s of the velocity-correction right-hand side is not in the compat vocabulary" % s)
    if out["c_u"] is None or len(out["grads"]) != 2:
        raise Unrecognised("velocity correction: a right-hand-side term is missing")
    return out


# ---- monolithic problems on the mixed space W (source/ns_bdf_solver.py:36-100, source/ns_solver_base.py:894-949) -----
#   F = -c_p div(u) q + sum_i alpha_i / k (u_i . w) + c_c (grad(u) u) . w - c_p div(w) p + c_nu grad(u) : grad(w)
#   J_newton = derivative(F, w_h) ;  J_picard = -c_p div(v) q + c_c (grad(v) u) . w - c_p div(w) r + c_nu grad(v) : grad(w)
def _sig_mixed(m, roles):
    """signature of a monomial whose leaves are parts of ``dlfn.split`` (ListTensor with split_of / index): the parts of
    the test function are w / q, of the trial function v / r, of the functions in ``roles`` <name>u / <name>p"""
    def names(leaf):
        from .dolfin import Argument
        parent = getattr(leaf, "split_of", None)
        if parent is None:
            return "?%s" % type(leaf).__name__
        if isinstance(parent, Argument):
            return (("w", "q") if parent.number == 0 else ("v", "r"))[leaf.index]
        for name, obj in roles.items():
            if obj is parent:
                return name + ("u", "p")[leaf.index]
        return "?Function"
    return X.atom_signature(m.atom, names)


def _split_parents(monos):
    from .dolfin import Function
    seen = []
    for m in monos:
        for leaf in X.atom_leaves(m.atom):
            parent = getattr(leaf, "split_of", None)
            if isinstance(parent, Function) and all(parent is not s for s in seen):
                seen.append(parent)
    return seen


def recognise_mixed_residual(F, w):
    """returns dict(solution, history {h1: Function, ...}, acc {"s" | "h<i>": [monomials alpha_i / k]} (empty for the
    stationary problem), c_div, c_conv, c_p, c_visc as monomials)"""
    monos = _monomials(F)
    parents = _split_parents(monos)
    if all(p is not w for p in parents):
        raise Unrecognised("mixed residual: the unknown does not appear in the form")
    roles = {"s": w}
    for i, f in enumerate(p for p in parents if p is not w):
        roles["h%d" % (i + 1)] = f
    out = dict(solution=w, acc={}, c_div=None, c_conv=None, c_p=None, c_visc=None)
    for m in monos:
        s = _sig_mixed(m, roles)
        if s.startswith("dot(") and s.endswith("u,w)") and s.count("(") == 1:
            out["acc"].setdefault(s[4:-4], []).append(m)
        elif s in ("mul(div(su),q)", "mul(q,div(su))") and out["c_div"] is None:
            out["c_div"] = m
        elif s == "dot(dot(grad(su),su),w)" and out["c_conv"] is None:
            out["c_conv"] = m
        elif s in ("mul(div(w),sp)", "mul(sp,div(w))") and out["c_p"] is None:
            out["c_p"] = m
        elif s == "inner(grad(su),grad(w))" and out["c_visc"] is None:
            out["c_visc"] = m
        else:
            raise Unrecognised("term %s of the mixed residual is not in the compat vocabulary (standard convective form, "
                               "reduced viscous form, no body force / tractions); the nsb200 mirror classes cover the "
                               "other weak-form variants" % s)
    if None in (out["c_div"], out["c_conv"], out["c_p"], out["c_visc"]):
        raise Unrecognised("mixed residual: a term of the mass / momentum balance is missing")
    if out["acc"] and "s" not in out["acc"]:
        raise Unrecognised("mixed residual: the acceleration term does not contain the unknown velocity")
    out["history"] = {k: roles[k] for k in out["acc"] if k != "s"}
    return out


def recognise_mixed_jacobian(J, F, w):
    """"newton" for dlfn.derivative(F, w), "picard" for the frozen-convection bilinear form of
    source/ns_solver_base.py:929-934 (its coefficients must be the residual's)"""
    if isinstance(J, X.Derivative):
        if J.form is not F or J.u is not w:
            raise Unrecognised("the Newton Jacobian must be dlfn.derivative(F, w) of the same residual")
        return "newton"
    res = recognise_mixed_residual(F, w)
    roles = {"s": w}
    seen = dict(c_div=None, c_conv=None, c_p=None, c_visc=None)
    for m in _monomials(J):
        s = _sig_mixed(m, roles)
        key = {"mul(div(v),q)": "c_div", "mul(q,div(v))": "c_div", "dot(dot(grad(v),su),w)": "c_conv",
               "mul(div(w),r)": "c_p", "mul(r,div(w))": "c_p", "inner(grad(v),grad(w))": "c_visc"}.get(s)
        if key is None or seen[key] is not None:
            raise Unrecognised("term %s of the Picard Jacobian is not in the compat vocabulary" % s)
        seen[key] = m
    for key, m in seen.items():
        if m is None or abs(m.value() - res[key].value()) > 1e-14 * abs(res[key].value()):
            raise Unrecognised("the Picard Jacobian's %s does not match the residual's" % key)
    return "picard"
