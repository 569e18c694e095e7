"""GPU parity: every assembly kernel of libnsb200.so (through the C-ABI) against the CPU oracle
on the same seeded inputs.  Tolerance: 1e-12 relative to the largest entry (north star: assembled
matrix and vector entries agree to 1e-12 relative in FP64); patterns and maps bit-exact."""
import numpy as np
import pytest
import scipy.sparse as sp

from helpers import meshes, random_fields, rel_max_err, bsr_to_csr

pytestmark = pytest.mark.gpu
TOL = 1e-12
COEF = dict(k=1e-2, alpha=(1.5, -2.0, 0.5), c_conv=1.0, c_p=1.0, c_visc=1.0 / 50.0)


@pytest.fixture(scope="module", params=[0, 1], ids=["2d", "3d"])
def case(request, built_library):
    import nsb200
    import oracle
    from nsb200.device import Device
    name, m, om = meshes()[request.param]
    dev = Device(m)
    osp = oracle.dofmap.Spaces(om)
    asm = oracle.forms.Assembler(osp)
    dev.set_coefficients(COEF["k"], COEF["alpha"], COEF["c_conv"], COEF["c_p"], COEF["c_visc"])
    fields = random_fields(dev.spaces)
    return dict(dev=dev, asm=asm, osp=osp, fields=fields, d=m.dim)


def up(dev, a):
    return dev.upload(np.ascontiguousarray(a, dtype=np.float64))


def test_patterns_bit_exact(case):
    dev, asm = case["dev"], case["asm"]
    for name, oname in (("P2P2", None), ("P1P1", "QQ")):
        pat = dev.patterns[name]
        if oname is None:
            import oracle
            s = case["osp"]
            orp, oc = oracle.dofmap.sparsity_pattern(s.cell_nodes, s.cell_nodes, s.n_nodes, s.n_nodes)
        else:
            orp, oc, _ = asm.pattern(oname)
        assert np.array_equal(pat.rowptr_host, orp)
        assert np.array_equal(pat.col_host, oc)
    # expanded velocity pattern == dolfin's full d x d block pattern on V
    d = case["d"]
    pat = dev.patterns["P2P2"]
    A = bsr_to_csr(pat.rowptr_host, pat.col_host, np.ones(pat.nnz * d * d), d)
    A.sort_indices()
    orp, oc, _ = asm.pattern("VV")
    assert np.array_equal(A.indptr, orp) and np.array_equal(A.indices, oc)


def test_momentum_residual(case):
    import nsb200._lib as L
    dev, asm = case["dev"], case["asm"]
    us, un, um, p = case["fields"]
    b = dev.empty(len(us))
    dev.assemble_ipcs_momentum(L.VEC, up(dev, us), up(dev, un), up(dev, um), up(dev, p), b=b)
    ref = asm.ipcs_momentum_residual(us, un, um, p, COEF["alpha"], COEF["k"], COEF["c_conv"], COEF["c_p"],
                                     COEF["c_visc"])
    assert rel_max_err(b.cpu().numpy(), ref) < TOL


@pytest.mark.parametrize("newton", [True, False], ids=["newton", "picard"])
def test_momentum_jacobian(case, newton):
    import nsb200._lib as L
    dev, asm, d = case["dev"], case["asm"], case["d"]
    us = case["fields"][0]
    pat = dev.patterns["P2P2"]
    A = dev.empty(pat.nnz * d * d)
    dev.assemble_ipcs_momentum(L.MAT | (0 if newton else L.PICARD), up(dev, us), A=A)
    got = bsr_to_csr(pat.rowptr_host, pat.col_host, A.cpu().numpy(), d)
    ref = asm.ipcs_momentum_jacobian(us, COEF["alpha"][0], COEF["k"], COEF["c_conv"], COEF["c_visc"], newton=newton)
    diff = (got - ref).tocoo()
    assert np.abs(diff.data).max() / np.abs(ref.data).max() < TOL


def test_jacobian_is_derivative_of_residual(case):
    """size-independent property: J(u) v ~ (F(u + eps v) - F(u - eps v)) / (2 eps)."""
    import nsb200._lib as L
    dev, d = case["dev"], case["d"]
    us, un, um, p = case["fields"]
    rng = np.random.default_rng(7)
    v = rng.uniform(-1, 1, len(us))
    pat = dev.patterns["P2P2"]
    A = dev.empty(pat.nnz * d * d)
    dev.assemble_ipcs_momentum(L.MAT, up(dev, us), A=A)
    J = bsr_to_csr(pat.rowptr_host, pat.col_host, A.cpu().numpy(), d)
    eps = 1e-5
    bp, bm = dev.empty(len(us)), dev.empty(len(us))
    dun, dum, dp = up(dev, un), up(dev, um), up(dev, p)
    dev.assemble_ipcs_momentum(L.VEC, up(dev, us + eps * v), dun, dum, dp, b=bp)
    dev.assemble_ipcs_momentum(L.VEC, up(dev, us - eps * v), dun, dum, dp, b=bm)
    fd = (bp - bm).cpu().numpy() / (2 * eps)
    assert rel_max_err(J @ v, fd) < 1e-7


def test_poisson(case):
    import nsb200._lib as L
    dev, asm = case["dev"], case["asm"]
    us, _, _, p = case["fields"]
    pat = dev.patterns["P1P1"]
    A, b = dev.empty(pat.nnz), dev.empty(len(p))
    dev.assemble_poisson(L.MAT | L.VEC, pn=up(dev, p), ustar=up(dev, us), b=b, A=A)
    ref = asm.poisson_matrix()
    got = sp.csr_matrix((A.cpu().numpy(), pat.col_host, pat.rowptr_host), shape=ref.shape)
    assert np.abs((got - ref).tocoo().data).max() / np.abs(ref.data).max() < TOL
    refb = asm.poisson_rhs(p, us, COEF["alpha"][0], COEF["k"])
    assert rel_max_err(b.cpu().numpy(), refb) < TOL
    # stiffness annihilates constants
    assert np.abs(got @ np.ones(got.shape[0])).max() < 1e-12 * np.abs(ref.data).max()


def test_correction(case):
    import nsb200._lib as L
    dev, asm, d = case["dev"], case["asm"], case["d"]
    us, _, _, p = case["fields"]
    p2 = p[::-1].copy()
    pat = dev.patterns["P2P2"]
    A, b = dev.empty(pat.nnz), dev.empty(len(us))
    dev.assemble_correction(L.MAT | L.VEC, ustar=up(dev, us), pnew=up(dev, p), pold=up(dev, p2), b=b, A=A)
    n = pat.nrows
    M = sp.csr_matrix((A.cpu().numpy(), pat.col_host, pat.rowptr_host), shape=(n, n))
    got = sp.kron(M, sp.identity(d), format="csr")
    ref = asm.velocity_mass_matrix()
    assert np.abs((got - ref).tocoo().data).max() / np.abs(ref.data).max() < TOL
    refb = asm.correction_rhs(us, p, p2, COEF["alpha"][0], COEF["k"])
    assert rel_max_err(b.cpu().numpy(), refb) < TOL
    # mass matrix sums to the volume of the domain
    assert abs(M.sum() - asm.volume()) < 1e-12 * asm.volume()


def test_p1_mass_and_weights(case):
    import nsb200._lib as L
    dev, asm = case["dev"], case["asm"]
    pat = dev.patterns["P1P1"]
    A, w = dev.empty(pat.nnz), dev.empty(pat.nrows)
    dev.assemble_p1_mass(L.MAT | L.VEC, w=w, A=A)
    ref = asm.pressure_mass_matrix()
    got = sp.csr_matrix((A.cpu().numpy(), pat.col_host, pat.rowptr_host), shape=ref.shape)
    assert np.abs((got - ref).tocoo().data).max() / np.abs(ref.data).max() < TOL
    p = case["fields"][3]
    assert abs(float(w.cpu().numpy() @ p) - asm.integral_Q(p)) < 1e-12 * max(1.0, abs(asm.integral_Q(p)))
    assert abs(dev.sum(w) - asm.volume()) < 1e-12 * asm.volume()


def test_assembly_is_bit_reproducible(case):
    """write-once row-gather: two launches give identical bits (no atomics)."""
    import nsb200._lib as L
    dev, d = case["dev"], case["d"]
    us = up(dev, case["fields"][0])
    pat = dev.patterns["P2P2"]
    A1, A2 = dev.empty(pat.nnz * d * d), dev.empty(pat.nnz * d * d)
    dev.assemble_ipcs_momentum(L.MAT, us, A=A1)
    dev.assemble_ipcs_momentum(L.MAT, us, A=A2)
    assert bool((A1 == A2).all())
