"""C-ABI boundary on a CPU-only box: libnsb200.so builds for sm_100a, loads, exports every symbol
include/nsb200.h declares, reports errors through status codes, has NO CPU fallback, and its
host-side pattern builder reproduces the oracle's sparsity patterns bit for bit."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "nsb200.h")) as f:
        text = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(nsb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported_and_bound(built_library):
    import nsb200
    names = declared_symbols()
    assert len(names) >= 35
    for name in names:
        assert hasattr(built_library, name), "libnsb200.so does not export %s" % name
    assert sorted(nsb200._lib.PROTOTYPES) == names, "ctypes prototypes and include/nsb200.h disagree"
    assert built_library.nsb_version() == 100


def test_no_cpu_fallback(built_library):
    import torch
    import nsb200
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    ctx = C.c_void_p()
    assert built_library.nsb_create(C.byref(ctx), 0) != 0
    assert "no CPU fallback" in nsb200._lib.last_error()
    mesh, markers = nsb200.hyper_cube(2, 2)
    ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=0.1)
    with pytest.raises(RuntimeError):
        nsb200.IPCSSolver(mesh, markers, "standard", ts)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "navierstokes-with-fenics_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", src, flags=re.M), fn
                assert "nsoracle" not in src, fn


@pytest.mark.parametrize("dim", [2, 3])
def test_pattern_builder_matches_oracle_bit_exact(built_library, dim):
    """nsb_pattern_build / _fill (host code, no GPU): rowptr, col, incidence and slot maps."""
    import nsb200
    import oracle
    L = nsb200._lib
    box = ((0.0, 0.0), (2.0, 1.0), (5, 3)) if dim == 2 else ((0.0, 0.0, 0.0), (1.0, 1.0, 0.5), (3, 2, 2))
    mesh, _ = nsb200.hyper_rectangle(*box)
    sp = nsb200.TaylorHoodSpaces(mesh)
    osp = oracle.dofmap.Spaces(oracle.mesh.hyper_rectangle(*box))
    for cd, n, ocd in ((sp.cell_nodes, sp.n_nodes, osp.cell_nodes), (mesh.cells, sp.n_vertices, osp.cell_dofs_Q),
                       (sp.cell_dofs_W, sp.ndofs_W, osp.cell_dofs_W)):
        assert np.array_equal(cd, ocd)                      # dof maps bit-exact
        cd = np.ascontiguousarray(cd, dtype=np.int32)
        nloc = cd.shape[1]
        h = C.c_void_p()
        L.check(built_library.nsb_pattern_build(n, cd.shape[0], nloc, cd.ctypes.data, nloc, cd.ctypes.data, C.byref(h)))
        nnz, npairs, mr = C.c_int64(), C.c_int64(), C.c_int32()
        L.check(built_library.nsb_pattern_sizes(h, C.byref(nnz), C.byref(npairs), C.byref(mr)))
        rowptr, col = np.empty(n + 1, np.int32), np.empty(nnz.value, np.int32)
        pp, pc = np.empty(n + 1, np.int32), np.empty(npairs.value, np.int32)
        pi, ps = np.empty(npairs.value, np.uint8), np.empty(npairs.value * nloc, np.uint8)
        L.check(built_library.nsb_pattern_fill(h, rowptr.ctypes.data, col.ctypes.data, pp.ctypes.data,
                                               pc.ctypes.data, pi.ctypes.data, ps.ctypes.data))
        built_library.nsb_pattern_free(h)
        orp, oc = oracle.dofmap.sparsity_pattern(ocd, ocd, n, n)
        assert np.array_equal(rowptr, orp) and np.array_equal(col, oc)
        assert npairs.value == cd.size and mr.value == np.diff(rowptr).max()
        # incidence: pair p of row r is (cell, iloc) with cd[cell, iloc] == r, cells ascending
        rows = np.repeat(np.arange(n), np.diff(pp))
        assert np.array_equal(cd[pc, pi], rows)
        assert all(np.all(np.diff(pc[pp[r]:pp[r + 1]]) >= 0) for r in range(n))
        # slots: col[rowptr[r] + slot] is the global column of local column j
        slots = ps.reshape(-1, nloc).astype(np.int64)
        assert np.array_equal(col[rowptr[rows][:, None] + slots], cd[pc])


def test_pattern_builder_rejects_bad_input(built_library):
    import nsb200
    h = C.c_void_p()
    cd = np.array([[0, 1, 5]], dtype=np.int32)
    assert built_library.nsb_pattern_build(3, 1, 3, cd.ctypes.data, 3, cd.ctypes.data, C.byref(h)) != 0
    assert "out of range" in nsb200._lib.last_error()
    with pytest.raises(RuntimeError):
        nsb200._lib.check(-1)
