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


@pytest.mark.parametrize("dim,n", [(2, (5, 4)), (3, (3, 3, 3))])
def test_scatter_schedule_invariants(built_library, dim, n):
    """nsb_sched_build (host-only; drives k_jacobian_rowgroups): every (pair, local column) contribution is scheduled exactly
    once, into its own entry or into one of that entry's shadow slots; no two lanes of a round share a slot; every shadow slot
    is folded into exactly one real entry of its group; groups respect the pair / slot limits"""
    import ctypes as C
    import nsb200
    from nsb200 import _lib as L
    from nsb200.spaces import TaylorHoodSpaces
    lib = built_library
    mesh = nsb200.hyper_rectangle((0.0,) * dim, (1.0,) * dim, n)[0]
    sp = TaylorHoodSpaces(mesh)
    cn = np.ascontiguousarray(sp.cell_nodes, dtype=np.int32)
    nloc = cn.shape[1]
    h = C.c_void_p()
    L.check(lib.nsb_pattern_build(sp.n_nodes, cn.shape[0], nloc, cn.ctypes.data, nloc, cn.ctypes.data, C.byref(h)))
    nnz, npairs, maxrow = C.c_int64(), C.c_int64(), C.c_int32()
    L.check(lib.nsb_pattern_sizes(h, C.byref(nnz), C.byref(npairs), C.byref(maxrow)))
    nrows = sp.n_nodes
    rowptr, col = np.empty(nrows + 1, np.int32), np.empty(nnz.value, np.int32)
    pair_ptr, pair_cell = np.empty(nrows + 1, np.int32), np.empty(npairs.value, np.int32)
    pair_iloc, pair_slots = np.empty(npairs.value, np.uint8), np.empty(npairs.value * nloc, np.uint8)
    L.check(lib.nsb_pattern_fill(h, rowptr.ctypes.data, col.ctypes.data, pair_ptr.ctypes.data, pair_cell.ctypes.data,
                                 pair_iloc.ctypes.data, pair_slots.ctypes.data))
    lib.nsb_pattern_free(h)
    cnt = (C.c_int64 * 4)()
    args = (nrows, rowptr.ctypes.data, pair_ptr.ctypes.data, pair_slots.ctypes.data, nloc, L.SCHED_MAX_BLOCKS)
    L.check(lib.nsb_sched_build(*args, cnt, None, None, None, None, None, None))
    ng, nc, nr, nf = (int(c) for c in cnt)
    g_row, c_first = np.empty(ng + 1, np.int32), np.empty(ng + 1, np.int32)
    r_ptr, sched = np.empty(nc + 1, np.int32), np.empty(32 * nr, np.uint16)
    f_ptr, fold = np.empty(ng + 1, np.int32), np.empty(max(nf, 1), np.uint32)
    L.check(lib.nsb_sched_build(*args, cnt, g_row.ctypes.data, c_first.ctypes.data, r_ptr.ctypes.data, sched.ctypes.data,
                                f_ptr.ctypes.data, fold.ctypes.data))
    assert g_row[0] == 0 and g_row[-1] == nrows and f_ptr[-1] == nf and r_ptr[-1] == nr and c_first[-1] == nc
    S = sched.reshape(nr, 32)
    seen_shadow_groups = 0
    for g in range(ng):
        r0, r1 = int(g_row[g]), int(g_row[g + 1])
        e_lo, nent = int(rowptr[r0]), int(rowptr[r1] - rowptr[r0])
        p_lo, p_hi = int(pair_ptr[r0]), int(pair_ptr[r1])
        nshadow = int(f_ptr[g + 1] - f_ptr[g])
        nslot = nent + nshadow
        assert nslot <= L.SCHED_MAX_BLOCKS or r1 == r0 + 1
        assert (p_hi - p_lo <= 32) or r1 == r0 + 1
        assert int(c_first[g + 1] - c_first[g]) == (p_hi - p_lo + 31) // 32
        # fold list: every shadow slot exactly once, target a real entry
        owner = {}
        for it in fold[f_ptr[g]:f_ptr[g + 1]]:
            real, shadow = int(it) >> 16, int(it) & 0xFFFF
            assert 0 <= real < nent and nent <= shadow < nslot and shadow not in owner
            owner[shadow] = real
        assert len(owner) == nshadow
        seen_shadow_groups += nshadow > 0
        # expected contributions: (pair, local column) -> real entry inside the group
        expected = {}
        row = r0
        for p in range(p_lo, p_hi):
            while pair_ptr[row + 1] <= p:
                row += 1
            for j in range(nloc):
                expected[(p, j)] = int(rowptr[row]) - e_lo + int(pair_slots[p * nloc + j])
        got = {}
        for ci, ch in enumerate(range(int(c_first[g]), int(c_first[g + 1]))):
            q0 = p_lo + 32 * ci
            for t in range(int(r_ptr[ch]), int(r_ptr[ch + 1])):
                busy = S[t] != 0xFFFF
                slots = (S[t][busy] >> 4).astype(int)
                assert len(np.unique(slots)) == len(slots), "two lanes of one round share a slot"
                for lane in np.nonzero(busy)[0]:
                    slot, j = int(S[t, lane]) >> 4, int(S[t, lane]) & 15
                    key = (q0 + int(lane), j)
                    assert key not in got and key in expected
                    got[key] = owner.get(slot, slot)
        assert got == expected
    assert seen_shadow_groups > 0 or dim == 2          # 3D vertex rows always have hot diagonals
