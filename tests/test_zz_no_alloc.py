"""The allocation-free residual entry point nsb_assemble_ipcs_momentum_residual_noalloc (include/nsb200.h; no counterpart in
the reference): same output as nsb_assemble_ipcs_momentum(NSB_VEC) on two
calls in a row with different inputs, no device memory allocated by the call, meshes beyond the static scratch refused.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from helpers import meshes, random_fields, rel_max_err

pytestmark = pytest.mark.gpu
TOL = 1e-12          # tests/test_gpu_assembly.py: assembled entries agree to 1e-12 relative


def _ptr(t):
    return C.c_void_p(t.data_ptr())


@pytest.mark.parametrize("which", [0, 1], ids=["2d", "3d"])
def test_noalloc_residual_equals_the_existing_entry_point_twice(built_library, which):
    import torch
    import nsb200._lib as L
    from nsb200.device import Device
    lib = built_library
    _, m, _ = meshes()[which]
    dev = Device(m)
    dev.set_coefficients(1e-2, (1.5, -2.0, 0.5), 1.0, 1.0, 1.0 / 50.0)
    sp = dev.spaces
    assert m.num_cells * sp.cell_nodes.shape[1] * m.dim <= lib.nsb_static_scratch_doubles()
    nV = sp.ndofs_V
    b_ref, b_new = dev.empty(nV), dev.empty(nV)
    for call, seed in enumerate((1234, 99)):            # two calls in a row, different inputs: stale scratch would show
        us, un, unm1, p = (dev.upload(np.ascontiguousarray(a)) for a in random_fields(sp, seed=seed))
        dev.assemble_ipcs_momentum(L.VEC, us, un, unm1, p, b=b_ref)
        b_new.fill_(float("nan"))
        torch.cuda.synchronize()
        free0, alloc0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_allocated()
        st = lib.nsb_assemble_ipcs_momentum_residual_noalloc(dev.ctx, _ptr(us), _ptr(un), _ptr(unm1), _ptr(p), _ptr(b_new),
                                                             dev.stream)
        dev.check(st)
        torch.cuda.synchronize()
        if call == 1:                                   # the first call may load the module lazily; the second may not move a byte
            assert torch.cuda.mem_get_info()[0] == free0
        assert torch.cuda.memory_allocated() == alloc0
        got, ref = b_new.cpu().numpy(), b_ref.cpu().numpy()
        assert np.isfinite(got).all()
        assert rel_max_err(got, ref) < TOL, (call, rel_max_err(got, ref))


def test_noalloc_residual_refuses_what_does_not_fit(built_library):
    import nsb200
    from nsb200.device import Device
    lib = built_library
    n = 23                                              # 23^3 x 6 = 73 002 tetrahedra > 69 905
    m, _ = nsb200.hyper_cube(3, n)
    assert m.num_cells * 30 > lib.nsb_static_scratch_doubles()
    dev = Device(m)
    dev.set_coefficients(1e-2, (1.5, -2.0, 0.5), 1.0, 1.0, 0.02)
    sp = dev.spaces
    u, p, b = dev.zeros(sp.ndofs_V), dev.zeros(sp.ndofs_Q), dev.zeros(sp.ndofs_V)
    st = lib.nsb_assemble_ipcs_momentum_residual_noalloc(dev.ctx, _ptr(u), _ptr(u), _ptr(u), _ptr(p), _ptr(b), dev.stream)
    assert st != 0
    with pytest.raises(RuntimeError, match="too large for the allocation-free residual"):
        dev.check(st)
    # a non-default weak form is refused as well
    m2 = meshes()[0][1]
    dev2 = Device(m2)
    dev2.set_coefficients(1e-2, (1.5, -2.0, 0.5), 1.0, 1.0, 0.02)
    dev2.set_forms(1, 0, None)
    s2 = dev2.spaces
    u, p, b = dev2.zeros(s2.ndofs_V), dev2.zeros(s2.ndofs_Q), dev2.zeros(s2.ndofs_V)
    st = lib.nsb_assemble_ipcs_momentum_residual_noalloc(dev2.ctx, _ptr(u), _ptr(u), _ptr(u), _ptr(p), _ptr(b), dev2.stream)
    assert st != 0
