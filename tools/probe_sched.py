"""Host-only study of the scatter schedule of the Jacobian assembly (nsb_sched_build): rounds per chunk and the
shared-memory bank conflicts of the scheduled read-modify-writes (8-byte words, 16 double banks per half warp:
lanes of a half warp whose entries agree mod 16 serialise)."""
import ctypes as C
import sys
import os
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nsb200
from nsb200 import _lib as L
from nsb200.spaces import TaylorHoodSpaces


def main(n=12):
    lib = L.load()
    mesh = nsb200.hyper_cube(3, n)[0]
    sp = TaylorHoodSpaces(mesh)
    cn = np.ascontiguousarray(sp.cell_nodes, dtype=np.int32)
    h = C.c_void_p()
    L.check(lib.nsb_pattern_build(sp.n_nodes, cn.shape[0], cn.shape[1], cn.ctypes.data, cn.shape[1], cn.ctypes.data, C.byref(h)))
    nnz, npairs, maxrow = C.c_int64(), C.c_int64(), C.c_int32()
    L.check(lib.nsb_pattern_sizes(h, C.byref(nnz), C.byref(npairs), C.byref(maxrow)))
    nrows, nlc = sp.n_nodes, cn.shape[1]
    rowptr = np.empty(nrows + 1, np.int32); col = np.empty(nnz.value, np.int32); pair_ptr = np.empty(nrows + 1, np.int32)
    pair_cell = np.empty(npairs.value, np.int32); pair_iloc = np.empty(npairs.value, np.uint8)
    pair_slots = np.empty(npairs.value * nlc, np.uint8)
    L.check(lib.nsb_pattern_fill(h, rowptr.ctypes.data, col.ctypes.data, pair_ptr.ctypes.data, pair_cell.ctypes.data,
                                 pair_iloc.ctypes.data, pair_slots.ctypes.data))
    lib.nsb_pattern_free(h)
    cnt = (C.c_int64 * 4)()
    a_ = (nrows, rowptr.ctypes.data, pair_ptr.ctypes.data, pair_slots.ctypes.data, nlc, L.SCHED_MAX_BLOCKS)
    L.check(lib.nsb_sched_build(*a_, cnt, None, None, None, None, None, None))
    g_row, c_first = np.empty(cnt[0] + 1, np.int32), np.empty(cnt[0] + 1, np.int32)
    r_ptr, sched = np.empty(cnt[1] + 1, np.int32), np.empty(32 * cnt[2], np.uint16)
    f_ptr, fold = np.empty(cnt[0] + 1, np.int32), np.empty(max(int(cnt[3]), 1), np.uint32)
    L.check(lib.nsb_sched_build(*a_, cnt, g_row.ctypes.data, c_first.ctypes.data, r_ptr.ctypes.data, sched.ctypes.data,
                                f_ptr.ctypes.data, fold.ctypes.data))
    S = sched.reshape(-1, 32)
    busy = S != 0xFFFF
    ent = (S >> 4).astype(np.int64)
    wave = 0
    for half in (slice(0, 16), slice(16, 32)):
        b = np.where(busy[:, half], ent[:, half] % 16, -1)
        mult = np.zeros(len(S), np.int64)
        for k in range(16):
            mult = np.maximum(mult, (b == k).sum(axis=1))
        wave += mult.sum()
    ideal = (busy[:, :16].any(axis=1).sum() + busy[:, 16:].any(axis=1).sum())
    print("n=%d pairs=%d groups=%d chunks=%d rounds=%d folds=%d rounds/chunk=%.2f rounds per 32 pairs=%.2f lane utilisation=%.3f" % (
        n, npairs.value, cnt[0], cnt[1], cnt[2], cnt[3], cnt[2] / cnt[1], 32.0 * cnt[2] / npairs.value, busy.mean()))
    print("RMW wavefronts per 32 pairs: %.1f (conflict-free at full lanes: %d)" % (32.0 * wave / npairs.value, 2 * nlc))
    print("wavefronts per 8-byte access: %.3f x the conflict-free count" % (wave / ideal))


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 12)
