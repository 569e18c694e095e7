#!/usr/bin/env python3
"""Gather locality of the P2 node numbering (CPU study for the L1TEX-sector-bound preconditioner SpMV):
distinct 32-byte sectors touched by the float4 gathers of one row / of one warp (8 rows), for the
reference numbering (vertices, then edges) and for a spatially sorted (Morton) numbering.
usage: probe_numbering.py [n]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nsb200                                  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
mesh, _ = nsb200.hyper_cube(3, n)
sp = nsb200.TaylorHoodSpaces(mesh)
cn = sp.cell_nodes
nn = sp.n_nodes
# scalar P2 pattern (CSR) from the cell -> node map
rows = np.repeat(cn, cn.shape[1], axis=1).reshape(-1)
cols = np.tile(cn, (1, cn.shape[1])).reshape(-1)
key = np.unique(rows.astype(np.int64) * nn + cols)
r, c = key // nn, key % nn
rowptr = np.searchsorted(r, np.arange(nn + 1))
print("nodes %d, entries %d, %.1f per row" % (nn, len(c), len(c) / nn))


def morton(x, bits=10):
    q = np.clip((x - x.min(0)) / (np.ptp(x, 0) + 1e-300) * (2 ** bits - 1), 0, 2 ** bits - 1).astype(np.uint64)
    code = np.zeros(len(x), dtype=np.uint64)
    for b in range(bits):
        for d in range(x.shape[1]):
            code |= ((q[:, d] >> np.uint64(b)) & np.uint64(1)) << np.uint64(b * x.shape[1] + d)
    return code


def sectors(perm, label):
    """perm[old] = new.  Rows are visited in NEW order, 8 consecutive rows share a warp (4 lanes per row)."""
    inv = np.argsort(perm)
    newc = perm[c]
    per_row, per_warp = [], []
    for w0 in range(0, nn - 8, max(1, nn // 4000) * 8):       # sample of warps
        secs = []
        for new_row in range(w0, w0 + 8):
            old = inv[new_row]
            s = np.unique(newc[rowptr[old]:rowptr[old + 1]] // 2)   # float4 = 16 B: two nodes per 32-B sector
            per_row.append(len(s))
            secs.append(s)
        per_warp.append(len(np.unique(np.concatenate(secs))))
    ent = len(c) / nn
    print("%-28s sectors per row %.1f (%.2f per entry), distinct per 8-row warp %.1f (%.2f per entry)"
          % (label, np.mean(per_row), np.mean(per_row) / ent, np.mean(per_warp), np.mean(per_warp) / (8 * ent)))


x = sp.node_coordinates()
sectors(np.arange(nn), "reference (verts, edges)")
sectors(np.argsort(np.argsort(morton(x), kind="stable")), "Morton order of coordinates")
lex = np.lexsort((x[:, 0], x[:, 1], x[:, 2]))
sectors(np.argsort(lex), "lexicographic (z, y, x)")
