// setup_host.cu -- CPU-side setup helpers of libnsb200 (no kernels): sparsity pattern and the
// row -> (cell, local row, column slots) incidence that drives the row-gather assembly.
//
// Replaces dolfin's SparsityPatternBuilder / PETSc AIJ preallocation that the reference
// triggers implicitly through its function spaces (source/ns_solver_base.py:501-524) and the
// first assemble() of each form.  Pattern semantics: union over cells of
// row dofs x col dofs of the whole element, ascending int32 columns, no extra diagonal.
#include <algorithm>
#include <vector>
#include "common.cuh"

thread_local std::string g_nsb_last_error;

struct nsb_pattern_builder {
    int32_t nrows = 0, ncells = 0, nlr = 0, nlc = 0;
    const int32_t* cell_rows = nullptr;   // borrowed, must stay alive until _fill
    const int32_t* cell_cols = nullptr;
    std::vector<int32_t> pair_ptr;        // [nrows+1]
    std::vector<int32_t> pair_cell;       // [npairs]
    std::vector<uint8_t> pair_iloc;       // [npairs]
    std::vector<int32_t> rowlen;          // [nrows]
    int64_t nnz = 0;
    int32_t max_row = 0;
};

static int row_columns(const nsb_pattern_builder* pb, int32_t row, std::vector<int32_t>& buf) {
    buf.clear();
    for (int32_t p = pb->pair_ptr[row]; p < pb->pair_ptr[row + 1]; ++p) {
        const int32_t* cc = pb->cell_cols + (size_t)pb->pair_cell[p] * pb->nlc;
        buf.insert(buf.end(), cc, cc + pb->nlc);
    }
    std::sort(buf.begin(), buf.end());
    buf.erase(std::unique(buf.begin(), buf.end()), buf.end());
    return (int)buf.size();
}

extern "C" int nsb_pattern_build(int32_t nrows, int32_t ncells, int32_t nloc_row,
                                 const int32_t* cell_rows, int32_t nloc_col,
                                 const int32_t* cell_cols, nsb_pattern_builder** out) {
    if (!out || !cell_rows || !cell_cols || nrows <= 0 || ncells <= 0 || nloc_row <= 0 ||
        nloc_col <= 0 || nloc_row > 255)
        return nsb_fail(nullptr, "nsb_pattern_build: bad arguments%s%s");
    auto* pb = new nsb_pattern_builder();
    pb->nrows = nrows; pb->ncells = ncells; pb->nlr = nloc_row; pb->nlc = nloc_col;
    pb->cell_rows = cell_rows; pb->cell_cols = cell_cols;
    const int64_t npairs = (int64_t)ncells * nloc_row;
    if (npairs >= INT32_MAX) { delete pb; return nsb_fail(nullptr, "nsb_pattern_build: too many (row,cell) pairs%s%s"); }
    pb->pair_ptr.assign((size_t)nrows + 1, 0);
    for (int64_t i = 0; i < npairs; ++i) {
        int32_t r = cell_rows[i];
        if (r < 0 || r >= nrows) { delete pb; return nsb_fail(nullptr, "nsb_pattern_build: row dof out of range%s%s"); }
        pb->pair_ptr[(size_t)r + 1]++;
    }
    for (int32_t r = 0; r < nrows; ++r) pb->pair_ptr[r + 1] += pb->pair_ptr[r];
    pb->pair_cell.resize((size_t)npairs);
    pb->pair_iloc.resize((size_t)npairs);
    {   // cells ascending inside each row: the summation order of dolfin's cell loop (App. A.4)
        std::vector<int32_t> cur(pb->pair_ptr.begin(), pb->pair_ptr.end() - 1);
        for (int32_t c = 0; c < ncells; ++c)
            for (int32_t i = 0; i < nloc_row; ++i) {
                int32_t r = cell_rows[(size_t)c * nloc_row + i];
                int32_t p = cur[r]++;
                pb->pair_cell[p] = c;
                pb->pair_iloc[p] = (uint8_t)i;
            }
    }
    pb->rowlen.assign((size_t)nrows, 0);
    int32_t maxrow = 0;
    int64_t nnz = 0;
#pragma omp parallel reduction(max : maxrow) reduction(+ : nnz)
    {
        std::vector<int32_t> buf;
#pragma omp for schedule(static, 4096)
        for (int32_t r = 0; r < nrows; ++r) {
            int n = row_columns(pb, r, buf);
            pb->rowlen[r] = n;
            nnz += n;
            if (n > maxrow) maxrow = n;
        }
    }
    pb->nnz = nnz;
    pb->max_row = maxrow;
    if (nnz >= INT32_MAX) { delete pb; return nsb_fail(nullptr, "nsb_pattern_build: nnz exceeds int32%s%s"); }
    if (maxrow > 256) { delete pb; return nsb_fail(nullptr, "nsb_pattern_build: row longer than 256 entries (uint8 slot map)%s%s"); }
    *out = pb;
    return 0;
}

extern "C" int nsb_pattern_sizes(const nsb_pattern_builder* pb, int64_t* nnz, int64_t* npairs,
                                 int32_t* max_row_len) {
    if (!pb) return nsb_fail(nullptr, "nsb_pattern_sizes: null builder%s%s");
    if (nnz) *nnz = pb->nnz;
    if (npairs) *npairs = (int64_t)pb->pair_cell.size();
    if (max_row_len) *max_row_len = pb->max_row;
    return 0;
}

extern "C" int nsb_pattern_fill(const nsb_pattern_builder* pb, int32_t* rowptr, int32_t* col,
                                int32_t* pair_ptr, int32_t* pair_cell, uint8_t* pair_iloc,
                                uint8_t* pair_slots) {
    if (!pb || !rowptr || !col) return nsb_fail(nullptr, "nsb_pattern_fill: null argument%s%s");
    rowptr[0] = 0;
    for (int32_t r = 0; r < pb->nrows; ++r) rowptr[r + 1] = rowptr[r] + pb->rowlen[r];
    if (pair_ptr) memcpy(pair_ptr, pb->pair_ptr.data(), sizeof(int32_t) * ((size_t)pb->nrows + 1));
    if (pair_cell) memcpy(pair_cell, pb->pair_cell.data(), sizeof(int32_t) * pb->pair_cell.size());
    if (pair_iloc) memcpy(pair_iloc, pb->pair_iloc.data(), pb->pair_iloc.size());
#pragma omp parallel
    {
        std::vector<int32_t> buf;
#pragma omp for schedule(static, 4096)
        for (int32_t r = 0; r < pb->nrows; ++r) {
            int n = row_columns(pb, r, buf);
            int32_t* dst = col + rowptr[r];
            for (int i = 0; i < n; ++i) dst[i] = buf[i];
            if (pair_slots) {
                for (int32_t p = pb->pair_ptr[r]; p < pb->pair_ptr[r + 1]; ++p) {
                    const int32_t* cc = pb->cell_cols + (size_t)pb->pair_cell[p] * pb->nlc;
                    uint8_t* s = pair_slots + (size_t)p * pb->nlc;
                    for (int j = 0; j < pb->nlc; ++j)
                        s[j] = (uint8_t)(std::lower_bound(buf.begin(), buf.end(), cc[j]) - buf.begin());
                }
            }
        }
    }
    return 0;
}

extern "C" int nsb_pattern_free(nsb_pattern_builder* pb) {
    delete pb;
    return 0;
}

// ---- TMA row tiles ------------------------------------------------------------------------------
// Partition of the rows of a CSR/BSR pattern into tiles whose value and column streams are one
// contiguous, 16-byte aligned chunk each, so that a CTA can fetch a whole tile with three bulk
// copies (spmv_tma.cuh).  Record = 64 uint32 (256 B):
//   [0] first row  [1] rows  [2] first entry (= rowptr[first row])  [3] entries
//   then uint16 off[rows + 1]: entry offset of every row inside the tile (off[rows] = entries).
// The kernels fetch the 4-entry aligned hull of [first entry, first entry + entries) (16-byte
// alignment of the column stream: 4 x 4 B; of the value stream: 4 x 8 B x bs^2), so tiles need
// no alignment themselves.  Returns -1 when a single row exceeds max_entries (the caller then
// keeps the generic kernel).
#define NSB_TILE_WORDS 64
#define NSB_TILE_MAX_ROWS 119

extern "C" int nsb_tiles_build(int32_t nrows, const int32_t* rowptr, int32_t max_entries, int32_t* ntiles_out,
                               uint32_t* records) {
    if (!rowptr || nrows <= 0 || max_entries < 4 || max_entries > 65535 || !ntiles_out)
        return nsb_fail(nullptr, "nsb_tiles_build: bad arguments%s%s");
    int32_t nt = 0;
    int32_t r0 = 0;
    while (r0 < nrows) {
        int32_t best = -1;
        for (int32_t r1 = r0 + 1; r1 <= nrows && r1 - r0 <= NSB_TILE_MAX_ROWS; ++r1) {
            if (rowptr[r1] - rowptr[r0] > max_entries) break;
            best = r1;
        }
        if (best < 0) return nsb_fail(nullptr, "nsb_tiles_build: a row is longer than the tile%s%s");
        if (records) {
            uint32_t* rec = records + (size_t)nt * NSB_TILE_WORDS;
            memset(rec, 0, sizeof(uint32_t) * NSB_TILE_WORDS);
            rec[0] = (uint32_t)r0;
            rec[1] = (uint32_t)(best - r0);
            rec[2] = (uint32_t)rowptr[r0];
            rec[3] = (uint32_t)(rowptr[best] - rowptr[r0]);
            uint16_t* off = reinterpret_cast<uint16_t*>(rec + 4);
            for (int32_t r = r0; r <= best; ++r) off[r - r0] = (uint16_t)(rowptr[r] - rowptr[r0]);
        }
        ++nt;
        r0 = best;
    }
    *ntiles_out = nt;
    return 0;
}

// ---- row groups of the warp-cooperative assembly kernels ------------------------------------------
// Consecutive rows with at most max_pairs (row, cell) incidence pairs per group (a single row
// with more pairs forms its own group).  group_row[ngroups + 1]; group_row == NULL only counts.
extern "C" int nsb_row_groups_build(int32_t nrows, const int32_t* pair_ptr, int32_t max_pairs, int32_t* ngroups_out,
                                    int32_t* group_row) {
    if (!pair_ptr || nrows <= 0 || max_pairs < 1 || !ngroups_out)
        return nsb_fail(nullptr, "nsb_row_groups_build: bad arguments%s%s");
    int32_t ng = 0, r0 = 0;
    while (r0 < nrows) {
        int32_t r1 = r0 + 1;
        while (r1 < nrows && pair_ptr[r1 + 1] - pair_ptr[r0] <= max_pairs) ++r1;
        if (group_row) group_row[ng] = r0;
        ++ng;
        r0 = r1;
    }
    if (group_row) group_row[ng] = nrows;
    *ngroups_out = ng;
    return 0;
}
