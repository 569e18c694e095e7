// setup_host.cu -- CPU-side setup helpers of libnsb200 (no kernels): sparsity pattern and the
// row -> (cell, local row, column slots) incidence that drives the row-gather assembly.
//
// Replaces dolfin's SparsityPatternBuilder / PETSc AIJ preallocation that the reference
// triggers implicitly through its function spaces (source/ns_solver_base.py:501-524) and the
// first assemble() of each form.  Pattern semantics: union over cells of
// row dofs x col dofs of the whole element, ascending int32 columns, no extra diagonal.
#include <algorithm>
#include <cstdlib>
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

// ---- conflict-free scatter schedule of the warp-cooperative matrix assembly (assemble_jac.cu) -----
// Rows are grouped (consecutive rows, <= 32 incidence pairs and <= max_blocks slots per group; a
// bigger single row forms its own group and is processed in chunks of 32 pairs).  In a chunk, lane
// q owns pair q and must add its nloc local columns into the shared copy of the group's rows.  The
// schedule orders these additions in ROUNDS such that within a round all lanes hit distinct slots
// (a proper edge colouring of the bipartite pair/slot graph): no atomics, fixed summation order.
//   * an entry that receives more than DEG_SPLIT contributions (the diagonal of a vertex row: one per
//     cell around the vertex) would force that many rounds; it gets SHADOW slots behind the group's
//     real entries, its contributions are dealt round-robin over the copies, and the copies are
//     folded into the real entry in a fixed order before the rows are written (fold list);
//   * among the free rounds the colouring prefers one in which no lane of the same half warp hits
//     the same shared-memory bank (slot mod 16 for 8-byte words at an odd stride).
//   group_row[ng+1]    first row of every group
//   chunk_first[ng+1]  first chunk of every group (chunk c covers pairs pair_ptr[row_lo] + 32 (c - first) ...)
//   round_ptr[nc+1]    first round of every chunk
//   sched[32 * rounds] per (round, lane): slot index inside the group << 4 | local column, 0xFFFF = idle
//   fold_ptr[ng+1], fold[nfold]   per group: (real entry << 16 | shadow slot), shadows of one entry consecutive
// Call with group_row == NULL to count (ngroups, nchunks, nrounds, nfold), then again to fill.
static int sched_env(const char* name, int dflt) {
    const char* v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

extern "C" int nsb_sched_build(int32_t nrows, const int32_t* rowptr, const int32_t* pair_ptr,
                               const uint8_t* pair_slots, int32_t nloc, int32_t max_blocks, int64_t* counts4,
                               int32_t* group_row, int32_t* chunk_first, int32_t* round_ptr, uint16_t* sched,
                               int32_t* fold_ptr, uint32_t* fold) {
    if (!rowptr || !pair_ptr || !pair_slots || !counts4 || nrows <= 0 || nloc <= 0 || nloc > 15 || max_blocks < 1 ||
        max_blocks > 4095)
        return nsb_fail(nullptr, "nsb_sched_build: bad arguments%s%s");
    const int DEG_SPLIT = sched_env("NSB_SCHED_SPLIT", 8);       // contributions per copy of a hot entry
    const bool bank_aware = sched_env("NSB_SCHED_BANKS", 1) != 0;
    const int max_pairs = 32;
    const int slack = sched_env("NSB_SCHED_SLACK", 0);          // rounds beyond nloc a bank-clean placement may open
    // static degree of every entry = number of (pair, local column) contributions; shadows per row
    const int64_t nnz = rowptr[nrows];
    std::vector<uint8_t> deg8((size_t)nnz, 0);
    std::vector<int32_t> row_shadows((size_t)nrows, 0);
#pragma omp parallel for schedule(static, 4096)
    for (int32_t r = 0; r < nrows; ++r) {
        for (int32_t p = pair_ptr[r]; p < pair_ptr[r + 1]; ++p)
            for (int32_t j = 0; j < nloc; ++j) {
                uint8_t& d = deg8[(size_t)rowptr[r] + pair_slots[(size_t)p * nloc + j]];
                if (d < 255) ++d;
            }
        int32_t sh = 0;
        for (int32_t e = rowptr[r]; e < rowptr[r + 1]; ++e) sh += (deg8[e] + DEG_SPLIT - 1) / DEG_SPLIT - (deg8[e] ? 1 : 0);
        row_shadows[r] = sh;
    }
    // pass 1: groups
    std::vector<int32_t> grow;
    for (int32_t r0 = 0; r0 < nrows;) {
        int32_t r1 = r0 + 1;
        int32_t slots = rowptr[r1] - rowptr[r0] + row_shadows[r0];
        while (r1 < nrows && pair_ptr[r1 + 1] - pair_ptr[r0] <= max_pairs &&
               slots + rowptr[r1 + 1] - rowptr[r1] + row_shadows[r1] <= max_blocks) {
            slots += rowptr[r1 + 1] - rowptr[r1] + row_shadows[r1];
            ++r1;
        }
        if (slots > 4095) return nsb_fail(nullptr, "nsb_sched_build: row too long%s%s");
        grow.push_back(r0);
        r0 = r1;
    }
    grow.push_back(nrows);
    const int32_t ng = (int32_t)grow.size() - 1;
    std::vector<int32_t> cfirst((size_t)ng + 1, 0), ffirst((size_t)ng + 1, 0);
    for (int32_t g = 0; g < ng; ++g) {
        const int32_t np = pair_ptr[grow[g + 1]] - pair_ptr[grow[g]];
        cfirst[g + 1] = cfirst[g] + (np + 31) / 32;
        int32_t sh = 0;
        for (int32_t r = grow[g]; r < grow[g + 1]; ++r) sh += row_shadows[r];
        ffirst[g + 1] = ffirst[g] + sh;
    }
    const int32_t nc = cfirst[ng];
    // pass 2: colour every chunk (greedy: slots by descending degree, smallest admissible round)
    std::vector<int32_t> nround((size_t)nc, 0);
    std::vector<std::vector<uint16_t>> tmp;
    const bool fill = group_row != nullptr;
    if (fill) {
        if (!chunk_first || !round_ptr || !sched || !fold_ptr || (ffirst[ng] > 0 && !fold))
            return nsb_fail(nullptr, "nsb_sched_build: null output%s%s");
        tmp.resize((size_t)nc);
    }
#pragma omp parallel
    {
        std::vector<int32_t> slot, lane_of, col_of, order, deg, shadow_base, seen;
        std::vector<uint64_t> lane_busy, slot_busy, bank_busy;   // bitsets over rounds (W x 64)
        std::vector<uint16_t> buf;
#pragma omp for schedule(dynamic, 256)
        for (int32_t g = 0; g < ng; ++g) {
            const int32_t row_lo = grow[g], row_hi = grow[g + 1];
            const int32_t p_lo = pair_ptr[row_lo], p_hi = pair_ptr[row_hi], e_lo = rowptr[row_lo];
            const int32_t nent = rowptr[row_hi] - e_lo;
            const int32_t nslot = nent + (ffirst[g + 1] - ffirst[g]);
            // shadow slots of the hot entries, in entry order
            shadow_base.assign((size_t)nent, -1);
            seen.assign((size_t)nent, 0);
            int32_t next_shadow = nent, nf = ffirst[g];
            for (int32_t e = 0; e < nent; ++e) {
                const int32_t copies = deg8[(size_t)e_lo + e] ? (deg8[(size_t)e_lo + e] + DEG_SPLIT - 1) / DEG_SPLIT : 1;
                if (copies > 1) {
                    shadow_base[e] = next_shadow;
                    if (fill)
                        for (int32_t k = 0; k + 1 < copies; ++k) fold[nf++] = ((uint32_t)e << 16) | (uint32_t)(next_shadow + k);
                    next_shadow += copies - 1;
                }
            }
            int32_t row = row_lo;
            for (int32_t c = cfirst[g]; c < cfirst[g + 1]; ++c) {
                const int32_t q0 = p_lo + 32 * (c - cfirst[g]), q1 = std::min(p_hi, q0 + 32);
                const int32_t nedge = (q1 - q0) * nloc;
                slot.resize(nedge); lane_of.resize(nedge); col_of.resize(nedge);
                deg.assign((size_t)nslot, 0);
                int32_t k = 0;
                for (int32_t p = q0; p < q1; ++p) {
                    while (pair_ptr[row + 1] <= p) ++row;
                    for (int32_t j = 0; j < nloc; ++j, ++k) {
                        const int32_t e = rowptr[row] - e_lo + pair_slots[(size_t)p * nloc + j];
                        int32_t s = e;
                        if (shadow_base[e] >= 0) {
                            const int32_t copies = (deg8[(size_t)e_lo + e] + DEG_SPLIT - 1) / DEG_SPLIT;
                            const int32_t cp = seen[e]++ % copies;
                            if (cp) s = shadow_base[e] + cp - 1;
                        }
                        slot[k] = s; lane_of[k] = p - q0; col_of[k] = j;
                        deg[s]++;
                    }
                }
                order.resize(nedge);
                for (int32_t i = 0; i < nedge; ++i) order[i] = i;
                std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return deg[slot[a]] > deg[slot[b]]; });
                const int W = 4;   // up to 256 rounds
                lane_busy.assign(32 * W, 0); slot_busy.assign((size_t)nslot * W, 0); bank_busy.assign(2 * 16 * W, 0);
                int32_t rounds = 0;
                std::vector<uint16_t>* out = fill ? &buf : nullptr;
                if (out) out->assign((size_t)32 * 256, 0xFFFFu);
                for (int32_t i = 0; i < nedge; ++i) {
                    const int32_t e = order[i];
                    const int32_t bk = ((lane_of[e] >> 4) * 16 + (slot[e] & 15)) * W;
                    int32_t t = -1, t_any = -1;
                    for (int wv = 0; wv < W && t < 0; ++wv) {
                        const uint64_t freebits = ~(lane_busy[lane_of[e] * W + wv] | slot_busy[(size_t)slot[e] * W + wv]);
                        if (freebits && t_any < 0) t_any = wv * 64 + __builtin_ctzll(freebits);
                        const uint64_t good = bank_aware ? freebits & ~bank_busy[bk + wv] : freebits;
                        if (good) t = wv * 64 + __builtin_ctzll(good);
                    }
                    // a bank-clean round beyond the rounds already open is not worth a new round
                    if (t < 0 || (t >= std::max(rounds, nloc + slack) && t_any < t)) t = t_any;
                    if (t < 0) t = 255;   // cannot happen: degree <= 255
                    lane_busy[lane_of[e] * W + t / 64] |= 1ull << (t % 64);
                    slot_busy[(size_t)slot[e] * W + t / 64] |= 1ull << (t % 64);
                    bank_busy[bk + t / 64] |= 1ull << (t % 64);
                    if (t + 1 > rounds) rounds = t + 1;
                    if (out) (*out)[(size_t)t * 32 + lane_of[e]] = (uint16_t)((slot[e] << 4) | col_of[e]);
                }
                nround[c] = rounds;
                if (out) tmp[c].assign(buf.begin(), buf.begin() + (size_t)rounds * 32);
            }
        }
    }
    int64_t total = 0;
    for (int32_t c = 0; c < nc; ++c) total += nround[c];
    counts4[0] = ng; counts4[1] = nc; counts4[2] = total; counts4[3] = ffirst[ng];
    if (!fill) return 0;
    memcpy(group_row, grow.data(), sizeof(int32_t) * ((size_t)ng + 1));
    memcpy(chunk_first, cfirst.data(), sizeof(int32_t) * ((size_t)ng + 1));
    memcpy(fold_ptr, ffirst.data(), sizeof(int32_t) * ((size_t)ng + 1));
    round_ptr[0] = 0;
    for (int32_t c = 0; c < nc; ++c) round_ptr[c + 1] = round_ptr[c] + nround[c];
#pragma omp parallel for schedule(static, 1024)
    for (int32_t c = 0; c < nc; ++c) memcpy(sched + (size_t)round_ptr[c] * 32, tmp[c].data(), sizeof(uint16_t) * tmp[c].size());
    return 0;
}
