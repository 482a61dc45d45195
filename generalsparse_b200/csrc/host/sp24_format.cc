// Host-side builder of the B200 2:4 image (sp24_format.hpp).
#include "sp24_format.hpp"

#include <algorithm>
#include <array>
#include <cstring>
#include <stdexcept>
#include <utility>

namespace gs {

void build_sp24_image(const uint32_t *rowptr, const uint32_t *col, const uint16_t *val, uint32_t n_rows, uint32_t n_cols, uint32_t group, sp24_image &im)
{
    if (group < 1 || group > 2) throw std::invalid_argument("2:4 image: group must be 1 or 2");
    im = sp24_image();
    im.n_rows = n_rows, im.n_cols = n_cols, im.group = group;
    im.n_panels = ((n_rows + SP24_ROWS - 1) / SP24_ROWS + group - 1) / group * group;
    im.k_tiles = (n_cols + SP24_COLS - 1) / SP24_COLS;
    const uint64_t T = (uint64_t)im.n_panels * im.k_tiles;
    if (T + 1 > 0xffffffffull) throw std::invalid_argument("2:4 image: too many tiles");
    im.bytes.assign(T * SP24_TILE_BYTES, 0);
    // every group starts as "keeps elements 0 and 1, both zero": a valid pattern whatever follows
    for (uint64_t t = 0; t < T; t++) std::fill_n(im.bytes.begin() + t * SP24_TILE_BYTES + SP24_A_BYTES, SP24_E_BYTES, (uint8_t)0x44);
    im.rem_rowptr.assign((size_t)n_rows + 1, 0);

    std::vector<std::pair<uint32_t, uint16_t>> row;
    for (uint32_t r = 0; r < n_rows; r++) {
        row.clear();
        for (uint32_t i = rowptr[r]; i < rowptr[r + 1]; i++) {
            if ((val[i] & 0x7fffu) == 0) continue; // +-0: the chain's padding
            if (col[i] >= n_cols) throw std::invalid_argument("2:4 image: column index out of range");
            row.emplace_back(col[i], val[i]);
        }
        std::sort(row.begin(), row.end(), [](const std::pair<uint32_t, uint16_t> &x, const std::pair<uint32_t, uint16_t> &y) { return x.first < y.first; });
        const uint64_t p = r / SP24_ROWS;
        const uint32_t rr = r % SP24_ROWS;
        size_t i = 0;
        while (i < row.size()) {
            const uint32_t g = row[i].first >> 2; // column group
            size_t j = i;
            while (j < row.size() && (row[j].first >> 2) == g) {
                if (j > i && row[j].first == row[j - 1].first) throw std::invalid_argument("2:4 image: duplicate column inside a row");
                j++;
            }
            uint8_t *tile = im.bytes.data() + sp24_tile_index(p, (uint64_t)g * 4u / SP24_COLS, im.k_tiles, group) * SP24_TILE_BYTES;
            const uint32_t k = (g * 4u) % SP24_COLS, slot = k / 2u; // first slot of the group in the compressed row
            uint32_t i0 = row[i].first & 3u, i1;
            uint16_t v0 = row[i].second, v1 = 0;
            if (j - i >= 2) i1 = row[i + 1].first & 3u, v1 = row[i + 1].second;
            else if (i0 == 3u) i1 = 3u, v1 = v0, i0 = 0u, v0 = 0; // a lone element in the last column: (0 holding zero, 3)
            else i1 = 3u;                                        // a lone element elsewhere: (itself, 3 holding zero)
            uint8_t *a0 = tile + sp24_a_offset(rr, slot), *a1 = tile + sp24_a_offset(rr, slot + 1);
            a0[0] = (uint8_t)(v0 & 0xff), a0[1] = (uint8_t)(v0 >> 8), a1[0] = (uint8_t)(v1 & 0xff), a1[1] = (uint8_t)(v1 >> 8);
            uint8_t &e = tile[sp24_e_byte(rr, k)];
            const uint32_t sh = sp24_e_shift(k);
            e = (uint8_t)((e & ~(0xfu << sh)) | ((i0 | (i1 << 2)) << sh));
            im.nnz += std::min<size_t>(j - i, 2);
            for (size_t q = i + 2; q < j; q++) im.rem_col.push_back(row[q].first), im.rem_val.push_back(row[q].second);
            i = j;
        }
        if (im.rem_col.size() > 0xfffffff0ull) throw std::invalid_argument("2:4 image: remainder too large");
        im.rem_rowptr[r + 1] = (uint32_t)im.rem_col.size();
    }
}

void expand_sp24_image(const sp24_image &im, std::vector<uint16_t> &dense)
{
    dense.assign((size_t)im.n_rows * im.n_cols, 0);
    for (uint32_t r = 0; r < im.n_rows; r++) {
        const uint64_t p = r / SP24_ROWS;
        const uint32_t rr = r % SP24_ROWS;
        for (uint32_t g = 0; g * 4u < im.n_cols; g++) {
            const uint8_t *tile = im.bytes.data() + sp24_tile_index(p, (uint64_t)g * 4u / SP24_COLS, im.k_tiles, im.group) * SP24_TILE_BYTES;
            const uint32_t k = (g * 4u) % SP24_COLS, slot = k / 2u;
            const uint32_t e = (tile[sp24_e_byte(rr, k)] >> sp24_e_shift(k)) & 0xfu;
            for (uint32_t s = 0; s < 2; s++) {
                const uint8_t *a = tile + sp24_a_offset(rr, slot + s);
                const uint16_t v = (uint16_t)(a[0] | (a[1] << 8));
                const uint64_t c = (uint64_t)g * 4u + ((e >> (2u * s)) & 3u);
                if ((v & 0x7fffu) && c < im.n_cols) dense[(size_t)r * im.n_cols + c] = v;
            }
        }
        for (uint32_t i = im.rem_rowptr[r]; i < im.rem_rowptr[r + 1]; i++) dense[(size_t)r * im.n_cols + im.rem_col[i]] = im.rem_val[i];
    }
}

void localize_sp24_remainder(const sp24_image &im, uint32_t row_bytes, sp24_tiled &out)
{
    if (row_bytes != 32 && row_bytes != 64 && row_bytes != 128) throw std::invalid_argument("2:4 image: B's rows are 32, 64 or 128 bytes");
    if (im.group != 1) throw std::invalid_argument("2:4 image: the tile-local remainder needs one panel per unit");
    const uint64_t T = (uint64_t)im.n_panels * im.k_tiles;
    std::vector<uint32_t> count(T, 0);
    for (uint32_t r = 0; r < im.n_rows; r++)
        for (uint32_t i = im.rem_rowptr[r]; i < im.rem_rowptr[r + 1]; i++) count[(uint64_t)(r / SP24_ROWS) * im.k_tiles + im.rem_col[i] / SP24_COLS]++;
    uint32_t mx = 0;
    for (uint64_t t = 0; t < T; t++) mx = std::max(mx, count[t]);
    if (mx > 0xffffu) throw std::invalid_argument("2:4 image: too many remainder entries in one tile");
    out = sp24_tiled();
    out.max_entries = mx, out.row_bytes = row_bytes;
    out.unit_bytes = SP24_TILE_BYTES + SP24_REM_HEADER + ((mx + 3u) & ~3u) * 4u;
    out.bytes.assign(T * out.unit_bytes, 0);
    for (uint64_t t = 0; t < T; t++) std::copy_n(im.bytes.begin() + t * SP24_TILE_BYTES, SP24_TILE_BYTES, out.bytes.begin() + t * out.unit_bytes);
    // Rows ascending; INSIDE a row the order of the entries is free (a sum), and it decides how the remainder warps' shared-memory
    // gathers collide: a warp holds 16 consecutive tile rows (a pair of threads per row), in round j every row reads the row of B its
    // j-th entry names, one 16-byte chunk per thread, i.e. 32 requests into the 8 sixteen-byte bank groups of a 128-byte line.  Entries
    // in column order land at random: ~7 of the 4 wavefronts a request needs at best (58 % of the kernel's shared-load wavefronts were
    // bank conflicts at N = 32, profiles/traffic.json).  So every row picks, position by position, the entry whose bank groups are
    // least used at that position by the rows of its warp so far (greedy; the groups come from the very offsets the threads will use).
    const uint32_t cpt = row_bytes >= 64 ? row_bytes / 32u : 1u; // 16-byte chunks a thread reads per entry (b200_sp24.cuh: CPT)
    auto groups_of = [&](uint32_t off, uint32_t g[2]) {
        if (row_bytes == 16) g[0] = g[1] = (off >> 4) & 7u; // NL = 8: the pair's threads take alternate entries, one chunk each
        else g[0] = (off >> 4) & 7u, g[1] = ((off ^ (cpt << 4)) >> 4) & 7u;
    };
    std::vector<std::vector<uint32_t>> rows_of_tile((size_t)im.k_tiles * SP24_ROWS);
    for (uint32_t p = 0; p < im.n_panels; p++) {
        for (auto &v : rows_of_tile) v.clear();
        for (uint32_t rr = 0; rr < SP24_ROWS; rr++) {
            const uint32_t r = p * SP24_ROWS + rr;
            if (r >= im.n_rows) break;
            for (uint32_t i = im.rem_rowptr[r]; i < im.rem_rowptr[r + 1]; i++) {
                const uint32_t kt = im.rem_col[i] / SP24_COLS;
                rows_of_tile[(size_t)kt * SP24_ROWS + rr].push_back((sp24_b_row_offset(im.rem_col[i] % SP24_COLS, row_bytes) << 16) | im.rem_val[i]);
            }
        }
        for (uint32_t kt = 0; kt < im.k_tiles; kt++) {
            uint8_t *unit = out.bytes.data() + ((uint64_t)p * im.k_tiles + kt) * out.unit_bytes;
            uint32_t n = 0;
            if (row_bytes == 64) {
                // N = 32: four threads read one 64-byte row of B (b200_sp24.cuh, NL = 32), a wavefront's eight lanes cover tile rows 2 j and 2 j + 1, and the
                // two reads collide exactly when both rows of B lie in the same 64-byte half of the banks, i.e. when the two column indices have the
                // same parity.  So an even tile row lists its even-column entries first, and its odd neighbour picks, position by position, an entry
                // of the other parity while it has one.
                auto half_of = [](uint32_t w) { return ((w >> 16) >> 6) & 1u; };
                for (uint32_t rr = 0; rr < SP24_ROWS; rr += 2) {
                    std::vector<uint32_t> &e0 = rows_of_tile[(size_t)kt * SP24_ROWS + rr], &e1 = rows_of_tile[(size_t)kt * SP24_ROWS + rr + 1];
                    std::stable_partition(e0.begin(), e0.end(), [&](uint32_t w) { return half_of(w) == 0; });
                    for (size_t pos = 0; pos < e1.size(); pos++) {
                        const uint32_t want = pos < e0.size() ? 1u - half_of(e0[pos]) : 1u;
                        for (size_t q = pos; q < e1.size(); q++)
                            if (half_of(e1[q]) == want) {
                                std::swap(e1[pos], e1[q]);
                                break;
                            }
                    }
                    for (uint32_t r2 = rr; r2 < rr + 2; r2++) {
                        unit[SP24_TILE_BYTES + 2u * r2] = (uint8_t)(n & 0xff), unit[SP24_TILE_BYTES + 2u * r2 + 1] = (uint8_t)(n >> 8);
                        for (uint32_t w : rows_of_tile[(size_t)kt * SP24_ROWS + r2]) {
                            uint8_t *dst = unit + SP24_TILE_BYTES + SP24_REM_HEADER + 4u * n++;
                            dst[0] = (uint8_t)w, dst[1] = (uint8_t)(w >> 8), dst[2] = (uint8_t)(w >> 16), dst[3] = (uint8_t)(w >> 24);
                        }
                    }
                }
            } else
            for (uint32_t w0 = 0; w0 < SP24_ROWS; w0 += 16) { // the 16 rows of one remainder warp
                std::vector<std::array<uint16_t, 8>> used; // used[position][bank group]
                for (uint32_t rr = w0; rr < w0 + 16; rr++) {
                    unit[SP24_TILE_BYTES + 2u * rr] = (uint8_t)(n & 0xff), unit[SP24_TILE_BYTES + 2u * rr + 1] = (uint8_t)(n >> 8);
                    std::vector<uint32_t> &e = rows_of_tile[(size_t)kt * SP24_ROWS + rr];
                    if (used.size() < e.size()) used.resize(e.size(), std::array<uint16_t, 8>{});
                    for (size_t pos = 0; pos < e.size(); pos++) {
                        size_t best = pos;
                        uint32_t best_cost = ~0u;
                        for (size_t q = pos; q < e.size(); q++) {
                            uint32_t g[2];
                            groups_of(e[q] >> 16, g);
                            const uint32_t cost = std::max<uint32_t>(used[pos][g[0]], used[pos][g[1]]) * 16u + used[pos][g[0]] + used[pos][g[1]];
                            if (cost < best_cost) best_cost = cost, best = q;
                        }
                        std::swap(e[pos], e[best]);
                        uint32_t g[2];
                        groups_of(e[pos] >> 16, g);
                        used[pos][g[0]]++;
                        if (g[1] != g[0]) used[pos][g[1]]++;
                        uint8_t *dst = unit + SP24_TILE_BYTES + SP24_REM_HEADER + 4u * n++;
                        const uint32_t w = e[pos];
                        dst[0] = (uint8_t)w, dst[1] = (uint8_t)(w >> 8), dst[2] = (uint8_t)(w >> 16), dst[3] = (uint8_t)(w >> 24);
                    }
                }
            }
            unit[SP24_TILE_BYTES + 2u * SP24_ROWS] = (uint8_t)(n & 0xff), unit[SP24_TILE_BYTES + 2u * SP24_ROWS + 1] = (uint8_t)(n >> 8);
        }
    }
}

void expand_sp24_tiled(const sp24_image &geo, const sp24_tiled &t, std::vector<uint16_t> &dense)
{
    sp24_image im; // the images back to back, no remainder: expand, then add the tile-local entries
    im.n_rows = geo.n_rows, im.n_cols = geo.n_cols, im.n_panels = geo.n_panels, im.k_tiles = geo.k_tiles, im.group = 1;
    const uint64_t T = (uint64_t)im.n_panels * im.k_tiles;
    im.bytes.resize(T * SP24_TILE_BYTES);
    for (uint64_t i = 0; i < T; i++) std::copy_n(t.bytes.begin() + i * t.unit_bytes, SP24_TILE_BYTES, im.bytes.begin() + i * SP24_TILE_BYTES);
    im.rem_rowptr.assign((size_t)im.n_rows + 1, 0);
    expand_sp24_image(im, dense);
    for (uint64_t i = 0; i < T; i++) {
        const uint8_t *blk = t.bytes.data() + i * t.unit_bytes + SP24_TILE_BYTES, *ent = blk + SP24_REM_HEADER;
        const uint64_t p = i / im.k_tiles, kt = i % im.k_tiles;
        auto start = [&](uint32_t rr) { return (uint32_t)(blk[2 * rr] | (blk[2 * rr + 1] << 8)); };
        for (uint32_t rr = 0; rr < SP24_ROWS; rr++)
            for (uint32_t e = start(rr); e < start(rr + 1); e++) {
                const uint32_t w = (uint32_t)ent[4 * e] | ((uint32_t)ent[4 * e + 1] << 8) | ((uint32_t)ent[4 * e + 2] << 16) | ((uint32_t)ent[4 * e + 3] << 24);
                const uint64_t r = p * SP24_ROWS + rr, c = kt * SP24_COLS + (w >> 16) / t.row_bytes;
                if (r >= im.n_rows || c >= im.n_cols || e >= t.max_entries || sp24_b_row_offset((w >> 16) / t.row_bytes, t.row_bytes) != (w >> 16)) throw std::invalid_argument("2:4 image: tile-local entry out of range");
                dense[r * im.n_cols + c] = (uint16_t)(w & 0xffffu);
            }
    }
}

bool compact_sp24_units(const uint8_t *units, size_t U, bool tiled, uint32_t n_panels, uint32_t k_tiles, double keep_frac, sp24_compact &out)
{
    out = sp24_compact();
    out.panel_ptr.assign(1, 0u);
    for (uint32_t p = 0; p < n_panels; p++) {
        const size_t before = out.unit_kt.size();
        for (uint32_t kt = 0; kt < k_tiles; kt++) {
            const uint8_t *t = units + ((size_t)p * k_tiles + kt) * U;
            bool any = false;
            for (size_t i = 0; i < SP24_A_BYTES && !any; i += 8) {
                uint64_t w;
                std::memcpy(&w, t + i, 8);
                any = w != 0;
            }
            if (!any && tiled) any = (t[SP24_TILE_BYTES + 2 * SP24_ROWS] | t[SP24_TILE_BYTES + 2 * SP24_ROWS + 1]) != 0; // start[128]: entries in the block
            if (any) out.unit_kt.push_back(kt);
        }
        if (out.unit_kt.size() == before) out.unit_kt.push_back(0u);
        out.panel_ptr.push_back((uint32_t)out.unit_kt.size());
    }
    if ((double)out.unit_kt.size() > keep_frac * (double)n_panels * (double)k_tiles) return false;
    out.bytes.resize(out.unit_kt.size() * U);
    for (uint32_t p = 0; p < n_panels; p++)
        for (uint32_t u = out.panel_ptr[p]; u < out.panel_ptr[p + 1]; u++) std::memcpy(out.bytes.data() + (size_t)u * U, units + ((size_t)p * k_tiles + out.unit_kt[u]) * U, U);
    return true;
}

} // namespace gs
