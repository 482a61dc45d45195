// Host-side builder of the B200 2:4 image (sp24_format.hpp).
#include "sp24_format.hpp"

#include <algorithm>
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

} // namespace gs
