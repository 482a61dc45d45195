// B200 2:4 image: the device layout of the sparse tensor-core path (GS_FAMILY_B200_SP24, tcgen05.mma.sp).  Built once
// per plan from what a row-direction chain of the reference produces (nz_col_indices, nz_vals and the row level's
// first_nz_indices -- total_BMT_result_reduce_to_one_register_token.cc:828-892 reads the same three); the
// reference-order arrays stay untouched.
//
// The matrix is cut into tiles of 128 rows x 128 columns.  `group` (1 or 2) consecutive row panels form a unit of work:
// tile (panel, k_tile) lies at index ((panel / group) * k_tiles + k_tile) * group + panel % group, so the `group` tiles a CTA
// multiplies with one piece of B are one contiguous run (group = 1: panel-major).  The panel count is padded to a multiple of
// `group` with empty tiles.  Inside a
// row every group of four consecutive columns keeps at most two elements (the 2:4 pattern of the sparse tensor core).
// A tile is SP24_TILE_BYTES = 18 432 bytes, copied to shared memory by ONE bulk copy and consumed as it lies:
//   [0, 16 384)      the compressed operand: 128 rows x 64 fp16 (slot 2g / 2g+1 = the two kept elements of column
//                    group g, in column order), in the K-major 128-byte-swizzled layout tcgen05.mma reads
//                    (row r, slot c -> r*128 + (((c>>3) ^ (r&7)) << 4) + (c&7)*2);
//   [16 384, 18 432) the metadata, 4 bits per group (index of the first kept element | index of the second << 2,
//                    first < second), arranged as 128 rows x 16 bytes so that `tcgen05.cp.128x128b` drops it into
//                    4 columns of tensor memory in the order the instruction expects (the layout CUTLASS names
//                    TensorEAtom_MMA_F16, cutlass/gemm/collective/builders/sm1xx_sparse_config.inl): the 32-bit word of
//                    16-byte row L = (r & 7) + 8*((k>>4)&1) + 16*(r>>4) and column j = k>>5 holds, in its low half,
//                    the four groups k&15 of tile row r with ((r>>3)&1) == 0, and in its high half those of row r + 8.
// Elements a group cannot hold (a third and fourth non-zero) go to the remainder, a CSR matrix over the same rows that
// a gather kernel multiplies: an exactly 2:4-pruned matrix has none, a 70 %-sparse unstructured one leaves 7.7 % of
// its non-zeros there.  Padding zeros the chain added (modify_*_by_col_pad_in_sub_matrix.cc) are dropped.
#pragma once
#include <cstdint>
#include <vector>

namespace gs {

constexpr uint32_t SP24_ROWS = 128, SP24_COLS = 128, SP24_A_BYTES = SP24_ROWS * (SP24_COLS / 2) * 2, SP24_E_BYTES = SP24_ROWS * 16,
                   SP24_TILE_BYTES = SP24_A_BYTES + SP24_E_BYTES;

struct sp24_image {
    uint32_t n_rows = 0, n_cols = 0;
    uint32_t n_panels = 0, k_tiles = 0, group = 1; // n_panels: padded to a multiple of group
    uint64_t nnz = 0;           // non-zeros held by the image
    std::vector<uint8_t> bytes; // n_panels * k_tiles * SP24_TILE_BYTES
    // what the 2:4 pattern could not hold, CSR over the same n_rows rows, columns ascending inside a row
    std::vector<uint32_t> rem_rowptr, rem_col;
    std::vector<uint16_t> rem_val;
};

inline uint32_t sp24_a_offset(uint32_t r, uint32_t slot) { return r * 128u + ((((slot >> 3) ^ (r & 7u)) & 7u) << 4) + (slot & 7u) * 2u; }
// byte and shift of the 4-bit metadata of column group (k >> 2) of tile row r
inline uint32_t sp24_e_byte(uint32_t r, uint32_t k)
{
    const uint32_t L = (r & 7u) + 8u * ((k >> 4) & 1u) + 16u * (r >> 4);
    return SP24_A_BYTES + 16u * L + 4u * (k >> 5) + 2u * ((r >> 3) & 1u) + ((k >> 3) & 1u);
}
inline uint32_t sp24_e_shift(uint32_t k) { return ((k >> 2) & 1u) * 4u; }
inline uint64_t sp24_tile_index(uint64_t panel, uint64_t k_tile, uint64_t k_tiles, uint64_t group) { return ((panel / group) * k_tiles + k_tile) * group + panel % group; }

// rowptr: n_rows + 1 offsets into col / val_f16 (row i of the row level = row i of the image).  Duplicate columns
// inside a row with a non-zero value are refused (std::invalid_argument).
void build_sp24_image(const uint32_t *rowptr, const uint32_t *col, const uint16_t *val_f16, uint32_t n_rows, uint32_t n_cols, uint32_t group, sp24_image &out);

// inverse, for tests: dense row-major n_rows x n_cols fp16 bits of image + remainder
void expand_sp24_image(const sp24_image &im, std::vector<uint16_t> &dense);

// The remainder cut along the tiles (group = 1 only): every tile image is followed by the remainder entries that fall into its
// 128 x 128 window, so the CTA that multiplies the tile finds them -- and the 128 rows of B they need, which the tile's own
// multiplication has brought to shared memory -- in the same stage of its copy ring.  A unit is `unit_bytes` =
// SP24_TILE_BYTES + SP24_REM_HEADER + 4 * (the largest entry count of any tile, rounded up to 4) bytes:
//   [0, 18 432)            the tile image as above;
//   [18 432, + 272)        u16 start[129] (+ padding): entries of tile row r are start[r] .. start[r+1];
//   [18 704, unit_bytes)   u32 entries, rows ascending, columns ascending inside a row: (where B's row lies << 16) | fp16 bits.
// "Where B's row lies": the byte offset of the row's first 16-byte chunk inside the 128 x N piece of B as the tensor map
// swizzles it into shared memory, k * row_bytes + (((k * row_bytes) >> 7) & (row_bytes / 16 - 1)) * 16 for column k of the
// tile and row_bytes = 2 N (32, 64 or 128): the thread that wants chunk c of the row reads at (offset ^ (c << 4)).
struct sp24_tiled {
    uint32_t unit_bytes = 0, max_entries = 0, row_bytes = 0;
    std::vector<uint8_t> bytes; // n_panels * k_tiles * unit_bytes
};
constexpr uint32_t SP24_REM_HEADER = 272;
inline uint32_t sp24_b_row_offset(uint32_t k, uint32_t row_bytes) { return k * row_bytes + ((((k * row_bytes) >> 7) & (row_bytes / 16u - 1u)) << 4); }
void localize_sp24_remainder(const sp24_image &im, uint32_t row_bytes, sp24_tiled &out);
void expand_sp24_tiled(const sp24_image &im, const sp24_tiled &t, std::vector<uint16_t> &dense); // im: geometry only, its remainder arrays are not read

// The stream without the tiles that hold nothing (group = 1): units of `unit_bytes` (SP24_TILE_BYTES, or a sp24_tiled unit when
// `tiled`), panel-major; the units of panel p are panel_ptr[p] .. panel_ptr[p+1] and unit u multiplies columns
// 128 * unit_kt[u] ...  A tile is kept when its compressed operand holds a non-zero bit or its tile-local block an entry;
// a panel with none keeps its first tile, so that its rows of the product are written.
struct sp24_compact {
    std::vector<uint32_t> panel_ptr, unit_kt;
    std::vector<uint8_t> bytes; // unit_kt.size() * unit_bytes
};
// fills panel_ptr / unit_kt always; copies the kept units (and returns true) only if they are at most keep_frac of all units
bool compact_sp24_units(const uint8_t *units, size_t unit_bytes, bool tiled, uint32_t n_panels, uint32_t k_tiles, double keep_frac, sp24_compact &out);

} // namespace gs
