// B200 tile stream: the device layout of a pruned-weight matrix for the decompress-to-shared-memory + tcgen05 path
// (GS_FAMILY_B200_TCGEN05).  Built once per plan from what a row-direction chain of the reference produces
// (nz_col_indices, nz_vals and the row level's first_nz_indices -- total_BMT_result_reduce_to_one_register_token.cc:828-892
// reads the same three); the reference-order arrays stay untouched.
//
// The matrix is cut into tiles of 128 rows x 64 columns, panel-major (tile t = panel * k_tiles + k_tile).  A tile is a
// run of (offset, value) pairs: `offset` is the BYTE offset of the element inside the 16 KB shared-memory image of the
// tile in the K-major 128-byte-swizzled operand layout tcgen05.mma reads (row r, column c ->
// r*128 + (((c>>3) ^ (r&7)) << 4) + (c&7)*2), so decompression is one 2-byte shared-memory store per non-zero.
// Runs start on multiples of 8 pairs (128-bit loads); a run is padded by repeating its last pair (an idempotent store).
// Padding zeros the chain added (modify_*_by_col_pad_in_sub_matrix.cc) carry no information here and are dropped.
#pragma once
#include <cstdint>
#include <vector>

namespace gs {

constexpr uint32_t TILE_ROWS = 128, TILE_COLS = 64, TILE_BYTES = TILE_ROWS * TILE_COLS * 2;

struct tile_stream {
    uint32_t n_rows = 0, n_cols = 0;
    uint32_t n_panels = 0, k_tiles = 0;
    uint64_t nnz = 0;               // pairs that carry a non-zero value (before run padding)
    std::vector<uint32_t> tile_ptr; // n_panels * k_tiles + 1 run starts, in pairs, multiples of 8
    std::vector<uint16_t> off;      // byte offset inside the tile image
    std::vector<uint16_t> val;      // fp16 bits
};

inline uint32_t tile_offset(uint32_t r, uint32_t c) { return r * 128u + ((((c >> 3) ^ (r & 7u)) & 7u) << 4) + (c & 7u) * 2u; }

// rowptr: n_rows + 1 offsets into col / val_f16 (row i of the row level = row i of the tiled matrix).
// pass_warps > 0: order every run so that the 32 pairs one warp stores together (pair j of lanes 0..31, 8 pairs per
// lane, pass_warps * 256 pairs per sweep of the producer warps) fall into distinct shared-memory banks where the
// tile's contents allow it; 0: row-major order inside a run.
void build_tile_stream(const uint32_t *rowptr, const uint32_t *col, const uint16_t *val_f16, uint32_t n_rows, uint32_t n_cols, uint32_t pass_warps, tile_stream &out);

// inverse, for tests: dense row-major n_rows x n_cols fp16 bits (zeros where nothing is stored)
void expand_tile_stream(const tile_stream &ts, std::vector<uint16_t> &dense);

} // namespace gs
