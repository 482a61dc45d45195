// The 2:4 image built on the device (sp24_build.cu): counts first (so that the host can choose the unit size and whether the empty tiles
// leave the stream, with the rules of kernels_sp24.cu / sp24_format.cc), then the fill.
#pragma once
#include "common.cuh"

namespace gsd {

struct sp24_dev_image {
    uint32_t n_panels = 0, k_tiles = 0, T = 0;
    uint64_t img_nnz = 0, rem_nnz = 0;
    uint32_t max_entries = 0; // the largest tile-local entry count
    bool bad = false;         // a row with unsorted or duplicate columns, or a column outside the matrix: the host builder refuses it
    // scratch of the build
    uint32_t *tile_img = nullptr, *rem_cnt = nullptr, *tile_rem = nullptr, *keep = nullptr, *unit_of_tile = nullptr;
    unsigned long long *scal = nullptr;
    void release();
    ~sp24_dev_image() { release(); }
};

cudaError_t sp24_dev_count(const launch_params &p, const uint32_t *rowptr, uint32_t n_rows, cudaStream_t s, sp24_dev_image &d);
// unit_bytes: SP24_TILE_BYTES (the plain image) or the tile-local unit size; compact: drop the tiles that hold nothing when at most keep_frac of them stay
cudaError_t sp24_dev_fill(const launch_params &p, const uint32_t *rowptr, uint32_t n_rows, uint32_t unit_bytes, uint32_t row_bytes, bool compact, double keep_frac, cudaStream_t s,
                          sp24_dev_image &d, unsigned char **img, uint32_t **panel_ptr, uint32_t **unit_kt, uint32_t *n_units);
cudaError_t sp24_dev_out_row(const launch_params &p, const uint32_t *rowid, uint32_t n_rows, uint32_t n_panels, cudaStream_t s, sp24_dev_image &d, uint32_t **out_row);

// 128 x 128 tiles that hold a stored entry (every row panel counts at least one): what the cost model prices the 2:4 image by
cudaError_t count_tiles128_dev(const void *col, bool col_is_u16, const uint32_t *rowptr, uint32_t n_rows, uint32_t n_cols, uint64_t nnz, cudaStream_t s, uint64_t *out);

} // namespace gsd
