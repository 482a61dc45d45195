// Transform steps of the SpMM path: flat-array re-implementations of the reference's
// data_transform_step classes (data_transform_step.hpp:249-4310, transform_step/*.cc).
// Every function is named after the reference class it replaces and produces bit-identical
// metadata arrays (tests/test_transform_parity.py compares against dumps of the reference
// pipeline).  Steps that differ only in the level they act on (BMTB/BMW/BMT) share one
// generic implementation taking POS_TYPE arguments; thin name-for-name wrappers sit in
// step_registry (steps.cc) so that records and tests can use the reference's names.
//
// Each step appends its reference-style record string ("<name>::{name:\"..\",target_matrix_id:..}")
// to `rec` when given, mirroring convert_to_string() of the reference classes.
#pragma once
#include "metadata.hpp"

namespace gs {
namespace steps {

using recs = std::vector<std::string>;

// ---- sort (operator/sort_operator.cc:82-105) -------------------------------------------------
void get_row_order_by_length(meta_data_set &ms, int id, recs *rec = nullptr);        // get_row_order_by_length.cc:14-126
void reorder_val_by_index(meta_data_set &ms, int id, recs *rec = nullptr);           // reorder_val_by_index.cc
void reorder_col_by_index(meta_data_set &ms, int id, recs *rec = nullptr);           // reorder_col_by_index.cc
void reorder_row_by_index(meta_data_set &ms, int id, recs *rec = nullptr);           // reorder_row_by_index.cc
void remove_empty_row_in_end_of_sub_matrix(meta_data_set &ms, int id, recs *rec = nullptr);

// ---- padding ------------------------------------------------------------------------------------
// pad every non-empty row to a multiple of m with (same row, last col of the row, 0)
// modify_{col_indices,vals,row_indices}_by_col_pad_in_sub_matrix.cc; the reference runs the three
// one after another on the same (unpadded) row array; here one pass does all three.
void modify_by_col_pad_in_sub_matrix(meta_data_set &ms, int id, uint64_t multiple, double padding_rate_up_bound, recs *rec = nullptr);
// pad the nnz stream tail to a multiple of target with (last row, last col, 0): modify_*_by_nnz_pad.cc
void modify_by_nnz_pad(meta_data_set &ms, int id, uint64_t nnz_target, double padding_rate_up_bound, recs *rec = nullptr);
// append rows so that the row count is a multiple: modify_*_by_row_pad_in_sub_matrix.cc
void modify_by_row_pad_in_sub_matrix(meta_data_set &ms, int id, uint64_t multiple, double padding_rate_up_bound, recs *rec = nullptr);
// pad each row of a parent block (GLOBAL = whole sub-matrix) to the block's max row length:
// modify_*_by_col_pad_parent_blk_to_max_row_size.cc
void modify_by_col_pad_parent_blk_to_max_row_size(meta_data_set &ms, POS_TYPE parent_pos, int id, bool padding_with_empty_row,
                                                  double padding_rate_up_bound, recs *rec = nullptr);
// one explicit zero for every empty row after a non-zero: modify_*_by_empty_pad_in_submatrix.cc
void modify_by_empty_pad_in_submatrix(meta_data_set &ms, int id, recs *rec = nullptr);

// ---- fixed blocking, row direction --------------------------------------------------------------
// top level (no parent): first_row_indices=[0,b,2b,..,rows], first_nz_indices = prefix nnz (both with sentinel)
//   BMTB: get_begin_{rows,nzs}_of_BMTBs_after_fixed_blocking_in_row_direction.cc
//   BMW : ..._of_BMW_after_fixed_blocking_in_row_direction_without_BMTB.cc
//   BMT : ..._of_BMT_after_fixed_blocking_in_row_direction.cc
void get_begin_rows_after_fixed_blocking_in_row_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t row_block, recs *rec = nullptr);
void get_begin_nzs_after_fixed_blocking_in_row_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t row_block, recs *rec = nullptr);
// inside a parent level (the stride restarts at every parent block): ..._in_BMTB.cc / ..._in_BMW.cc
void get_begin_rows_after_fixed_blocking_in_row_direction_in_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t row_block, recs *rec = nullptr);
void get_begin_nzs_after_fixed_blocking_in_row_direction_in_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t row_block, recs *rec = nullptr);
// ..._relative_to_BMTB.cc / ..._relative_to_BMW.cc (no sentinel)
void get_begin_rows_after_fixed_blocking_in_row_direction_relative_to_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t row_block, recs *rec = nullptr);
void get_begin_nzs_after_fixed_blocking_in_row_direction_relative_to_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t row_block, recs *rec = nullptr);
// first child index of every parent by row ranges: get_begin_BMWs_of_BMTB_after_blocking_in_row_direction.cc,
// get_begin_BMTs_of_specific_parent_after_blocking_in_row_direction.cc
void get_begin_children_of_parent_after_blocking_in_row_direction(meta_data_set &ms, POS_TYPE child, POS_TYPE parent, int id, recs *rec = nullptr);
// ... by nz ranges: get_begin_BMTs_of_specific_parent_after_blocking.cc, get_begin_BMWs_of_BMTB_after_blocking.cc
void get_begin_children_of_parent_after_blocking(meta_data_set &ms, POS_TYPE child, POS_TYPE parent, int id, recs *rec = nullptr);

// ---- fixed blocking, col direction (row chunks of <= c nnz) --------------------------------------
// get_begin_{rows,nzs}_of_{BMT,BMW,BMTB}_after_fixed_blocking_in_col_direction.cc
void get_begin_rows_after_fixed_blocking_in_col_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t col_size, recs *rec = nullptr);
void get_begin_nzs_after_fixed_blocking_in_col_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t col_size, recs *rec = nullptr);
void get_begin_rows_after_fixed_blocking_in_col_direction_relative_to_parent(meta_data_set &ms, POS_TYPE parent, int id, uint64_t col_size, recs *rec = nullptr);
void get_begin_nzs_after_fixed_blocking_in_col_direction_relative_to_parents(meta_data_set &ms, POS_TYPE parent, int id, uint64_t col_size, recs *rec = nullptr);

// ---- fixed blocking, nnz direction (every k entries regardless of rows) --------------------------
// get_begin_{rows,nzs}_of_{BMT,BMW,BMTB}_after_fixed_blocking_in_nnz_direction.cc
void get_begin_rows_after_fixed_blocking_in_nnz_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t nnz_per_blk, recs *rec = nullptr);
void get_begin_nzs_after_fixed_blocking_in_nnz_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t nnz_per_blk, recs *rec = nullptr);
// get_begin_{rows,nzs}_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB.cc (no sentinel; the parent index moves on by
// at most one block per BMW, as in the reference)
void get_begin_rows_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB(meta_data_set &ms, int id, uint64_t nnz_per_blk, recs *rec = nullptr);
void get_begin_nzs_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB(meta_data_set &ms, int id, uint64_t nnz_per_blk, recs *rec = nullptr);

// ---- sizes ---------------------------------------------------------------------------------------
// get_BMT_size_of_each_parent.cc: emits <parent>/BMT_size_of_each_blk only when all BMTs of a parent have
// the same nnz; returns silently (no output) otherwise.
void get_BMT_size_of_each_parent(meta_data_set &ms, POS_TYPE parent_pos, int id, bool row_direction_blocking, recs *rec = nullptr);
// get_BMW_size_of_each_parent.cc with parent GLOBAL_META: GLOBAL/BMW_size_of_each_blk = [size] when all BMWs hold the same nnz,
// nothing otherwise.  get_BMTB_size.cc: GLOBAL/BMTB_size_of_each_blk = [size]; BMTBs of different sizes are an error there.
void get_BMW_size_of_global(meta_data_set &ms, int id, recs *rec = nullptr);
void get_BMTB_size(meta_data_set &ms, int id, recs *rec = nullptr);

// ---- bit maps and segments (thread_bit_map.cc, parent_bit_map_of_thread.cc, segment_*.cc) ----------
void thread_bit_map(meta_data_set &ms, bool parent_flag, uint64_t parent_size, int id, recs *rec = nullptr);
void parent_bit_map_of_thread(meta_data_set &ms, POS_TYPE pos, int id, uint64_t vector_width, recs *rec = nullptr);
void segment_empty_flag(meta_data_set &ms, POS_TYPE pos, uint64_t size, int id, recs *rec = nullptr);
void segment_empty_row_indices(meta_data_set &ms, POS_TYPE pos, int id, recs *rec = nullptr);
void segment_offset(meta_data_set &ms, bool parent_flag, uint64_t size, int id, recs *rec = nullptr);
void segment_ptr(meta_data_set &ms, POS_TYPE pos, int id, recs *rec = nullptr);

// ---- merge-thread grouping (get_begin_*_after_merge_thread.cc) -------------------------------------
void get_begin_rows_after_merge_thread(meta_data_set &ms, POS_TYPE pos, uint64_t merge_num, int id, recs *rec = nullptr);
void get_begin_nzs_after_merge_thread(meta_data_set &ms, POS_TYPE pos, uint64_t merge_num, int id, recs *rec = nullptr);
void get_begin_rows_relative_to_parent_after_merge_thread(meta_data_set &ms, POS_TYPE pos, uint64_t merge_num, int id, recs *rec = nullptr);
void get_begin_nzs_relative_to_parent_after_merge_thread(meta_data_set &ms, POS_TYPE pos, uint64_t merge_num, int id, recs *rec = nullptr);
void get_begin_BMTs_after_merge_thread(meta_data_set &ms, POS_TYPE pos, uint64_t merge_num, int id, recs *rec = nullptr);

// ---- balanced (nnz) blocking in row direction (data_transform_common.cc:934-988) -------------------
std::vector<uint64_t> balanced_begin_rows(const std::vector<uint64_t> &nnz_of_each_row, uint64_t nnz_per_interval);
std::vector<uint64_t> balanced_begin_nzs(const std::vector<uint64_t> &nnz_of_each_row, uint64_t nnz_per_interval);
void get_begin_rows_after_nnz_blocking_in_row_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t nnz_per_interval, recs *rec = nullptr);
void get_begin_nzs_after_nnz_blocking_in_row_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t nnz_per_interval, recs *rec = nullptr);

// ---- merge path (get_begin_{rows,nzs}_of_level_after_merge_path.cc), linear two-pointer walk --------
void get_begin_rows_of_level_after_merge_path(meta_data_set &ms, POS_TYPE pos, int id, uint64_t work_size, recs *rec = nullptr);
void get_begin_nzs_of_level_after_merge_path(meta_data_set &ms, POS_TYPE pos, int id, uint64_t work_size, recs *rec = nullptr);

// ---- child levels relative to their parent, one level up / other directions -------------------------
void get_begin_rows_of_BMW_after_fixed_blocking_in_col_direction_relative_to_BMTB(meta_data_set &ms, int id, uint64_t col_size, recs *rec = nullptr);
void get_begin_nzs_of_BMW_after_fixed_blocking_in_col_direction_relative_to_BMTB(meta_data_set &ms, int id, uint64_t col_size, recs *rec = nullptr);
void get_begin_rows_of_BMT_after_fixed_blocking_in_nnz_direction_relative_to_parent(meta_data_set &ms, POS_TYPE parent, int id, uint64_t nnz_per_blk, recs *rec = nullptr);
void get_begin_nzs_of_BMT_after_fixed_blocking_in_nnz_direction_relative_to_parent(meta_data_set &ms, POS_TYPE parent, int id, uint64_t nnz_per_blk, recs *rec = nullptr);
void get_BMW_size_of_each_parent(meta_data_set &ms, POS_TYPE parent, int id, recs *rec = nullptr);
// balanced (nnz) blocking inside a parent level: get_begin_{rows,nzs}_of_{BMT,BMW}_after_nnz_blocking_in_row_direction_{in,relative_to}_{BMW,BMTB}.cc
void get_begin_rows_after_nnz_blocking_in_row_direction_in_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t nnz_per_interval, bool relative, recs *rec = nullptr);
void get_begin_nzs_after_nnz_blocking_in_row_direction_in_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t nnz_per_interval, bool relative, recs *rec = nullptr);

// ---- sub-matrix division (the 7 steps of fixed_interval_row_matrix_div_operator / row_nz_matrix_div_operator) --------
void fixed_div_in_row_direction(meta_data_set &ms, int id, uint64_t fixed_row_gap_size, recs *rec = nullptr);
std::vector<uint64_t> div_positions_by_row_nnz(const std::vector<uint64_t> &nnz_of_each_row, uint64_t nz_gap, uint64_t max_gap, uint64_t expansion);
void div_according_to_row_nz(meta_data_set &ms, int id, uint64_t nz_gap_size, uint64_t max_gap, uint64_t expansion_rate, recs *rec = nullptr);

// ---- interleave (modify_*_by_interlance_storage.cc) -------------------------------------------------
void modify_by_interlance_storage(meta_data_set &ms, POS_TYPE parent_pos, int id, recs *rec = nullptr);

void remove_item_of_metadata(meta_data_set &ms, int id, const std::string &name, POS_TYPE pos, recs *rec = nullptr);

// the reference's step-class names covered by the functions above (for docs/tests)
const std::vector<std::string> &covered_reference_step_names();

} // namespace steps
} // namespace gs
