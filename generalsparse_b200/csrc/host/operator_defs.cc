// Concrete operators (operator.hpp:268-1297, operator/*.cc).  Constructor argument order follows the
// reference's code_generator flavour; args arrive as int64 through make_operator().
#include "operators.hpp"

#include "../../../include/gsparse_b200.h"

#include <algorithm>

namespace gs {

namespace {

using opv = std::vector<std::shared_ptr<basic_operator>>;

bool any_name_has(const opv &ops, const char *needle)
{
    for (const auto &o : ops)
        if (o->get_name().find(needle) != std::string::npos) return true;
    return false;
}

void append(std::vector<std::string> &dst, const steps::recs &src) { dst.insert(dst.end(), src.begin(), src.end()); }

// shared by the row-direction blocking operators: drop a parent level's arrays, re-run the operators
// that produced them without their own padding (fixed_interval_row_direction_thread_blocking_operator.cc:258-269)
void drop_level(meta_data_set &ms, POS_TYPE pos, int id, steps::recs &rec)
{
    for (const std::string &n : ms.all_item_of_metadata_of_diff_pos(pos, id)) steps::remove_item_of_metadata(ms, id, n, pos, &rec);
}

// after a child operator padded the matrix, the parent levels are derived again without their own padding; the steps of
// the re-run land in the child's record sequence (fixed_interval_col_direction_thread_blocking_operator.cc run())
void rerun_former(const opv &former, bool check, steps::recs &into)
{
    for (const auto &o : former) {
        size_t before = o->get_data_transform_sequence().size();
        o->set_padding_to_false();
        o->run(check);
        const auto &seq = o->get_data_transform_sequence();
        into.insert(into.end(), seq.begin() + before, seq.end());
    }
}

// ------------------------------------------------------------------------------------------- CONVERTING
class sort_operator : public basic_operator { // operator.hpp:268, operator/sort_operator.cc:70-108
  public:
    explicit sort_operator(std::shared_ptr<code_generator> cg) : basic_operator("sort_operator", std::move(cg), CONVERTING_OP) {}
    bool is_valid_according_to_metadata() override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        return ms.is_exist(GLOBAL_META, "nz_row_indices", id) && ms.is_exist(GLOBAL_META, "nz_col_indices", id) &&
               ms.is_exist(GLOBAL_META, "nz_vals", id) && !ms.is_exist(GLOBAL_META, "original_nz_row_indices", id);
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // sorting is only legal before any distribution (sort_operator.cc is_valid_according_to_operator)
        return h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id).empty() &&
               h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty();
    }
    void run(bool check) override
    {
        if (check) GS_REQUIRE(is_valid_according_to_metadata(), "sort_operator: metadata not sortable (already sorted?)");
        steps::recs r;
        steps::get_row_order_by_length(*meta_data_set_ptr, target_matrix_id, &r);
        steps::reorder_val_by_index(*meta_data_set_ptr, target_matrix_id, &r);
        steps::reorder_col_by_index(*meta_data_set_ptr, target_matrix_id, &r);
        steps::reorder_row_by_index(*meta_data_set_ptr, target_matrix_id, &r);
        steps::remove_empty_row_in_end_of_sub_matrix(*meta_data_set_ptr, target_matrix_id, &r);
        append(transform_seq, r);
        is_run = true;
    }
};

// ----------------------------------------------------------------------------------------- DISTRIBUTING
class fixed_interval_row_direction_tblock_blocking_operator : public basic_operator { // operator.hpp:356
  public:
    fixed_interval_row_direction_tblock_blocking_operator(std::shared_ptr<code_generator> cg, int b, bool is_padding)
        : basic_operator("fixed_interval_row_direction_tblock_blocking_operator", std::move(cg), DISTRIBUTING_OP), b_(b), pad_(is_padding)
    {
        GS_REQUIRE(b > 0, "fixed_row_block_size must be > 0");
        params = ",fixed_row_block_size:" + std::to_string(b);
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        return h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id).empty() &&
               h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty();
    }
    void set_padding_to_false() override { pad_ = false; }
    void run(bool) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        steps::recs r;
        if (pad_) steps::modify_by_row_pad_in_sub_matrix(ms, id, b_, pad_bound(), &r);
        steps::get_begin_rows_after_fixed_blocking_in_row_direction(ms, TBLOCK_META, id, b_, &r);
        steps::get_begin_nzs_after_fixed_blocking_in_row_direction(ms, TBLOCK_META, id, b_, &r);
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(TBLOCK_META);
        is_run = true;
    }

  private:
    uint64_t b_;
    bool pad_;
};

class fixed_interval_row_direction_warp_blocking_operator : public basic_operator { // operator.hpp:398
  public:
    fixed_interval_row_direction_warp_blocking_operator(std::shared_ptr<code_generator> cg, int b, bool row_rel, bool nz_rel, bool is_padding)
        : basic_operator("fixed_interval_row_direction_warp_blocking_operator", std::move(cg), DISTRIBUTING_OP), b_(b), row_rel_(row_rel),
          nz_rel_(nz_rel), pad_(is_padding)
    {
        GS_REQUIRE(b > 0, "fixed_row_block_size must be > 0");
        if (is_padding) GS_REQUIRE(!row_rel && !nz_rel, "row padding excludes relative indices");
        params = ",fixed_row_block_size:" + std::to_string(b) + ",is_padding:" + std::to_string((int)is_padding);
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        return h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty() && !any_name_has(d, "warp") &&
               !any_name_has(d, "thread") && !any_name_has(d, "col_direction") && !any_name_has(d, "interlance");
    }
    void set_padding_to_false() override { pad_ = false; }
    void run(bool) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        steps::recs r;
        // is_valid_according_to_metadata (fixed_interval_row_direction_warp_blocking_operator.cc:154-157): row padding only
        // while no thread-block level exists -- it would shift rows under the blocks already cut
        GS_REQUIRE(!(pad_ && ms.count_of_metadata_of_diff_pos(TBLOCK_META, id) != 0), "warp row padding after a tblock level exists");
        if (ms.is_exist(TBLOCK_META, "first_row_indices", id)) {
            steps::get_begin_rows_after_fixed_blocking_in_row_direction_in_parent(ms, WARP_META, TBLOCK_META, id, b_, &r);
            if (row_rel_) steps::get_begin_rows_after_fixed_blocking_in_row_direction_relative_to_parent(ms, WARP_META, TBLOCK_META, id, b_, &r);
            steps::get_begin_nzs_after_fixed_blocking_in_row_direction_in_parent(ms, WARP_META, TBLOCK_META, id, b_, &r);
            if (nz_rel_) steps::get_begin_nzs_after_fixed_blocking_in_row_direction_relative_to_parent(ms, WARP_META, TBLOCK_META, id, b_, &r);
            steps::get_begin_children_of_parent_after_blocking_in_row_direction(ms, WARP_META, TBLOCK_META, id, &r);
        } else {
            GS_REQUIRE(!row_rel_ && !nz_rel_, "relative indices need a BMTB level");
            if (pad_) steps::modify_by_row_pad_in_sub_matrix(ms, id, b_, pad_bound(), &r);
            steps::get_begin_rows_after_fixed_blocking_in_row_direction(ms, WARP_META, id, b_, &r);
            steps::get_begin_nzs_after_fixed_blocking_in_row_direction(ms, WARP_META, id, b_, &r);
        }
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(WARP_META);
        is_run = true;
    }

  private:
    uint64_t b_;
    bool row_rel_, nz_rel_, pad_;
};

class fixed_interval_row_direction_thread_blocking_operator : public basic_operator { // operator.hpp:598
  public:
    fixed_interval_row_direction_thread_blocking_operator(std::shared_ptr<code_generator> cg, int b, bool row_rel, bool nz_rel, bool is_row_padding,
                                                          bool pad_to_row_max, bool pad_to_col_size, int col_size,
                                                          std::shared_ptr<operator_context> history)
        : basic_operator("fixed_interval_row_direction_thread_blocking_operator", std::move(cg), DISTRIBUTING_OP), b_(b), row_rel_(row_rel),
          nz_rel_(nz_rel), row_pad_(is_row_padding), max_pad_(pad_to_row_max), col_pad_(pad_to_col_size), col_size_(col_size)
    {
        GS_REQUIRE(b > 0, "fixed_row_block_size must be > 0");
        if (is_row_padding) GS_REQUIRE(!row_rel && !nz_rel, "row padding excludes relative indices");
        former_ = history->read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        params = ",fixed_row_block_size:" + std::to_string(b) + ",is_row_padding:" + std::to_string((int)is_row_padding);
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // fixed_interval_row_direction_thread_blocking_operator.cc: no implementing operator yet, and no earlier
        // thread / col-direction / interleave operator (balanced_interval only without max padding)
        opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        if (!h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty()) return false;
        bool balanced = any_name_has(d, "balanced_interval") && max_pad_;
        return !any_name_has(d, "thread") && !any_name_has(d, "col") && !any_name_has(d, "interlance") && !balanced;
    }
    void set_padding_to_false() override { row_pad_ = max_pad_ = col_pad_ = false; }
    bool get_is_col_padding_with_row_max_size_with_empty_row() const override { return max_pad_; }
    void run(bool check) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        steps::recs r;
        bool has_tb = ms.is_exist(TBLOCK_META, "first_row_indices", id), has_w = ms.is_exist(WARP_META, "first_row_indices", id);
        if (has_w || has_tb) {
            POS_TYPE parent = has_w ? WARP_META : TBLOCK_META;
            if (max_pad_) {
                steps::modify_by_col_pad_parent_blk_to_max_row_size(ms, parent, id, true, pad_bound(), &r);
                if (has_w) drop_level(ms, WARP_META, id, r);
                if (has_tb) drop_level(ms, TBLOCK_META, id, r);
                rerun_former(former_, check, r);
            }
            if (col_pad_) {
                steps::modify_by_col_pad_in_sub_matrix(ms, id, col_size_, pad_bound(), &r);
                if (has_w) drop_level(ms, WARP_META, id, r);
                if (has_tb) drop_level(ms, TBLOCK_META, id, r);
                rerun_former(former_, check, r);
            }
            steps::get_begin_rows_after_fixed_blocking_in_row_direction_in_parent(ms, THREAD_META, parent, id, b_, &r);
            if (row_rel_) steps::get_begin_rows_after_fixed_blocking_in_row_direction_relative_to_parent(ms, THREAD_META, parent, id, b_, &r);
            steps::get_begin_nzs_after_fixed_blocking_in_row_direction_in_parent(ms, THREAD_META, parent, id, b_, &r);
            if (nz_rel_) steps::get_begin_nzs_after_fixed_blocking_in_row_direction_relative_to_parent(ms, THREAD_META, parent, id, b_, &r);
            steps::get_begin_children_of_parent_after_blocking_in_row_direction(ms, THREAD_META, parent, id, &r);
            if (max_pad_ && b_ == 1) steps::get_BMT_size_of_each_parent(ms, parent, id, true, &r);
        } else {
            GS_REQUIRE(!row_rel_ && !nz_rel_, "relative indices need a parent level");
            if (row_pad_) steps::modify_by_row_pad_in_sub_matrix(ms, id, b_, pad_bound(), &r);
            if (max_pad_) steps::modify_by_col_pad_parent_blk_to_max_row_size(ms, GLOBAL_META, id, true, pad_bound(), &r);
            if (col_pad_) steps::modify_by_col_pad_in_sub_matrix(ms, id, col_size_, pad_bound(), &r);
            steps::get_begin_rows_after_fixed_blocking_in_row_direction(ms, THREAD_META, id, b_, &r);
            steps::get_begin_nzs_after_fixed_blocking_in_row_direction(ms, THREAD_META, id, b_, &r);
            if (max_pad_ && b_ == 1) steps::get_BMT_size_of_each_parent(ms, GLOBAL_META, id, true, &r);
        }
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(THREAD_META);
        code_generator_ptr->set_thread_for_row(true);
        is_run = true;
    }

  private:
    uint64_t b_;
    bool row_rel_, nz_rel_, row_pad_, max_pad_, col_pad_;
    uint64_t col_size_;
    opv former_;
};

class fixed_interval_col_direction_thread_blocking_operator : public basic_operator { // operator.hpp:477
  public:
    fixed_interval_col_direction_thread_blocking_operator(std::shared_ptr<code_generator> cg, int col_block, bool row_rel_bmtb, bool nz_rel_bmtb,
                                                          bool pad_to_col_size, bool pad_to_row_max, std::shared_ptr<operator_context> history)
        : basic_operator("fixed_interval_col_direction_thread_blocking_operator", std::move(cg), DISTRIBUTING_OP), c_(col_block),
          row_rel_(row_rel_bmtb), nz_rel_(nz_rel_bmtb), col_pad_(pad_to_col_size), max_pad_(pad_to_row_max)
    {
        GS_REQUIRE(col_block > 0, "fixed_col_block_size must be > 0");
        former_ = history->read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        params = ",fixed_col_block_size:" + std::to_string(col_block) + ",is_padding:" + std::to_string((int)(pad_to_col_size || pad_to_row_max));
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        return h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty() && !any_name_has(d, "thread") && !any_name_has(d, "interlance");
    }
    void set_padding_to_false() override { col_pad_ = max_pad_ = false; }
    bool get_is_padding_with_col_size_in_bmt() const override { return col_pad_; }
    void run(bool check) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        steps::recs r;
        bool has_tb = ms.is_exist(TBLOCK_META, "first_row_indices", id), has_w = ms.is_exist(WARP_META, "first_row_indices", id);
        if (max_pad_) steps::modify_by_col_pad_parent_blk_to_max_row_size(ms, has_w ? WARP_META : (has_tb ? TBLOCK_META : GLOBAL_META), id, false, pad_bound(), &r);
        if (col_pad_) steps::modify_by_col_pad_in_sub_matrix(ms, id, c_, pad_bound(), &r);
        if (col_pad_ || max_pad_) {
            if (has_tb) drop_level(ms, TBLOCK_META, id, r);
            if (has_w) drop_level(ms, WARP_META, id, r);
            rerun_former(former_, check, r);
        }
        steps::get_begin_rows_after_fixed_blocking_in_col_direction(ms, THREAD_META, id, c_, &r);
        steps::get_begin_nzs_after_fixed_blocking_in_col_direction(ms, THREAD_META, id, c_, &r);
        // relative to the innermost parent an earlier operator built: BMW if a warp operator ran, else BMTB (ctor, :43-82)
        const bool w_former = any_name_has(former_, "warp"), tb_former = !w_former && any_name_has(former_, "tblock");
        if (tb_former && row_rel_) steps::get_begin_rows_after_fixed_blocking_in_col_direction_relative_to_parent(ms, TBLOCK_META, id, c_, &r);
        if (tb_former && nz_rel_) steps::get_begin_nzs_after_fixed_blocking_in_col_direction_relative_to_parents(ms, TBLOCK_META, id, c_, &r);
        if (has_tb) {
            steps::get_begin_children_of_parent_after_blocking(ms, THREAD_META, TBLOCK_META, id, &r);
            if (col_pad_) steps::get_BMT_size_of_each_parent(ms, TBLOCK_META, id, false, &r);
        }
        if (w_former && row_rel_) steps::get_begin_rows_after_fixed_blocking_in_col_direction_relative_to_parent(ms, WARP_META, id, c_, &r);
        if (w_former && nz_rel_) steps::get_begin_nzs_after_fixed_blocking_in_col_direction_relative_to_parents(ms, WARP_META, id, c_, &r);
        if (has_w) {
            steps::get_begin_children_of_parent_after_blocking(ms, THREAD_META, WARP_META, id, &r);
            if (col_pad_) steps::get_BMT_size_of_each_parent(ms, WARP_META, id, false, &r);
        }
        if (col_pad_) steps::get_BMT_size_of_each_parent(ms, GLOBAL_META, id, false, &r);
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(THREAD_META);
        is_run = true;
    }

  private:
    uint64_t c_;
    bool row_rel_, nz_rel_, col_pad_, max_pad_;
    opv former_;
};

class fixed_interval_nnz_direction_thread_blocking_operator : public basic_operator { // operator.hpp:906
  public:
    fixed_interval_nnz_direction_thread_blocking_operator(std::shared_ptr<code_generator> cg, int nnz_per_bmt, bool row_rel, bool nz_rel, bool nnz_padding)
        : basic_operator("fixed_interval_nnz_direction_thread_blocking_operator", std::move(cg), DISTRIBUTING_OP), k_(nnz_per_bmt), row_rel_(row_rel),
          nz_rel_(nz_rel), pad_(nnz_padding)
    {
        GS_REQUIRE(nnz_per_bmt > 0, "nnz_per_BMT must be > 0");
        params = ",nnz_per_BMT:" + std::to_string(nnz_per_bmt);
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        return h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty() && !any_name_has(d, "thread") && !any_name_has(d, "interlance");
    }
    void set_padding_to_false() override { pad_ = false; }
    void run(bool) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        steps::recs r;
        bool has_tb = ms.is_exist(TBLOCK_META, "first_row_indices", id), has_w = ms.is_exist(WARP_META, "first_row_indices", id);
        if (pad_) steps::modify_by_nnz_pad(ms, id, k_, pad_bound(), &r);
        steps::get_begin_rows_after_fixed_blocking_in_nnz_direction(ms, THREAD_META, id, k_, &r);
        // relative to the innermost parent level: BMW, else BMTB (fixed_interval_nnz_direction_thread_blocking_operator.cc run())
        GS_REQUIRE(!(row_rel_ || nz_rel_) || has_w || has_tb, "relative indices need a parent level");
        if (row_rel_) steps::get_begin_rows_of_BMT_after_fixed_blocking_in_nnz_direction_relative_to_parent(ms, has_w ? WARP_META : TBLOCK_META, id, k_, &r);
        steps::get_begin_nzs_after_fixed_blocking_in_nnz_direction(ms, THREAD_META, id, k_, &r);
        if (nz_rel_) steps::get_begin_nzs_of_BMT_after_fixed_blocking_in_nnz_direction_relative_to_parent(ms, has_w ? WARP_META : TBLOCK_META, id, k_, &r);
        if (has_w) steps::get_begin_children_of_parent_after_blocking(ms, THREAD_META, WARP_META, id, &r);
        else if (has_tb) steps::get_begin_children_of_parent_after_blocking(ms, THREAD_META, TBLOCK_META, id, &r);
        steps::get_BMT_size_of_each_parent(ms, GLOBAL_META, id, false, &r);
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(THREAD_META);
        is_run = true;
    }

  private:
    uint64_t k_;
    bool row_rel_, nz_rel_, pad_;
};

class fixed_interval_nnz_direction_tblock_blocking_operator : public basic_operator { // operator/fixed_interval_nnz_direction_tblock_blocking_operator.cc
  public:
    fixed_interval_nnz_direction_tblock_blocking_operator(std::shared_ptr<code_generator> cg, int nnz_per_bmtb, bool nnz_padding)
        : basic_operator("fixed_interval_nnz_direction_tblock_blocking_operator", std::move(cg), DISTRIBUTING_OP), k_(nnz_per_bmtb), pad_(nnz_padding)
    {
        GS_REQUIRE(nnz_per_bmtb > 0, "nnz_per_BMTB must be > 0");
        params = ",nnz_per_BMTB:" + std::to_string(nnz_per_bmtb);
    }
    uint64_t get_nnz_per_BMTB() const { return k_; }
    bool is_valid_according_to_operator(const operator_context &h) override
    { // :38-55: the first distributing operator, nothing implemented yet
        return h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty() && h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id).empty();
    }
    bool is_valid_according_to_metadata() override
    { // :57-94: no level opened yet, no interleaved storage
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        return !ms.is_exist(TBLOCK_META, "first_row_indices", id) && !ms.is_exist(WARP_META, "first_row_indices", id) && !ms.is_exist(THREAD_META, "first_row_indices", id) &&
               !ms.is_exist(GLOBAL_META, "nz_row_indices_after_interlance_storage", id);
    }
    void set_padding_to_false() override { pad_ = false; }
    void run(bool) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        steps::recs r;
        if (pad_) steps::modify_by_nnz_pad(ms, id, k_, pad_bound(), &r);
        steps::get_begin_rows_after_fixed_blocking_in_nnz_direction(ms, TBLOCK_META, id, k_, &r);
        steps::get_begin_nzs_after_fixed_blocking_in_nnz_direction(ms, TBLOCK_META, id, k_, &r);
        if (pad_) steps::get_BMTB_size(ms, id, &r);
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(TBLOCK_META);
        is_run = true;
    }

  private:
    uint64_t k_;
    bool pad_;
};

class fixed_interval_nnz_direction_warp_blocking_operator : public basic_operator { // operator/fixed_interval_nnz_direction_warp_blocking_operator.cc
  public:
    fixed_interval_nnz_direction_warp_blocking_operator(std::shared_ptr<code_generator> cg, int nnz_per_bmw, bool row_rel, bool nz_rel, bool nnz_padding)
        : basic_operator("fixed_interval_nnz_direction_warp_blocking_operator", std::move(cg), DISTRIBUTING_OP), k_(nnz_per_bmw), row_rel_(row_rel), nz_rel_(nz_rel),
          pad_(nnz_padding)
    {
        GS_REQUIRE(nnz_per_bmw > 0, "nnz_per_BMW must be > 0");
        params = ",nnz_per_BMW:" + std::to_string(nnz_per_bmw);
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    { // :44-92: only nnz-direction tblock blocking may precede, and its interval must be a multiple of this one
        if (!h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty()) return false;
        for (const auto &op : h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id)) {
            auto tb = std::dynamic_pointer_cast<fixed_interval_nnz_direction_tblock_blocking_operator>(op);
            if (!tb || tb->get_nnz_per_BMTB() % k_ != 0) return false;
        }
        return true;
    }
    bool is_valid_according_to_metadata() override
    { // :94-143: relative indices need the tblock level, padding forbids it
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        const bool has_tb = ms.is_exist(TBLOCK_META, "first_row_indices", id);
        if (ms.is_exist(WARP_META, "first_row_indices", id) || ms.is_exist(THREAD_META, "first_row_indices", id)) return false;
        if ((row_rel_ || nz_rel_) && !has_tb) return false;
        if (pad_ && has_tb) return false;
        return !ms.is_exist(GLOBAL_META, "nz_row_indices_after_interlance_storage", id);
    }
    void set_padding_to_false() override { pad_ = false; }
    void run(bool) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        steps::recs r;
        const bool has_tb = ms.is_exist(TBLOCK_META, "first_row_indices", id);
        if (pad_) steps::modify_by_nnz_pad(ms, id, k_, pad_bound(), &r);
        steps::get_begin_rows_after_fixed_blocking_in_nnz_direction(ms, WARP_META, id, k_, &r);
        steps::get_begin_nzs_after_fixed_blocking_in_nnz_direction(ms, WARP_META, id, k_, &r);
        if (row_rel_) steps::get_begin_rows_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB(ms, id, k_, &r);
        if (nz_rel_) steps::get_begin_nzs_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB(ms, id, k_, &r);
        if (has_tb) steps::get_begin_children_of_parent_after_blocking(ms, WARP_META, TBLOCK_META, id, &r);
        steps::get_BMW_size_of_global(ms, id, &r);
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(WARP_META);
        is_run = true;
    }

  private:
    uint64_t k_;
    bool row_rel_, nz_rel_, pad_;
};

// balanced_interval_row_direction_{tblock,warp,thread}_blocking_operator (operator.hpp:871,823,777; operator/balanced_interval_*.cc)
class balanced_interval_row_direction_blocking_operator : public basic_operator {
  public:
    balanced_interval_row_direction_blocking_operator(std::shared_ptr<code_generator> cg, POS_TYPE pos, int nnz_per_interval, bool row_rel, bool nz_rel)
        : basic_operator(std::string("balanced_interval_row_direction_") + (pos == TBLOCK_META ? "tblock" : pos == WARP_META ? "warp" : "thread") +
                             "_blocking_operator", std::move(cg), DISTRIBUTING_OP), pos_(pos), per_(nnz_per_interval), row_rel_(row_rel), nz_rel_(nz_rel)
    {
        GS_REQUIRE(nnz_per_interval > 0, "nnz_per_interval must be > 0");
        params = ",nnz_per_interval:" + std::to_string(nnz_per_interval);
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // nothing implemented yet; tblock: the first distributing operator; warp: no warp / thread / col / interlance / nnz_direction operator
        // before it; thread: no thread / col / interlance / nnz_direction operator before it
        if (!h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty()) return false;
        const opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        if (pos_ == TBLOCK_META) return d.empty();
        if (any_name_has(d, "thread") || any_name_has(d, "col") || any_name_has(d, "interlance") || any_name_has(d, "nnz_direction")) return false;
        return pos_ == THREAD_META || !any_name_has(d, "warp");
    }
    void run(bool) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        steps::recs r;
        const bool has_tb = ms.is_exist(TBLOCK_META, "first_row_indices", id), has_w = ms.is_exist(WARP_META, "first_row_indices", id);
        const POS_TYPE parent = pos_ == THREAD_META && has_w ? WARP_META : (pos_ != TBLOCK_META && has_tb) ? TBLOCK_META : NONE_META;
        GS_REQUIRE(ms.count_of_metadata_of_diff_pos(pos_, id) == 0, "balanced blocking: the level already has metadata");
        if (parent != NONE_META) {
            steps::get_begin_rows_after_nnz_blocking_in_row_direction_in_parent(ms, pos_, parent, id, per_, false, &r);
            if (row_rel_) steps::get_begin_rows_after_nnz_blocking_in_row_direction_in_parent(ms, pos_, parent, id, per_, true, &r);
            steps::get_begin_nzs_after_nnz_blocking_in_row_direction_in_parent(ms, pos_, parent, id, per_, false, &r);
            if (nz_rel_) steps::get_begin_nzs_after_nnz_blocking_in_row_direction_in_parent(ms, pos_, parent, id, per_, true, &r);
            steps::get_begin_children_of_parent_after_blocking_in_row_direction(ms, pos_, parent, id, &r);
        } else {
            GS_REQUIRE(!row_rel_ && !nz_rel_, "relative indices need a parent level");
            steps::get_begin_rows_after_nnz_blocking_in_row_direction(ms, pos_, id, per_, &r);
            steps::get_begin_nzs_after_nnz_blocking_in_row_direction(ms, pos_, id, per_, &r);
        }
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(pos_);
        if (pos_ == THREAD_META) code_generator_ptr->set_thread_for_row(true);
        is_run = true;
    }

  private:
    POS_TYPE pos_;
    uint64_t per_;
    bool row_rel_, nz_rel_;
};

// fixed_interval_col_direction_tblock_blocking_operator (operator.hpp:727): the first distributing operator of a chain
class fixed_interval_col_direction_tblock_blocking_operator : public basic_operator {
  public:
    fixed_interval_col_direction_tblock_blocking_operator(std::shared_ptr<code_generator> cg, int col_block, bool pad_to_col_size, bool pad_to_row_max)
        : basic_operator("fixed_interval_col_direction_tblock_blocking_operator", std::move(cg), DISTRIBUTING_OP), c_(col_block), col_pad_(pad_to_col_size), max_pad_(pad_to_row_max)
    {
        GS_REQUIRE(col_block > 0, "fixed_col_block_size must be > 0");
        params = ",fixed_col_block_size:" + std::to_string(col_block) + ",is_padding:" + std::to_string((int)(pad_to_col_size || pad_to_row_max));
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        return h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty() && h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id).empty();
    }
    void set_padding_to_false() override { col_pad_ = max_pad_ = false; }
    void run(bool) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        GS_REQUIRE(ms.count_of_metadata_of_diff_pos(TBLOCK_META, id) + ms.count_of_metadata_of_diff_pos(WARP_META, id) + ms.count_of_metadata_of_diff_pos(THREAD_META, id) == 0 &&
                       !ms.is_exist(GLOBAL_META, "nz_row_indices_after_interlance_storage", id),
                   "col-direction tblock blocking: a level or interleaved storage already exists");
        steps::recs r;
        if (max_pad_) steps::modify_by_col_pad_parent_blk_to_max_row_size(ms, GLOBAL_META, id, false, pad_bound(), &r);
        if (col_pad_) steps::modify_by_col_pad_in_sub_matrix(ms, id, c_, pad_bound(), &r);
        steps::get_begin_rows_after_fixed_blocking_in_col_direction(ms, TBLOCK_META, id, c_, &r);
        steps::get_begin_nzs_after_fixed_blocking_in_col_direction(ms, TBLOCK_META, id, c_, &r);
        if (col_pad_) steps::get_BMTB_size(ms, id, &r);
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(TBLOCK_META);
        is_run = true;
    }

  private:
    uint64_t c_;
    bool col_pad_, max_pad_;
};

// fixed_interval_col_direction_warp_blocking_operator (operator.hpp:652)
class fixed_interval_col_direction_warp_blocking_operator : public basic_operator {
  public:
    fixed_interval_col_direction_warp_blocking_operator(std::shared_ptr<code_generator> cg, int col_block, bool row_rel_bmtb, bool nz_rel_bmtb, bool pad_to_col_size, bool pad_to_row_max,
                                                        std::shared_ptr<operator_context> history)
        : basic_operator("fixed_interval_col_direction_warp_blocking_operator", std::move(cg), DISTRIBUTING_OP), c_(col_block), row_rel_(row_rel_bmtb), nz_rel_(nz_rel_bmtb),
          col_pad_(pad_to_col_size), max_pad_(pad_to_row_max)
    {
        GS_REQUIRE(col_block > 0, "fixed_col_block_size must be > 0");
        former_ = history->read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        pad_pos_ = any_name_has(former_, "tblock") ? TBLOCK_META : GLOBAL_META; // ctor, :38-54
        params = ",fixed_col_block_size:" + std::to_string(col_block) + ",is_padding:" + std::to_string((int)(pad_to_col_size || pad_to_row_max));
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        if (!h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty()) return false;
        const opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        const bool balanced = any_name_has(d, "balanced_interval") && max_pad_;
        return !any_name_has(d, "thread") && !any_name_has(d, "warp") && !any_name_has(d, "col") && !any_name_has(d, "interlance") && !balanced;
    }
    void set_padding_to_false() override { col_pad_ = max_pad_ = false; }
    void run(bool check) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        const bool has_tb = ms.is_exist(TBLOCK_META, "first_row_indices", id);
        GS_REQUIRE(ms.count_of_metadata_of_diff_pos(WARP_META, id) + ms.count_of_metadata_of_diff_pos(THREAD_META, id) == 0, "col-direction warp blocking: a warp or thread level already exists");
        GS_REQUIRE(!(row_rel_ || nz_rel_) || has_tb, "indices relative to BMTB need a tblock level");
        GS_REQUIRE(!(max_pad_ && pad_pos_ == TBLOCK_META) || has_tb, "padding to the BMTB's row maximum needs a tblock level");
        GS_REQUIRE(has_tb == !former_.empty(), "col-direction warp blocking: the tblock level and the operator history disagree");
        GS_REQUIRE(!ms.is_exist(GLOBAL_META, "nz_row_indices_after_interlance_storage", id), "interleaved storage already exists");
        steps::recs r;
        if (max_pad_) steps::modify_by_col_pad_parent_blk_to_max_row_size(ms, pad_pos_, id, false, pad_bound(), &r);
        if (col_pad_) steps::modify_by_col_pad_in_sub_matrix(ms, id, c_, pad_bound(), &r);
        if (col_pad_ || max_pad_) {
            if (has_tb) drop_level(ms, TBLOCK_META, id, r);
            rerun_former(former_, check, r);
        }
        steps::get_begin_rows_after_fixed_blocking_in_col_direction(ms, WARP_META, id, c_, &r);
        steps::get_begin_nzs_after_fixed_blocking_in_col_direction(ms, WARP_META, id, c_, &r);
        if (row_rel_) steps::get_begin_rows_of_BMW_after_fixed_blocking_in_col_direction_relative_to_BMTB(ms, id, c_, &r);
        if (nz_rel_) steps::get_begin_nzs_of_BMW_after_fixed_blocking_in_col_direction_relative_to_BMTB(ms, id, c_, &r);
        if (has_tb) steps::get_begin_children_of_parent_after_blocking(ms, WARP_META, TBLOCK_META, id, &r);
        if (col_pad_ && has_tb) steps::get_BMW_size_of_each_parent(ms, TBLOCK_META, id, &r);
        if (col_pad_) steps::get_BMW_size_of_each_parent(ms, GLOBAL_META, id, &r);
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(WARP_META);
        is_run = true;
    }

  private:
    uint64_t c_;
    bool row_rel_, nz_rel_, col_pad_, max_pad_;
    POS_TYPE pad_pos_ = GLOBAL_META;
    opv former_;
};

// interlance_storage_operator (operator.hpp:567, operator/interlance_storage_operator.cc)
class interlance_storage_operator : public basic_operator {
  public:
    interlance_storage_operator(std::shared_ptr<code_generator> cg, std::shared_ptr<operator_context> history)
        : basic_operator("interlance_storage_operator", std::move(cg), DISTRIBUTING_OP)
    {
        // the innermost parent level an earlier operator built (ctor, :9-38)
        const opv d = history->read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        bool warp = false, tblock = false;
        for (const auto &o : d) {
            if (o->get_name().find("warp") != std::string::npos) warp = true;
            else if (o->get_name().find("tblock") != std::string::npos) tblock = true;
        }
        pos_ = warp ? WARP_META : tblock ? TBLOCK_META : GLOBAL_META;
        code_generator_ptr->set_interleave_storage();
    }
    std::string convert_to_string() const override { return "interlance_storage_operator::{name:\"" + name + ",target_matrix_id:" + std::to_string(target_matrix_id) + "}"; } // (sic: no closing quote)
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // after a thread-level operator that made the BMTs of a parent equally long (column blocking padded to the block size, or row
        // blocking padded to the parent's longest row), once, before anything is implemented
        if (!h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty()) return false;
        bool ok = false;
        for (const auto &o : h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id)) {
            if (o->get_name().find("interlance") != std::string::npos) return false;
            if (o->get_name().find("fixed_interval_col_direction_thread_blocking_operator") != std::string::npos && o->get_is_padding_with_col_size_in_bmt()) ok = true;
            if (o->get_name().find("fixed_interval_row_direction_thread_blocking_operator") != std::string::npos && o->get_is_col_padding_with_row_max_size_with_empty_row()) ok = true;
        }
        return ok;
    }
    bool is_valid_according_to_metadata() override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        if (ms.is_exist(GLOBAL_META, "nz_row_indices_after_interlance_storage", id) || ms.is_exist(GLOBAL_META, "nz_col_indices_after_interlance_storage", id) ||
            ms.is_exist(GLOBAL_META, "nz_vals_after_interlance_storage", id))
            return false;
        if (pos_ != GLOBAL_META && !ms.is_exist(pos_, "first_BMT_indices", id)) return false;
        return ms.is_exist(pos_, "BMT_size_of_each_blk", id);
    }
    void run(bool) override
    {
        GS_REQUIRE(is_valid_according_to_metadata(), "interlance_storage_operator: needs BMT_size_of_each_blk (equally long BMTs) at the parent level and no interleaved arrays yet");
        steps::recs r;
        steps::modify_by_interlance_storage(*meta_data_set_ptr, pos_, target_matrix_id, &r);
        append(transform_seq, r);
        is_run = true;
    }

  private:
    POS_TYPE pos_ = GLOBAL_META;
};

// fixed_interval_row_matrix_div_operator (operator.hpp:289) and row_nz_matrix_div_operator (:319): CONVERTING operators that cut the
// sub-matrix into new sub-matrices (ids max + 1 ...); the parent's streams are removed, its boundaries stay
class matrix_div_operator : public basic_operator {
  public:
    matrix_div_operator(std::shared_ptr<code_generator> cg, int fixed_row_interval_size)
        : basic_operator("fixed_interval_row_matrix_div_operator", std::move(cg), CONVERTING_OP), fixed_(true), a_(fixed_row_interval_size)
    {
        GS_REQUIRE(fixed_row_interval_size > 0, "fixed_row_interval_size must be > 0");
        params = ",fixed_row_interval_size:" + std::to_string(fixed_row_interval_size);
    }
    matrix_div_operator(std::shared_ptr<code_generator> cg, int init_upper, int max_upper, int expansion)
        : basic_operator("row_nz_matrix_div_operator", std::move(cg), CONVERTING_OP), fixed_(false), a_(init_upper), b_(max_upper), c_(expansion)
    {
        GS_REQUIRE(init_upper > 0 && expansion > 0, "init_row_size_upper_boundary and expansion_rate must be > 0");
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // (sic) only a division of the same sub-matrix, met while nothing is distributed or implemented, makes the operator invalid
        if (!h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty() || !h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id).empty()) return true;
        for (const auto &o : h.read_operator_context_arr(CONVERTING_OP, target_matrix_id))
            if (o->get_name().find("div_operator") != std::string::npos) return false;
        return true;
    }
    void run(bool) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        GS_REQUIRE(ms.is_exist(GLOBAL_META, "nz_row_indices", id) && ms.is_exist(GLOBAL_META, "nz_col_indices", id) && ms.is_exist(GLOBAL_META, "nz_vals", id) &&
                       ms.is_exist(GLOBAL_META, "begin_row_index", id) && ms.is_exist(GLOBAL_META, "end_row_index", id) && ms.is_exist(GLOBAL_META, "begin_col_index", id) &&
                       ms.is_exist(GLOBAL_META, "end_col_index", id) && !ms.is_exist(GLOBAL_META, "nz_row_indices_after_interlance_storage", id),
                   name + ": the sub-matrix has no streams or boundaries to divide");
        const uint64_t row_num = ms.scalar(GLOBAL_META, "end_row_index", id) - ms.scalar(GLOBAL_META, "begin_row_index", id) + 1;
        const uint64_t max_div = (uint64_t)get_config().as_integer("MAX_DIV_TIMES_OF_DIV");
        steps::recs r;
        if (fixed_) {
            const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
            std::vector<bool> used((row_num + a_ - 1) / a_, false);
            uint64_t n = 0;
            for (uint64_t x : R.u) {
                GS_REQUIRE(x / a_ < used.size(), name + ": row index beyond end_row_index");
                if (!used[x / a_]) used[x / a_] = true, n++;
            }
            GS_REQUIRE(row_num > a_ && n <= max_div, name + ": needs more rows than one interval and at most MAX_DIV_TIMES_OF_DIV non-empty parts");
            steps::fixed_div_in_row_direction(ms, id, a_, &r);
        } else {
            const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
            std::vector<uint64_t> len = get_nnz_of_each_row_in_spec_range(R, 0, row_num - 1, 0, R.get_len() - 1);
            GS_REQUIRE(steps::div_positions_by_row_nnz(len, a_, b_, c_).size() <= max_div, name + ": more parts than MAX_DIV_TIMES_OF_DIV");
            steps::div_according_to_row_nz(ms, id, a_, b_, c_, &r);
        }
        append(transform_seq, r);
        is_run = true;
    }

  private:
    bool fixed_;
    uint64_t a_, b_ = 0, c_ = 0;
};

// merge_path_{tblock,warp,thread}_operator (operator.hpp:1205,1251,1297; operator/merge_path_*_operator.cc)
class merge_path_operator : public basic_operator {
  public:
    merge_path_operator(std::shared_ptr<code_generator> cg, POS_TYPE pos, int work_size)
        : basic_operator(std::string("merge_path_") + (pos == TBLOCK_META ? "tblock" : pos == WARP_META ? "warp" : "thread") + "_operator", std::move(cg), DISTRIBUTING_OP),
          pos_(pos), work_(work_size)
    {
        GS_REQUIRE(work_size > 0, "work_size must be > 0");
        params = ",work_size:" + std::to_string(work_size);
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // merge_path_thread_operator.cc:30-62: nothing implementing yet, and no thread-level, column-direction or
        // interlance operator before it (the tblock and warp flavours exclude their own level the same way)
        if (!h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty()) return false;
        const auto d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        if (any_name_has(d, "thread") || any_name_has(d, "col") || any_name_has(d, "interlance")) return false;
        if (pos_ != THREAD_META && any_name_has(d, "warp")) return false;
        if (pos_ == TBLOCK_META && any_name_has(d, "tblock")) return false;
        return true;
    }
    void run(bool) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        GS_REQUIRE(ms.count_of_metadata_of_diff_pos(THREAD_META, id) == 0, "merge path: thread-level metadata already exists");
        if (pos_ != THREAD_META) GS_REQUIRE(ms.count_of_metadata_of_diff_pos(WARP_META, id) == 0, "merge path: warp-level metadata already exists");
        if (pos_ == TBLOCK_META) GS_REQUIRE(ms.count_of_metadata_of_diff_pos(TBLOCK_META, id) == 0, "merge path: tblock-level metadata already exists");
        steps::recs r;
        steps::get_begin_rows_of_level_after_merge_path(ms, pos_, id, work_, &r);
        steps::get_begin_nzs_of_level_after_merge_path(ms, pos_, id, work_, &r);
        append(transform_seq, r);
        code_generator_ptr->open_spec_level_of_paral(TBLOCK_META); // all three flavours open the tblock level (merge_path_*_operator.cc)
        is_run = true;
    }

  private:
    POS_TYPE pos_;
    uint64_t work_;
};

class empty_row_pad_operator : public basic_operator { // operator.hpp:449
  public:
    explicit empty_row_pad_operator(std::shared_ptr<code_generator> cg) : basic_operator("empty_row_pad_operator", std::move(cg), CONVERTING_OP) {}
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        return h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id).empty() && h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty();
    }
    void run(bool) override
    {
        steps::recs r;
        steps::modify_by_empty_pad_in_submatrix(*meta_data_set_ptr, target_matrix_id, &r);
        append(transform_seq, r);
        is_run = true;
    }
};

// ----------------------------------------------------------------------------------------- IMPLEMENTING
class thread_total_reduce_operator : public basic_operator { // operator.hpp:1070
  public:
    thread_total_reduce_operator(std::shared_ptr<code_generator> cg, bool need_warp_reduction, int S, int C)
        : basic_operator("thread_total_reduce_operator", std::move(cg), IMPLEMENTING_OP), nwr_(need_warp_reduction), S_(S), C_(C)
    {
        GS_REQUIRE(S >= 1 && C >= 1, "coarsen factors must be >= 1");
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // thread_total_reduce_operator.cc:14-40: a thread-level distribution, no nnz-direction operator before it (an nnz chunk is not a
        // row: the token stores its sum), nothing implemented yet
        const opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        return h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty() && !any_name_has(d, "nnz") && any_name_has(d, "thread");
    }
    bool is_valid_according_to_metadata() override { return meta_data_set_ptr->is_exist(THREAD_META, "first_nz_indices", target_matrix_id); } // :41-55
    void run(bool) override
    {
        GS_REQUIRE(is_valid_according_to_metadata(), "thread_total_reduce_operator: no THREAD first_nz_indices");
        reduction_token t;
        t.kind = RED_TOTAL_BMT, t.need_warp_reduction = nwr_, t.sparse_coarsen_factor = S_, t.coarsen_factor = C_;
        code_generator_ptr->set_reduction_token(THREAD_META, t);
        code_generator_ptr->open_spec_level_of_paral(THREAD_META);
        is_run = true;
    }

  private:
    bool nwr_;
    int S_, C_;
};

class thread_bit_map_operator : public basic_operator { // operator.hpp:1093, operator/thread_bit_map_operator.cc
  public:
    thread_bit_map_operator(std::shared_ptr<code_generator> cg, POS_TYPE pos, int size, int S, int C)
        : basic_operator("thread_bit_map_operator", std::move(cg), IMPLEMENTING_OP), pos_(pos), size_(size), S_(S), C_(C)
    {
        GS_REQUIRE(size >= 1, "size must be >= 1");
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // thread_bit_map_operator.cc: nothing implemented yet, a thread-level AND an nnz-direction distribution
        const opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        return h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty() && any_name_has(d, "thread") && any_name_has(d, "nnz");
    }
    bool is_valid_according_to_metadata() override { return meta_data_set_ptr->is_exist(THREAD_META, "first_nz_indices", target_matrix_id); }
    void run(bool) override
    {
        GS_REQUIRE(is_valid_according_to_metadata(), "thread_bit_map_operator: no THREAD first_nz_indices");
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        steps::recs r;
        steps::thread_bit_map(ms, pos_ != THREAD_META, size_, id, &r);
        steps::segment_empty_flag(ms, pos_, size_, id, &r);
        steps::segment_empty_row_indices(ms, pos_, id, &r);
        steps::segment_offset(ms, false, size_, id, &r);
        steps::segment_ptr(ms, pos_, id, &r);
        append(transform_seq, r);
        reduction_token t;
        t.kind = RED_THREAD_BIT_MAP, t.need_warp_reduction = (pos_ == WARP_META), t.sparse_coarsen_factor = S_, t.coarsen_factor = C_;
        code_generator_ptr->set_reduction_token(THREAD_META, t);
        code_generator_ptr->open_spec_level_of_paral(THREAD_META);
        is_run = true;
    }

  private:
    POS_TYPE pos_;
    uint64_t size_;
    int S_, C_;
};

class warp_total_reduce_operator : public basic_operator { // operator.hpp:1117
  public:
    warp_total_reduce_operator(std::shared_ptr<code_generator> cg, int C) : basic_operator("warp_total_reduce_operator", std::move(cg), IMPLEMENTING_OP), C_(C) {}
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // warp_total_reduce_operator.cc: nothing implemented yet, a warp-level distribution, no nnz-direction operator
        const opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        return h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty() && !any_name_has(d, "nnz") && any_name_has(d, "warp");
    }
    bool is_valid_according_to_metadata() override { return meta_data_set_ptr->is_exist(WARP_META, "first_nz_indices", target_matrix_id); }
    void run(bool) override
    {
        GS_REQUIRE(is_valid_according_to_metadata(), "warp_total_reduce_operator: no WARP first_nz_indices");
        reduction_token t;
        t.kind = RED_TOTAL_WARP, t.coarsen_factor = C_;
        code_generator_ptr->set_reduction_token(WARP_META, t);
        code_generator_ptr->open_spec_level_of_paral(WARP_META);
        is_run = true;
    }

  private:
    int C_;
};

// warp_bit_map_operator (operator.hpp:1139) and warp_segment_reduce_operator (:1162): group VECTOR_WIDTH BMTs into a BMW
class warp_group_operator : public basic_operator {
  public:
    warp_group_operator(std::shared_ptr<code_generator> cg, bool bitmap, int C, bool relative_nz, bool relative_row)
        : basic_operator(bitmap ? "warp_bit_map_operator" : "warp_segment_reduce_operator", std::move(cg), IMPLEMENTING_OP), bitmap_(bitmap), C_(C),
          rel_nz_(relative_nz), rel_row_(relative_row)
    {
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // warp_bit_map_operator.cc: a thread-level distribution, no warp-level one, no nnz-direction operator;
        // warp_segment_reduce_operator.cc: thread-level AND nnz-direction distribution, no warp-level one, thread_bit_map implemented.
        // Both consume what the thread-level reduction before them leaves, so that one must have run (the reference fails later, in the token)
        const opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id), im = h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id);
        if (!any_name_has(d, "thread") || any_name_has(d, "warp")) return false;
        if (bitmap_) return !any_name_has(d, "nnz") && any_name_has(im, "thread_total_reduce");
        return any_name_has(d, "nnz") && any_name_has(im, "thread_bit_map");
    }
    bool is_valid_according_to_metadata() override
    {
        return meta_data_set_ptr->is_exist(THREAD_META, "first_nz_indices", target_matrix_id) && !meta_data_set_ptr->is_exist(WARP_META, "first_nz_indices", target_matrix_id);
    }
    void run(bool) override
    {
        GS_REQUIRE(is_valid_according_to_metadata(), name + ": needs THREAD first_nz_indices and no WARP first_nz_indices");
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        uint64_t vw = (uint64_t)get_config().as_float("VECTOR_WIDTH");
        steps::recs r;
        steps::get_begin_rows_after_merge_thread(ms, WARP_META, vw, id, &r);
        steps::get_begin_nzs_after_merge_thread(ms, WARP_META, vw, id, &r);
        if (rel_row_) steps::get_begin_rows_relative_to_parent_after_merge_thread(ms, WARP_META, vw, id, &r);
        if (rel_nz_) steps::get_begin_nzs_relative_to_parent_after_merge_thread(ms, WARP_META, vw, id, &r);
        steps::get_begin_BMTs_after_merge_thread(ms, WARP_META, vw, id, &r);
        if (bitmap_) steps::parent_bit_map_of_thread(ms, WARP_META, id, vw, &r);
        append(transform_seq, r);
        reduction_token t;
        t.kind = bitmap_ ? RED_WARP_BIT_MAP : RED_WARP_SEGMENT, t.coarsen_factor = C_;
        code_generator_ptr->set_reduction_token(WARP_META, t);
        code_generator_ptr->open_spec_level_of_paral(WARP_META);
        is_run = true;
    }

  private:
    bool bitmap_;
    int C_;
    bool rel_nz_, rel_row_;
};

class tblock_total_reduce_operator : public basic_operator { // operator.hpp:1186
  public:
    tblock_total_reduce_operator(std::shared_ptr<code_generator> cg, int C) : basic_operator("tblock_total_reduce_operator", std::move(cg), IMPLEMENTING_OP), C_(C) {}
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // tblock_total_reduce_operator.cc: nothing implemented yet, a tblock-level distribution, no nnz-direction operator
        const opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        return h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id).empty() && !any_name_has(d, "nnz") && any_name_has(d, "tblock");
    }
    bool is_valid_according_to_metadata() override { return meta_data_set_ptr->is_exist(TBLOCK_META, "first_nz_indices", target_matrix_id); }
    void run(bool) override
    {
        GS_REQUIRE(is_valid_according_to_metadata(), "tblock_total_reduce_operator: no TBLOCK first_nz_indices");
        reduction_token t;
        t.kind = RED_TOTAL_TBLOCK, t.coarsen_factor = C_;
        code_generator_ptr->set_reduction_token(TBLOCK_META, t);
        code_generator_ptr->open_spec_level_of_paral(TBLOCK_META);
        is_run = true;
    }

  private:
    int C_;
};

class tblock_thread_bit_map_operator : public basic_operator { // operator.hpp:1209
  public:
    tblock_thread_bit_map_operator(std::shared_ptr<code_generator> cg, int C, int block_size, bool relative_nz, bool relative_row)
        : basic_operator("tblock_thread_bit_map_operator", std::move(cg), IMPLEMENTING_OP), C_(C), bs_(block_size), rel_nz_(relative_nz), rel_row_(relative_row)
    {
        GS_REQUIRE(block_size >= 1, "block_size must be >= 1");
    }
    bool is_valid_according_to_operator(const operator_context &h) override
    {
        // tblock_thread_bit_map_operator.cc: a thread-level distribution, no tblock-level one, no nnz-direction operator; it consumes
        // the BMT sums of thread_total_reduce
        const opv d = h.read_operator_context_arr(DISTRIBUTING_OP, target_matrix_id);
        return any_name_has(d, "thread") && !any_name_has(d, "tblock") && !any_name_has(d, "nnz") &&
               any_name_has(h.read_operator_context_arr(IMPLEMENTING_OP, target_matrix_id), "thread_total_reduce");
    }
    void run(bool) override
    {
        auto &ms = *meta_data_set_ptr;
        int id = target_matrix_id;
        steps::recs r;
        steps::get_begin_rows_after_merge_thread(ms, TBLOCK_META, bs_, id, &r);
        steps::get_begin_nzs_after_merge_thread(ms, TBLOCK_META, bs_, id, &r);
        if (rel_row_) steps::get_begin_rows_relative_to_parent_after_merge_thread(ms, TBLOCK_META, bs_, id, &r);
        if (rel_nz_) steps::get_begin_nzs_relative_to_parent_after_merge_thread(ms, TBLOCK_META, bs_, id, &r);
        steps::get_begin_BMTs_after_merge_thread(ms, TBLOCK_META, bs_, id, &r);
        steps::parent_bit_map_of_thread(ms, TBLOCK_META, id, (uint64_t)get_config().as_float("VECTOR_WIDTH"), &r);
        steps::segment_offset(ms, true, bs_, id, &r);
        append(transform_seq, r);
        reduction_token t;
        t.kind = RED_TBLOCK_BIT_MAP, t.coarsen_factor = C_;
        code_generator_ptr->set_reduction_token(TBLOCK_META, t);
        code_generator_ptr->open_spec_level_of_paral(TBLOCK_META);
        is_run = true;
    }

  private:
    int C_;
    uint64_t bs_;
    bool rel_nz_, rel_row_;
};

class grid_block_operator : public basic_operator { // operator.hpp:1047, operator/grid_block_operator.cc:2-34
  public:
    grid_block_operator(std::shared_ptr<code_generator> cg, unsigned grid_x, std::vector<unsigned> block, unsigned C)
        : basic_operator("grid_block_operator", std::move(cg), IMPLEMENTING_OP)
    {
        GS_REQUIRE(block.size() == 2 && block[0] > 0 && block[1] > 0 && C > 0, "block must be (x,y) > 0");
        if (block[1] % 2 != 0) block[1]++;
        long n = get_config().as_integer("DENSE_MATRIX_SIZE");
        unsigned per = (meta_data_set_ptr->is_exist(WARP_META, "first_nz_indices", target_matrix_id) ? block[1] : block[0]) * C;
        unsigned grid_y = (unsigned)((n + per - 1) / per);
        grid_ = {grid_x, grid_y};
        block_ = block;
    }
    void run(bool) override
    {
        code_generator_ptr->set_thread_grid(grid_, block_);
        is_run = true;
    }

  private:
    std::vector<unsigned> grid_, block_;
};

} // namespace

// ------------------------------------------------------------------------------------------------ factory
const std::vector<std::string> &operator_names()
{
    static const std::vector<std::string> n = {"sort_operator",
                                               "fixed_interval_row_direction_tblock_blocking_operator",
                                               "fixed_interval_row_direction_warp_blocking_operator",
                                               "fixed_interval_row_direction_thread_blocking_operator",
                                               "fixed_interval_col_direction_thread_blocking_operator",
                                               "fixed_interval_nnz_direction_thread_blocking_operator",
                                               "fixed_interval_nnz_direction_warp_blocking_operator",
                                               "fixed_interval_nnz_direction_tblock_blocking_operator",
                                               "balanced_interval_row_direction_tblock_blocking_operator",
                                               "balanced_interval_row_direction_warp_blocking_operator",
                                               "balanced_interval_row_direction_thread_blocking_operator",
                                               "empty_row_pad_operator",
                                               "fixed_interval_col_direction_tblock_blocking_operator",
                                               "fixed_interval_col_direction_warp_blocking_operator",
                                               "interlance_storage_operator",
                                               "fixed_interval_row_matrix_div_operator",
                                               "row_nz_matrix_div_operator",
                                               "merge_path_tblock_operator",
                                               "merge_path_warp_operator",
                                               "merge_path_thread_operator",
                                               "thread_total_reduce_operator",
                                               "thread_bit_map_operator",
                                               "warp_total_reduce_operator",
                                               "warp_bit_map_operator",
                                               "warp_segment_reduce_operator",
                                               "tblock_total_reduce_operator",
                                               "tblock_thread_bit_map_operator",
                                               "grid_block_operator"};
    return n;
}

std::shared_ptr<basic_operator> make_operator(const std::string &op, std::shared_ptr<code_generator> cg, std::shared_ptr<operator_context> h,
                                              const std::vector<int64_t> &a)
{
    auto need = [&](size_t n) { GS_REQUIRE(a.size() == n, op + " takes " + std::to_string(n) + " arguments after code_generator_ptr"); };
    if (op == "sort_operator") return need(0), std::make_shared<sort_operator>(cg);
    if (op == "fixed_interval_row_direction_tblock_blocking_operator")
        return need(2), std::make_shared<fixed_interval_row_direction_tblock_blocking_operator>(cg, (int)a[0], a[1] != 0);
    if (op == "fixed_interval_row_direction_warp_blocking_operator")
        return need(4), std::make_shared<fixed_interval_row_direction_warp_blocking_operator>(cg, (int)a[0], a[1] != 0, a[2] != 0, a[3] != 0);
    if (op == "fixed_interval_row_direction_thread_blocking_operator")
        return need(7), std::make_shared<fixed_interval_row_direction_thread_blocking_operator>(cg, (int)a[0], a[1] != 0, a[2] != 0, a[3] != 0, a[4] != 0,
                                                                                                 a[5] != 0, (int)a[6], h);
    if (op == "fixed_interval_col_direction_thread_blocking_operator")
        return need(5), std::make_shared<fixed_interval_col_direction_thread_blocking_operator>(cg, (int)a[0], a[1] != 0, a[2] != 0, a[3] != 0, a[4] != 0, h);
    if (op == "fixed_interval_nnz_direction_thread_blocking_operator")
        return need(4), std::make_shared<fixed_interval_nnz_direction_thread_blocking_operator>(cg, (int)a[0], a[1] != 0, a[2] != 0, a[3] != 0);
    if (op == "fixed_interval_nnz_direction_warp_blocking_operator")
        return need(4), std::make_shared<fixed_interval_nnz_direction_warp_blocking_operator>(cg, (int)a[0], a[1] != 0, a[2] != 0, a[3] != 0);
    if (op == "fixed_interval_nnz_direction_tblock_blocking_operator")
        return need(2), std::make_shared<fixed_interval_nnz_direction_tblock_blocking_operator>(cg, (int)a[0], a[1] != 0);
    if (op == "balanced_interval_row_direction_tblock_blocking_operator")
        return need(1), std::make_shared<balanced_interval_row_direction_blocking_operator>(cg, TBLOCK_META, (int)a[0], false, false);
    if (op == "fixed_interval_col_direction_tblock_blocking_operator")
        return need(3), std::make_shared<fixed_interval_col_direction_tblock_blocking_operator>(cg, (int)a[0], a[1] != 0, a[2] != 0);
    if (op == "fixed_interval_col_direction_warp_blocking_operator")
        return need(5), std::make_shared<fixed_interval_col_direction_warp_blocking_operator>(cg, (int)a[0], a[1] != 0, a[2] != 0, a[3] != 0, a[4] != 0, h);
    if (op == "interlance_storage_operator") return need(0), std::make_shared<interlance_storage_operator>(cg, h);
    if (op == "fixed_interval_row_matrix_div_operator") return need(1), std::make_shared<matrix_div_operator>(cg, (int)a[0]);
    if (op == "row_nz_matrix_div_operator") return need(3), std::make_shared<matrix_div_operator>(cg, (int)a[0], (int)a[1], (int)a[2]);
    // the warp and thread flavours take (nnz_per_interval, row_index_is_relative, nz_index_is_relative) like the reference;
    // the one-argument form stays for callers that block at the top level only
    if (op == "balanced_interval_row_direction_warp_blocking_operator" || op == "balanced_interval_row_direction_thread_blocking_operator") {
        GS_REQUIRE(a.size() == 1 || a.size() == 3, op + " takes 1 or 3 arguments after code_generator_ptr");
        return std::make_shared<balanced_interval_row_direction_blocking_operator>(cg, op.find("warp") != std::string::npos ? WARP_META : THREAD_META, (int)a[0], a.size() == 3 && a[1] != 0,
                                                                                   a.size() == 3 && a[2] != 0);
    }
    if (op == "merge_path_tblock_operator") return need(1), std::make_shared<merge_path_operator>(cg, TBLOCK_META, (int)a[0]);
    if (op == "merge_path_warp_operator") return need(1), std::make_shared<merge_path_operator>(cg, WARP_META, (int)a[0]);
    if (op == "merge_path_thread_operator") return need(1), std::make_shared<merge_path_operator>(cg, THREAD_META, (int)a[0]);
    if (op == "empty_row_pad_operator") return need(0), std::make_shared<empty_row_pad_operator>(cg);
    if (op == "thread_total_reduce_operator") return need(3), std::make_shared<thread_total_reduce_operator>(cg, a[0] != 0, (int)a[1], (int)a[2]);
    if (op == "thread_bit_map_operator") return need(4), std::make_shared<thread_bit_map_operator>(cg, (POS_TYPE)a[0], (int)a[1], (int)a[2], (int)a[3]);
    if (op == "warp_total_reduce_operator") return need(1), std::make_shared<warp_total_reduce_operator>(cg, (int)a[0]);
    if (op == "warp_bit_map_operator") return need(3), std::make_shared<warp_group_operator>(cg, true, (int)a[0], a[1] != 0, a[2] != 0);
    if (op == "warp_segment_reduce_operator") return need(3), std::make_shared<warp_group_operator>(cg, false, (int)a[0], a[1] != 0, a[2] != 0);
    if (op == "tblock_total_reduce_operator") return need(1), std::make_shared<tblock_total_reduce_operator>(cg, (int)a[0]);
    if (op == "tblock_thread_bit_map_operator")
        return need(4), std::make_shared<tblock_thread_bit_map_operator>(cg, (int)a[0], (int)a[1], a[2] != 0, a[3] != 0);
    if (op == "grid_block_operator")
        return need(4), std::make_shared<grid_block_operator>(cg, (unsigned)a[0], std::vector<unsigned>{(unsigned)a[1], (unsigned)a[2]}, (unsigned)a[3]);
    throw Error("unknown or unbuilt operator: " + op);
}

// ------------------------------------------------------------------------------- the driver's seven chains
void run_family_chain(int family, std::shared_ptr<code_generator> cg, operator_executer &ex, int N, int S, int C)
{
    GS_REQUIRE(N >= 1 && S >= 1 && C >= 1, "N, sparse_coarsen and coarsen must be >= 1");
    auto h = ex.get_operator_context();
    auto ms = cg->get_metadata_set();
    auto run = [&](const char *name, std::vector<int64_t> args) { ex.add_and_run(make_operator(name, cg, h, args)); };
    set_config("DENSE_MATRIX_SIZE", (double)N); // token_test.cc:1630
    int64_t rows = (int64_t)ms->scalar(GLOBAL_META, "origin_row_num", -1), nnz = (int64_t)ms->scalar(GLOBAL_META, "origin_nnz_num", -1);
    int x, y;
    switch (family) {
    case GS_FAMILY_THREAD_TOTAL: // token_test.cc:1003-1092
        run("sort_operator", {});
        run("fixed_interval_row_direction_thread_blocking_operator", {1, 0, 0, 0, 0, 1, S});
        x = std::min(N / C, 32);
        GS_REQUIRE(x >= 1, "N / coarsen must be >= 1");
        set_config("VECTOR_WIDTH", (double)x);
        y = 128 / x;
        GS_REQUIRE(ms->check(), "metadata check failed");
        run("thread_total_reduce_operator", {0, S, C});
        run("grid_block_operator", {rows / y + 1, x, y, C});
        break;
    case GS_FAMILY_WARP_TOTAL: // :1188-1249
        run("fixed_interval_row_direction_warp_blocking_operator", {1, 0, 0, 0});
        y = std::min(N / C, 32);
        GS_REQUIRE(y >= 1, "N / coarsen must be >= 1");
        x = std::max(256 / y, 32);
        set_config("VECTOR_WIDTH", (double)x);
        run("warp_total_reduce_operator", {C});
        run("grid_block_operator", {rows, x, y, C});
        break;
    case GS_FAMILY_WARP_BITMAP: // :1253-1315
        run("fixed_interval_col_direction_thread_blocking_operator", {64, 0, 0, 1, 0});
        y = std::min(N / C, 32);
        GS_REQUIRE(y >= 1, "N / coarsen must be >= 1");
        x = std::max(128 / y, 32);
        set_config("VECTOR_WIDTH", (double)x);
        run("thread_total_reduce_operator", {1, S, C});
        run("warp_bit_map_operator", {C, 1, 1});
        run("grid_block_operator", {rows, x, y, C});
        break;
    case GS_FAMILY_THREAD_BITMAP: // :1319-1391
        run("fixed_interval_nnz_direction_thread_blocking_operator", {32, 0, 0, 1});
        x = std::min(N / C, 32);
        GS_REQUIRE(x >= 1, "N / coarsen must be >= 1");
        set_config("VECTOR_WIDTH", (double)x);
        y = 128 / x;
        cg->open_spec_level_of_paral(THREAD_META);
        run("thread_bit_map_operator", {THREAD_META, x, S, C});
        run("grid_block_operator", {nnz / y + 1, x, y, C});
        break;
    case GS_FAMILY_WARP_SEGMENT: // :1393-1455
        run("fixed_interval_nnz_direction_thread_blocking_operator", {32, 0, 0, 1});
        x = std::min(N, 32);
        y = 256 / x;
        set_config("VECTOR_WIDTH", (double)x);
        run("thread_bit_map_operator", {WARP_META, x, S, C});
        run("warp_segment_reduce_operator", {C, 0, 0});
        run("grid_block_operator", {nnz / 128 + 1, x, y, C});
        break;
    case GS_FAMILY_TBLOCK_TOTAL: // :1458-1514
        run("fixed_interval_row_direction_tblock_blocking_operator", {1, 0});
        x = std::min(N, 32);
        y = 256 / x;
        set_config("VECTOR_WIDTH", (double)x);
        run("tblock_total_reduce_operator", {C});
        run("grid_block_operator", {rows, x, y, C});
        break;
    case GS_FAMILY_TBLOCK_BITMAP: { // :1518-1582
        run("fixed_interval_col_direction_thread_blocking_operator", {64, 0, 0, 1, 0});
        x = std::min(N / C, 32);
        GS_REQUIRE(x >= 1, "N / coarsen must be >= 1");
        y = 256 / x;
        set_config("VECTOR_WIDTH", (double)x);
        run("thread_total_reduce_operator", {0, S, C});
        run("grid_block_operator", {rows, x, y, C});
        int y_even = y % 2 ? y + 1 : y; // the driver passes block[1] before grid_block_operator's rounding of its own copy
        (void)y_even;
        run("tblock_thread_bit_map_operator", {C, y, 0, 0});
        break;
    }
    default: throw Error("run_family_chain: not one of the seven driver families");
    }
}

} // namespace gs
