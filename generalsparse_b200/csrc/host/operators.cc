#include "operators.hpp"

#include "../../../include/gsparse_b200.h"

#include <algorithm>
#include <cerrno>
#include <cstdlib>
#include <ctime>
#include <fstream>
#include <sys/stat.h>

namespace gs {

const char *stage_name(operator_stage_type s)
{
    switch (s) {
    case CONVERTING_OP: return "CONVERTING_OP";
    case DISTRIBUTING_OP: return "DISTRIBUTING_OP";
    case IMPLEMENTING_OP: return "IMPLEMENTING_OP";
    default: return "NONE_OP";
    }
}

// ------------------------------------------------------------------------------------ code_generator
code_generator::code_generator(std::shared_ptr<meta_data_set> ms, int sub_matrix_id) : ms_(std::move(ms)), id_(sub_matrix_id)
{
    GS_REQUIRE(ms_ != nullptr, "null metadata set");
    GS_REQUIRE(sub_matrix_id >= 0, "sub_matrix_id must be >= 0");
    GS_REQUIRE(ms_->is_exist(GLOBAL_META, "nz_row_indices", id_), "sub-matrix has no nz_row_indices");
}

void code_generator::open_spec_level_of_paral(POS_TYPE pos)
{
    GS_REQUIRE(pos == TBLOCK_META || pos == WARP_META || pos == THREAD_META, "level must be TBLOCK/WARP/THREAD");
    open_[pos] = true;
}

void code_generator::set_reduction_token(POS_TYPE pos, const reduction_token &tok)
{
    GS_REQUIRE(pos == TBLOCK_META || pos == WARP_META || pos == THREAD_META, "level must be TBLOCK/WARP/THREAD");
    tokens_[pos] = tok;
    compiled_ = false;
}

void code_generator::set_thread_grid(const std::vector<unsigned> &grid, const std::vector<unsigned> &block)
{
    GS_REQUIRE(grid.size() == 2 && block.size() == 2, "grid and block are 2-vectors");
    grid_ = grid;
    block_ = block;
}

void code_generator::add_new_metadata_dependency(POS_TYPE pos, const std::string &name, int id)
{
    GS_REQUIRE(ms_->is_exist(pos, name, id), "dependency on a missing item " + meta_data_set::item_name(pos, name, id));
    for (const auto &d : extra_deps_)
        if (d.pos == pos && d.name == name && d.sub_matrix_id == id) return;
    extra_deps_.push_back({pos, name, id});
}

const kernel_plan &code_generator::plan() const
{
    GS_REQUIRE(compiled_, "compile() has not run");
    return plan_;
}

// compile(): the installed reduction tokens fix the family; the arrays each family reads are the
// ones the reference's tokens register through add_new_metadata_dependency (the files its
// generate_final_program writes to data_source/<id>/ -- pinned by tests/golden "registered" lists).
void code_generator::compile()
{
    const reduction_token &t = tokens_[THREAD_META], &w = tokens_[WARP_META], &b = tokens_[TBLOCK_META];
    kernel_plan p;
    auto dep = [&](POS_TYPE pos, const char *name) {
        GS_REQUIRE(ms_->is_exist(pos, name, id_), std::string("the chain did not produce ") + meta_data_set::item_name(pos, name, id_));
        p.deps.push_back({pos, name, id_});
    };
    dep(GLOBAL_META, "nz_col_indices");
    dep(GLOBAL_META, "nz_vals");
    p.has_row_permutation = ms_->is_exist(GLOBAL_META, "original_nz_row_indices", id_);
    if (p.has_row_permutation) dep(GLOBAL_META, "original_nz_row_indices");

    // a total reduction over blocks that are not one row each -- merge-path parts (first_row_indices_without_ending only), nnz-balanced row
    // groups, row blocks of several rows -- has no reference-shaped kernel (those store a block's sum as one row): the range kernel
    // takes the level's first_nz_indices as ranges and needs the row of every non-zero
    auto one_row_per_block = [&](POS_TYPE pos) {
        if (!ms_->is_exist(pos, "first_row_indices", id_)) return false;
        const std::vector<uint64_t> &fr = ms_->get(pos, "first_row_indices", id_).u;
        for (uint64_t i = 0; i < fr.size(); i++)
            if (fr[i] != i) return false;
        return true;
    };
    const POS_TYPE total_level = (t.kind == RED_TOTAL_BMT && w.kind == RED_NONE && b.kind == RED_NONE) ? THREAD_META
                                 : (w.kind == RED_TOTAL_WARP && t.kind == RED_NONE)                     ? WARP_META
                                 : (b.kind == RED_TOTAL_TBLOCK && w.kind == RED_NONE && t.kind == RED_NONE) ? TBLOCK_META
                                                                                                         : NONE_META;
    if (total_level != NONE_META && ms_->is_exist(total_level, "first_nz_indices", id_) && !one_row_per_block(total_level)) {
        p.family = GS_FAMILY_B200_MERGE;
        p.coarsen = total_level == THREAD_META ? t.coarsen_factor : total_level == WARP_META ? w.coarsen_factor : b.coarsen_factor;
        dep(GLOBAL_META, "nz_row_indices");
        dep(total_level, "first_nz_indices");
        dep(total_level, ms_->is_exist(total_level, "first_row_indices", id_) ? "first_row_indices" : "first_row_indices_without_ending");
    } else if (b.kind == RED_TOTAL_TBLOCK && w.kind == RED_NONE && t.kind == RED_NONE) {
        p.family = GS_FAMILY_TBLOCK_TOTAL;
        p.coarsen = b.coarsen_factor;
        dep(TBLOCK_META, "first_row_indices");
        dep(TBLOCK_META, "first_nz_indices");
    } else if (b.kind == RED_TBLOCK_BIT_MAP && t.kind == RED_TOTAL_BMT) {
        p.family = GS_FAMILY_TBLOCK_BITMAP;
        p.coarsen = b.coarsen_factor;
        p.sparse_coarsen = t.sparse_coarsen_factor;
        dep(TBLOCK_META, "first_BMT_indices");
        dep(THREAD_META, "first_nz_indices");
        dep(THREAD_META, "first_row_indices_without_ending");
        dep(THREAD_META, "bit_map_of_thread");
        dep(THREAD_META, "segment_offset");
    } else if (w.kind == RED_TOTAL_WARP && t.kind == RED_NONE) {
        p.family = GS_FAMILY_WARP_TOTAL;
        p.coarsen = w.coarsen_factor;
        dep(WARP_META, "first_row_indices");
        dep(WARP_META, "first_nz_indices");
    } else if (w.kind == RED_WARP_BIT_MAP && t.kind == RED_TOTAL_BMT) {
        p.family = GS_FAMILY_WARP_BITMAP;
        p.coarsen = w.coarsen_factor;
        p.sparse_coarsen = t.sparse_coarsen_factor;
        p.need_warp_reduction = true;
        dep(WARP_META, "first_row_indices");
        dep(WARP_META, "first_nz_indices");
        dep(WARP_META, "first_BMT_indices");
        dep(WARP_META, "bit_map_of_thread");
        dep(THREAD_META, "first_row_indices_without_ending");
        dep(THREAD_META, ms_->is_exist(THREAD_META, "first_nz_indices_relative_to_BMW", id_) ? "first_nz_indices_relative_to_BMW" : "first_nz_indices");
    } else if (w.kind == RED_WARP_SEGMENT && t.kind == RED_THREAD_BIT_MAP) {
        p.family = GS_FAMILY_WARP_SEGMENT;
        p.coarsen = w.coarsen_factor;
        p.sparse_coarsen = t.sparse_coarsen_factor;
        p.need_warp_reduction = true;
        dep(WARP_META, "first_BMT_indices");
        dep(THREAD_META, "first_row_indices");
        dep(THREAD_META, "first_nz_indices");
        dep(THREAD_META, "thread_bit_map");
        dep(THREAD_META, "segment_ptr");
        dep(THREAD_META, "segment_offset");
        dep(THREAD_META, "segment_empty_flag");
        dep(THREAD_META, "segment_empty_row_indices");
    } else if (t.kind == RED_THREAD_BIT_MAP && w.kind == RED_NONE && b.kind == RED_NONE) {
        p.family = GS_FAMILY_THREAD_BITMAP;
        p.coarsen = t.coarsen_factor;
        p.sparse_coarsen = t.sparse_coarsen_factor;
        dep(THREAD_META, "first_row_indices");
        dep(THREAD_META, "first_nz_indices");
        dep(THREAD_META, "thread_bit_map");
        dep(THREAD_META, "segment_ptr");
        dep(THREAD_META, "segment_empty_flag");
        dep(THREAD_META, "segment_empty_row_indices");
    } else if (t.kind == RED_TOTAL_BMT && w.kind == RED_NONE && b.kind == RED_NONE) {
        p.family = GS_FAMILY_THREAD_TOTAL;
        p.coarsen = t.coarsen_factor;
        p.sparse_coarsen = t.sparse_coarsen_factor;
        dep(THREAD_META, "first_row_indices");
        dep(THREAD_META, "first_nz_indices");
    } else {
        throw Error("compile: the installed reduction tokens do not form one of the kernel families");
    }
    for (const auto &d : extra_deps_) {
        bool seen = false;
        for (const auto &e : p.deps) seen = seen || (e.pos == d.pos && e.name == d.name);
        if (!seen) p.deps.push_back(d);
    }
    GS_REQUIRE(grid_.size() == 2, "grid_block_operator has not set the launch shape");
    p.grid[0] = grid_[0], p.grid[1] = grid_[1], p.block[0] = block_[0], p.block[1] = block_[1];
    p.vector_width = (int)get_config().as_integer("VECTOR_WIDTH");
    p.thread_for_row = thread_for_row_;
    p.rows = ms_->scalar(GLOBAL_META, "origin_row_num", -1);
    p.cols = ms_->scalar(GLOBAL_META, "origin_col_num", -1);
    p.origin_nnz = ms_->scalar(GLOBAL_META, "origin_nnz_num", -1);
    p.stored_nnz = ms_->get(GLOBAL_META, "nz_col_indices", id_).get_len();
    plan_ = p;
    compiled_ = true;
}

unsigned long code_generator::generate_final_program(unsigned int repeat_num)
{
    GS_REQUIRE(repeat_num >= 1, "repeat_num must be >= 1");
    if (!compiled_) compile();
    // the directory id: rand() + time % 1000 in the reference (metadata_set.cc:529-530); any id not taken yet will do
    const std::string root = get_config().as_string("ROOT_PATH_STR") + "/data_source";
    mkdir(root.c_str(), 0755);
    unsigned long id = (unsigned long)time(nullptr) % 1000 + (unsigned long)rand();
    std::string dir;
    for (;; id++) {
        dir = root + "/" + std::to_string(id);
        if (mkdir(dir.c_str(), 0755) == 0) break;
        GS_REQUIRE(errno == EEXIST, "cannot create " + dir);
    }
    for (const metadata_dependency &d : plan_.deps) {
        const std::string key = meta_data_set::item_name(d.pos, d.name, d.sub_matrix_id);
        ms_->output_item_to_file(key, dir + "/" + key);
    }
    std::ofstream(dir + "/gs_repeat") << repeat_num << "\n"; // the reference bakes the count into the program text
    return id;
}

// ---------------------------------------------------------------------------------- context / executer
std::vector<std::shared_ptr<basic_operator>> operator_context::read_operator_context_arr(operator_stage_type stage, int id) const
{
    std::vector<std::shared_ptr<basic_operator>> out;
    for (const auto &o : all_)
        if (o->get_stage() == stage && o->get_target_matrix_id() == id) out.push_back(o);
    return out;
}

void operator_context::add(std::shared_ptr<basic_operator> op) { all_.push_back(std::move(op)); }

std::string basic_operator::convert_to_string() const
{
    return name + "::{name:\"" + name + "\",stage:" + stage_name(stage) + ",target_matrix_id:" + std::to_string(target_matrix_id) + params + "}";
}

void operator_executer::add_and_run(std::shared_ptr<basic_operator> op)
{
    GS_REQUIRE(op != nullptr, "null operator");
    // operator_executer.cc:19-26: the order whitelist is asserted, then run, then recorded
    GS_REQUIRE(op->is_valid_according_to_operator(*ctx_), op->get_name() + " is not valid after the operators already run");
    op->run(get_config().as_bool("OPERATOR_RUNTIME_CHECK"));
    ctx_->add(op);
}

} // namespace gs
