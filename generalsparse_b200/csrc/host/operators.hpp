// Operator graph of the SpMM path: B200-native counterpart of operator.hpp / operator/*.cc /
// operator_executer.{hpp,cc} / the operator-facing part of code_generator.hpp.
// Operator class names, constructor argument order (the reference's third flavour, the one that
// takes a code_generator), stages, step order and records follow the reference; what they drive
// is a kernel plan (family + launch shape + list of metadata arrays) instead of CUDA source text.
#pragma once
#include "config.hpp"
#include "metadata.hpp"
#include "steps.hpp"

#include <memory>
#include <string>
#include <vector>

namespace gs {

enum operator_stage_type { CONVERTING_OP = 0, DISTRIBUTING_OP = 1, IMPLEMENTING_OP = 2, NONE_OP = 3 }; // operator.hpp:16-23
const char *stage_name(operator_stage_type s);

// what set_reduction_token installs: a descriptor of the reduction, not an emitter of text
// (reduction_token.hpp:7-110)
enum reduction_kind {
    RED_NONE = 0,
    RED_TOTAL_BMT,       // total_BMT_result_reduce_to_one_register_token
    RED_THREAD_BIT_MAP,  // thread_bit_map_reduce_to_two_register_token
    RED_TOTAL_WARP,      // total_warp_result_reduce_to_one_register_token
    RED_WARP_BIT_MAP,    // warp_bit_map_reduce_token
    RED_WARP_SEGMENT,    // warp_segment_reduce_token
    RED_TOTAL_TBLOCK,    // total_block_result_reduce_to_one_register_token
    RED_TBLOCK_BIT_MAP   // tblock_bit_map_reduce_token
};
struct reduction_token {
    reduction_kind kind = RED_NONE;
    bool need_warp_reduction = false;
    int sparse_coarsen_factor = 1;
    int coarsen_factor = 1;
};

struct metadata_dependency {
    POS_TYPE pos;
    std::string name;
    int sub_matrix_id;
};

// kernel plan: what compile() derives (mirrors gs_plan_desc of include/gsparse_b200.h)
struct kernel_plan {
    int family = -1;
    int sparse_coarsen = 1, coarsen = 1, vector_width = 1;
    bool need_warp_reduction = false, thread_for_row = false, has_row_permutation = false;
    unsigned grid[2] = {0, 0}, block[2] = {0, 0};
    uint64_t rows = 0, cols = 0, origin_nnz = 0, stored_nnz = 0;
    std::vector<metadata_dependency> deps;
};

// code_generator (code_generator.hpp:44-457): the surface operators touch, unchanged in name
class code_generator {
  public:
    code_generator(std::shared_ptr<meta_data_set> ms, int sub_matrix_id);
    std::shared_ptr<meta_data_set> get_metadata_set() const { return ms_; }      // code_generator.hpp:287
    int get_sub_matrix_id() const { return id_; }                                // :60
    void open_spec_level_of_paral(POS_TYPE pos);                                 // :194
    void set_reduction_token(POS_TYPE pos, const reduction_token &tok);          // :244
    void set_thread_grid(const std::vector<unsigned> &grid, const std::vector<unsigned> &block); // :115
    void set_thread_for_row(bool v) { thread_for_row_ = v; }                     // :260
    void set_interleave_storage() { interleave_ = true; }                        // :209
    void add_new_metadata_dependency(POS_TYPE pos, const std::string &name, int id); // :66
    void compile();                                                              // :265
    // :271-280: writes the arrays the kernel reads to ROOT_PATH_STR/data_source/<id>/ (one text file per array, the reference's hand-off,
    // metadata_set.cc:517-571) and returns <id>; what the reference also writes there -- the generated program text -- has no counterpart:
    // execute_binary() runs pre-built kernels on those files
    unsigned long generate_final_program(unsigned int repeat_num);
    bool compiled() const { return compiled_; }
    const kernel_plan &plan() const;
    bool level_open(POS_TYPE pos) const { return open_[pos]; }

  private:
    std::shared_ptr<meta_data_set> ms_;
    int id_;
    bool open_[5] = {false, false, false, false, false};
    reduction_token tokens_[5];
    std::vector<unsigned> grid_, block_;
    bool thread_for_row_ = false, interleave_ = false, compiled_ = false;
    std::vector<metadata_dependency> extra_deps_;
    kernel_plan plan_;
};

class basic_operator;
// operator_context (operator.hpp:191-265): history per stage and sub-matrix
class operator_context {
  public:
    std::vector<std::shared_ptr<basic_operator>> read_operator_context_arr(operator_stage_type stage, int sub_matrix_id) const;
    void add(std::shared_ptr<basic_operator> op);

  private:
    std::vector<std::shared_ptr<basic_operator>> all_;
};

class basic_operator { // operator.hpp:64-189
  public:
    basic_operator(std::string name, std::shared_ptr<code_generator> cg, operator_stage_type stage)
        : name(std::move(name)), stage(stage), code_generator_ptr(std::move(cg)),
          meta_data_set_ptr(code_generator_ptr->get_metadata_set()), target_matrix_id(code_generator_ptr->get_sub_matrix_id())
    {
    }
    virtual ~basic_operator() = default;
    virtual void run(bool check = true) = 0;
    virtual bool is_valid_according_to_metadata() { return true; }
    virtual bool is_valid_according_to_operator(const operator_context &) { return true; }
    virtual void set_padding_to_false() {}
    // what interlance_storage_operator asks the thread-level operators before it (operator.hpp:567)
    virtual bool get_is_padding_with_col_size_in_bmt() const { return false; }
    virtual bool get_is_col_padding_with_row_max_size_with_empty_row() const { return false; }
    virtual std::string convert_to_string() const;
    const std::string &get_name() const { return name; }
    operator_stage_type get_stage() const { return stage; }
    int get_target_matrix_id() const { return target_matrix_id; }
    const std::vector<std::string> &get_data_transform_sequence() const { return transform_seq; }
    bool is_run = false;

  protected:
    std::string name;
    operator_stage_type stage;
    std::shared_ptr<code_generator> code_generator_ptr;
    std::shared_ptr<meta_data_set> meta_data_set_ptr;
    int target_matrix_id;
    std::vector<std::string> transform_seq; // step records, in run order
    std::string params;                      // ",key:value" pairs the reference's convert_to_string() adds for this operator
    double pad_bound() const { return (double)get_config().as_integer("PADDING_RATE_UP_BOUND"); }
};

// operator_executer (operator_executer.hpp:10-80): validity assert -> run -> record
class operator_executer {
  public:
    operator_executer() : ctx_(std::make_shared<operator_context>()) {}
    std::shared_ptr<operator_context> get_operator_context() const { return ctx_; }
    void add_and_run(std::shared_ptr<basic_operator> op);

  private:
    std::shared_ptr<operator_context> ctx_;
};

// factory over the reference's operator class names; args = ctor arguments after code_generator_ptr
std::shared_ptr<basic_operator> make_operator(const std::string &op_name, std::shared_ptr<code_generator> cg,
                                              std::shared_ptr<operator_context> history, const std::vector<int64_t> &args);
const std::vector<std::string> &operator_names();

// execute_binary (executor.hpp:15): the arrays of data_source/<id>/ and the matrix file -> B200 kernels (csrc/handoff.cc), timing and
// GFLOP/s as code_generator.cc:582-624, the host check of :631-638; false on failure (the reference asserts)
bool execute_binary(const std::string &execute_path, const std::string &filename, float &total_execute_time, float &gflops);

// the driver's seven chains (token_test.cc:1003-1582)
void run_family_chain(int family, std::shared_ptr<code_generator> cg, operator_executer &ex, int N, int sparse_coarsen, int coarsen);

} // namespace gs
