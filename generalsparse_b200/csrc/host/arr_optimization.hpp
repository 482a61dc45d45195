// Model-driven index compression: the reference's arr_optimization.{hpp,cc} -- an offset array that follows a
// closed form is replaced by the formula in the generated kernel (code_generator.cc:2618-3240, MODEL_DRIVEN_INDEX_COMPRESS) and leaves
// the byte stream.  Here: the same five models with the same acceptance rules (checked mode), fitted on the host at upload time; the
// engine keeps the model beside the device array, kernels that can take a formula instead of a load get it (b200_nnz: first_nz_indices
// of a padded nnz-direction chain is 32 * i), and gs_plan_model_bytes() reports how many of the chain's registered bytes a model covers.
#pragma once
#include <cstdint>
#include <vector>

namespace gs {

enum arr_compress_type { NONE_COMPRESS = 0, LINEAR_COMPRESS, CONSTANT_COMPRESS, BRANCH_COMPRESS, CYCLE_LINEAR_COMPRESS, CYCLE_INCREASE_COMPRESS }; // arr_optimization.hpp:12-20

struct arr_model {
    arr_compress_type type = NONE_COMPRESS;
    uint64_t coefficient = 0, intercept = 0, cycle = 0; // linear: c * i + b; constant: b; cycle_linear: (i % cycle) * c + b; cycle_increase: i / cycle + b
    std::vector<uint64_t> low, up, value;               // branch: value[k] for low[k] <= i <= up[k]
    uint64_t eval(uint64_t i) const;
};

// the reference's init_*_compressor with need_checked = true (arr_optimization.cc): false when the array does not follow the model
bool fit_linear(const std::vector<uint64_t> &a, arr_model &m);
bool fit_constant(const std::vector<uint64_t> &a, arr_model &m);
bool fit_branch(const std::vector<uint64_t> &a, uint64_t max_branches, arr_model &m);
bool fit_cycle_linear(const std::vector<uint64_t> &a, uint64_t cycle_num, arr_model &m);
bool fit_cycle_increase(const std::vector<uint64_t> &a, arr_model &m);
// the first model that fits, in the order constant, linear, cycle_increase, cycle_linear (cycle = the first index at which the array
// falls), branch; NONE_COMPRESS if none does
arr_model fit_arr_model(const std::vector<uint64_t> &a, uint64_t max_branches);

} // namespace gs
