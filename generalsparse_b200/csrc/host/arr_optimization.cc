#include "arr_optimization.hpp"

namespace gs {

uint64_t arr_model::eval(uint64_t i) const
{
    switch (type) {
    case LINEAR_COMPRESS: return coefficient * i + intercept;
    case CONSTANT_COMPRESS: return intercept;
    case CYCLE_LINEAR_COMPRESS: return (i % cycle) * coefficient + intercept;
    case CYCLE_INCREASE_COMPRESS: return i / cycle + intercept;
    case BRANCH_COMPRESS:
        for (size_t k = 0; k < low.size(); k++)
            if (i >= low[k] && i <= up[k]) return value[k];
        return 0;
    default: return 0;
    }
}

bool fit_linear(const std::vector<uint64_t> &a, arr_model &m)
{ // init_linear_compressor: at least two entries, non-decreasing, one step throughout
    if (a.size() <= 1 || a[1] < a[0]) return false;
    const uint64_t step = a[1] - a[0];
    for (size_t i = 0; i + 1 < a.size(); i++)
        if (a[i + 1] < a[i] || a[i + 1] - a[i] != step) return false;
    m = arr_model();
    m.type = LINEAR_COMPRESS, m.coefficient = step, m.intercept = a[0];
    return true;
}

bool fit_constant(const std::vector<uint64_t> &a, arr_model &m)
{
    if (a.empty()) return false;
    for (uint64_t v : a)
        if (v != a[0]) return false;
    m = arr_model();
    m.type = CONSTANT_COMPRESS, m.intercept = a[0];
    return true;
}

bool fit_branch(const std::vector<uint64_t> &a, uint64_t max_branches, arr_model &m)
{ // init_branch_compressor: runs of equal values; gives up once the number of closed runs reaches BRANCH_COMPRESS_MAX_SIZE
    if (a.size() <= 2) return false;
    arr_model r;
    r.type = BRANCH_COMPRESS;
    r.low.push_back(0), r.value.push_back(a[0]);
    for (size_t i = 0; i < a.size(); i++)
        if (a[i] != r.value.back()) {
            r.up.push_back(i - 1), r.low.push_back(i), r.value.push_back(a[i]);
            if (r.up.size() >= max_branches) return false;
        }
    r.up.push_back(a.size() - 1);
    m = r;
    return true;
}

bool fit_cycle_linear(const std::vector<uint64_t> &a, uint64_t cycle_num, arr_model &m)
{ // init_cycle_linear_compressor: whole cycles of length cycle_num (1 < cycle_num < size), every entry = (i % cycle) * c + b
    if (cycle_num <= 1 || cycle_num >= a.size() || a.size() % cycle_num != 0 || a[1] < a[0]) return false;
    const uint64_t b = a[0], c = a[1] - a[0];
    for (size_t i = 0; i < a.size(); i++)
        if (a[i] != (i % cycle_num) * c + b) return false; // (the reference's integer-division test accepts a few more arrays; the kernel formula needs equality)
    m = arr_model();
    m.type = CYCLE_LINEAR_COMPRESS, m.cycle = cycle_num, m.coefficient = c, m.intercept = b;
    return true;
}

bool fit_cycle_increase(const std::vector<uint64_t> &a, arr_model &m)
{ // init_cycle_increase_compressor: 0 0 1 1 2 2 ...: the cycle is the length of the first run, one more per cycle
    if (a.size() <= 2) return false;
    uint64_t cycle = 0;
    for (size_t i = 0; i < a.size(); i++)
        if (a[i] != a[0]) {
            if (a[i] < a[0]) return false;
            cycle = i;
            break;
        }
    if (cycle == 0 || a.size() % cycle != 0) return false;
    for (size_t i = 0; i < a.size(); i++)
        if (a[i] != a[0] + i / cycle) return false;
    m = arr_model();
    m.type = CYCLE_INCREASE_COMPRESS, m.cycle = cycle, m.intercept = a[0];
    return true;
}

arr_model fit_arr_model(const std::vector<uint64_t> &a, uint64_t max_branches)
{
    arr_model m;
    if (fit_constant(a, m) || fit_linear(a, m) || fit_cycle_increase(a, m)) return m;
    for (size_t i = 1; i < a.size(); i++)
        if (a[i] < a[i - 1]) {
            if (fit_cycle_linear(a, i, m)) return m;
            break;
        }
    if (fit_branch(a, max_branches, m)) return m;
    return arr_model();
}

} // namespace gs
