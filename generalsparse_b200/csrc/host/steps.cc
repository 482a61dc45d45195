#include "steps.hpp"

#include <algorithm>

namespace gs {
namespace steps {

namespace {

const char *level_tag(POS_TYPE p)
{
    switch (p) {
    case TBLOCK_META: return "BMTB";
    case WARP_META: return "BMW";
    case THREAD_META: return "BMT";
    default: return "GLOBAL";
    }
}

void record(recs *rec, const std::string &name, int id, const std::string &extra = "")
{
    if (rec) rec->push_back(name + "::{name:\"" + name + "\",target_matrix_id:" + std::to_string(id) + extra + "}");
}

std::vector<uint64_t> row_nnz(const meta_data_set &ms, int id, uint64_t *row_num_out = nullptr)
{
    uint64_t row_num = ms.real_row_num(id);
    const universal_array &r = ms.get(GLOBAL_META, "nz_row_indices", id);
    if (row_num_out) *row_num_out = row_num;
    return get_nnz_of_each_row_in_spec_range(r, 0, row_num - 1, 0, r.get_len() - 1);
}

universal_array vals_like(const universal_array &old, std::vector<double> v) { return universal_array(std::move(v), old.type); }

void check_pad_rate(uint64_t after, uint64_t before, double bound, const char *who)
{
    // the reference asserts (aborts) when nnz_after/nnz_before >= PADDING_RATE_UP_BOUND
    if ((double)after / (double)before >= bound)
        throw Error(std::string(who) + ": current padding rate is " + std::to_string((double)after / (double)before) +
                    ", higher than " + std::to_string(bound));
}

} // namespace

// ------------------------------------------------------------------------------------------------ sort
void get_row_order_by_length(meta_data_set &ms, int id, recs *rec)
{
    GS_REQUIRE(!ms.is_exist(GLOBAL_META, "original_nz_row_indices", id), "the sub-matrix is already sorted");
    uint64_t row_num;
    std::vector<uint64_t> len = row_nnz(ms, id, &row_num);
    uint64_t max_len = *std::max_element(len.begin(), len.end());
    // counting sort by length, longest first, stable inside a length class (get_row_order_by_length.cc:67-90)
    std::vector<uint64_t> start(max_len + 2, 0);
    for (uint64_t l : len) start[max_len - l + 1]++;
    for (size_t i = 1; i < start.size(); i++) start[i] += start[i - 1];
    std::vector<uint64_t> order(row_num);
    for (uint64_t r = 0; r < row_num; r++) order[start[max_len - len[r]]++] = r;
    ms.add_u(GLOBAL_META, "original_nz_row_indices", id, std::move(order));
    record(rec, "get_row_order_by_length", id);
}

namespace {
// CSR offsets of the current (row-sorted) stream over the logical row count end-begin+1
std::vector<uint64_t> logical_rowptr(const meta_data_set &ms, int id, uint64_t *rows_out)
{
    uint64_t rows = ms.scalar(GLOBAL_META, "end_row_index", id) - ms.scalar(GLOBAL_META, "begin_row_index", id) + 1;
    const universal_array &r = ms.get(GLOBAL_META, "nz_row_indices", id);
    std::vector<uint64_t> n = get_nnz_of_each_row_in_spec_range(r, 0, rows - 1, 0, r.get_len() - 1);
    std::vector<uint64_t> ptr(rows + 1, 0);
    for (uint64_t i = 0; i < rows; i++) ptr[i + 1] = ptr[i] + n[i];
    *rows_out = rows;
    return ptr;
}
} // namespace

void reorder_val_by_index(meta_data_set &ms, int id, recs *rec)
{
    uint64_t rows;
    std::vector<uint64_t> ptr = logical_rowptr(ms, id, &rows);
    const universal_array &order = ms.get(GLOBAL_META, "original_nz_row_indices", id);
    GS_REQUIRE(order.get_len() == rows, "original_nz_row_indices length != row count");
    const universal_array &old = ms.get(GLOBAL_META, "nz_vals", id);
    std::vector<double> out;
    out.reserve(old.f.size());
    for (uint64_t nr = 0; nr < rows; nr++) {
        uint64_t o = order.u[nr];
        GS_REQUIRE(o < rows, "row order entry out of range");
        out.insert(out.end(), old.f.begin() + ptr[o], old.f.begin() + ptr[o + 1]);
    }
    universal_array na = vals_like(old, std::move(out));
    ms.replace(GLOBAL_META, "nz_vals", id, std::move(na));
    record(rec, "reorder_val_by_index", id);
}

void reorder_col_by_index(meta_data_set &ms, int id, recs *rec)
{
    uint64_t rows;
    std::vector<uint64_t> ptr = logical_rowptr(ms, id, &rows);
    const universal_array &order = ms.get(GLOBAL_META, "original_nz_row_indices", id);
    GS_REQUIRE(order.get_len() == rows, "original_nz_row_indices length != row count");
    const universal_array &old = ms.get(GLOBAL_META, "nz_col_indices", id);
    std::vector<uint64_t> out;
    out.reserve(old.u.size());
    for (uint64_t nr = 0; nr < rows; nr++) {
        uint64_t o = order.u[nr];
        GS_REQUIRE(o < rows, "row order entry out of range");
        out.insert(out.end(), old.u.begin() + ptr[o], old.u.begin() + ptr[o + 1]);
    }
    ms.replace_u(GLOBAL_META, "nz_col_indices", id, std::move(out));
    record(rec, "reorder_col_by_index", id);
}

void reorder_row_by_index(meta_data_set &ms, int id, recs *rec)
{
    uint64_t rows;
    std::vector<uint64_t> ptr = logical_rowptr(ms, id, &rows);
    const universal_array &order = ms.get(GLOBAL_META, "original_nz_row_indices", id);
    GS_REQUIRE(order.get_len() == rows, "original_nz_row_indices length != row count");
    std::vector<uint64_t> out;
    out.reserve(ptr[rows]);
    for (uint64_t nr = 0; nr < rows; nr++) {
        uint64_t o = order.u[nr];
        out.insert(out.end(), ptr[o + 1] - ptr[o], nr);
    }
    ms.replace_u(GLOBAL_META, "nz_row_indices", id, std::move(out));
    record(rec, "reorder_row_by_index", id);
}

void remove_empty_row_in_end_of_sub_matrix(meta_data_set &ms, int id, recs *rec)
{
    uint64_t begin = ms.scalar(GLOBAL_META, "begin_row_index", id), end = ms.scalar(GLOBAL_META, "end_row_index", id);
    uint64_t rows = end - begin + 1;
    uint64_t max_row = ms.get(GLOBAL_META, "nz_row_indices", id).u.back();
    GS_REQUIRE(max_row <= rows - 1, "row index past end_row_index");
    if (max_row >= rows - 1) return; // a step that touched nothing leaves no record (operator.hpp:163-175 set_transform_seq)
    ms.replace_u(GLOBAL_META, "end_row_index", id, {begin + max_row});
    record(rec, "remove_empty_row_in_end_of_sub_matrix", id);
}

// --------------------------------------------------------------------------------------------- padding
void modify_by_col_pad_in_sub_matrix(meta_data_set &ms, int id, uint64_t m, double bound, recs *rec)
{
    GS_REQUIRE(m >= 2, "multiple_of_each_row_size must be >= 2");
    uint64_t row_num;
    std::vector<uint64_t> len = row_nnz(ms, id, &row_num);
    const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    const universal_array &C = ms.get(GLOBAL_META, "nz_col_indices", id);
    const universal_array &V = ms.get(GLOBAL_META, "nz_vals", id);
    uint64_t nnz = R.get_len(), after = nnz;
    bool padded = false;
    for (uint64_t i = 0; i < row_num; i++)
        if (len[i] % m != 0) {
            after += (len[i] / m + 1) * m - len[i];
            check_pad_rate(after, nnz, bound, "modify_col_indices_by_col_pad_in_sub_matrix");
            padded = true;
        }
    if (padded) {
        std::vector<uint64_t> nr, nc;
        std::vector<double> nv;
        nr.reserve(after), nc.reserve(after), nv.reserve(after);
        uint64_t p = 0;
        for (uint64_t i = 0; i < row_num; i++) {
            uint64_t l = len[i];
            for (uint64_t k = 0; k < l; k++, p++) nr.push_back(i), nc.push_back(C.u[p]), nv.push_back(V.f[p]);
            if (l % m != 0) {
                uint64_t target = (l / m + 1) * m;
                uint64_t last_col = nc.back();
                for (uint64_t k = l; k < target; k++) nr.push_back(i), nc.push_back(last_col), nv.push_back(0.0);
            }
        }
        universal_array va = vals_like(V, std::move(nv));
        ms.replace_u(GLOBAL_META, "nz_col_indices", id, std::move(nc));
        ms.replace(GLOBAL_META, "nz_vals", id, std::move(va));
        ms.replace_u(GLOBAL_META, "nz_row_indices", id, std::move(nr));
    }
    if (!padded) return;
    std::string extra = ",multiple_of_each_row_size:" + std::to_string(m);
    record(rec, "modify_col_indices_by_col_pad_in_sub_matrix", id, extra);
    record(rec, "modify_vals_by_col_pad_in_sub_matrix", id, extra);
    record(rec, "modify_row_indices_by_col_pad_in_sub_matrix", id, extra);
}

void modify_by_nnz_pad(meta_data_set &ms, int id, uint64_t target, double bound, recs *rec)
{
    GS_REQUIRE(target >= 1, "nnz_target must be >= 1");
    universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    universal_array &C = ms.get(GLOBAL_META, "nz_col_indices", id);
    universal_array &V = ms.get(GLOBAL_META, "nz_vals", id);
    uint64_t nnz = C.get_len();
    if (nnz % target == 0) return;
    {
        uint64_t n2 = (nnz / target + 1) * target;
        check_pad_rate(n2, nnz, bound, "modify_col_indices_by_nnz_pad");
        C.u.resize(n2, C.u.back());
        V.f.resize(n2, 0.0);
        R.u.resize(n2, R.u.back());
    }
    std::string extra = "nnz_target:" + std::to_string(target); // (no comma: modify_*_by_nnz_pad.cc convert_to_string)
    record(rec, "modify_col_indices_by_nnz_pad", id, extra);
    record(rec, "modify_vals_by_nnz_pad", id, extra);
    record(rec, "modify_row_indices_by_nnz_pad", id, extra);
}

void modify_by_row_pad_in_sub_matrix(meta_data_set &ms, int id, uint64_t multiple, double bound, recs *rec)
{
    GS_REQUIRE(multiple >= 1, "multiple must be >= 1");
    // NB: this step uses the LOGICAL row count end-begin+1 (modify_row_indices_by_row_pad_in_sub_matrix.cc)
    uint64_t rows = ms.scalar(GLOBAL_META, "end_row_index", id) - ms.scalar(GLOBAL_META, "begin_row_index", id) + 1;
    universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    universal_array &C = ms.get(GLOBAL_META, "nz_col_indices", id);
    universal_array &V = ms.get(GLOBAL_META, "nz_vals", id);
    if (rows % multiple == 0) return;
    {
        uint64_t added = (rows / multiple + 1) * multiple - rows;
        uint64_t nnz = R.get_len();
        check_pad_rate(nnz + added, nnz, bound, "modify_col_indices_by_row_pad_in_sub_matrix");
        uint64_t last_col = C.u.back();
        for (uint64_t i = 0; i < added; i++) {
            C.u.push_back(last_col);
            V.f.push_back(0.0);
            R.u.push_back(rows + i);
        }
    }
    std::string extra = ",multiple:" + std::to_string(multiple);
    record(rec, "modify_col_indices_by_row_pad_in_sub_matrix", id, extra);
    record(rec, "modify_vals_by_row_pad_in_sub_matrix", id, extra);
    record(rec, "modify_row_indices_by_row_pad_in_sub_matrix", id, extra);
}

void modify_by_empty_pad_in_submatrix(meta_data_set &ms, int id, recs *rec)
{
    uint64_t row_num = ms.real_row_num(id);
    const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    const universal_array &C = ms.get(GLOBAL_META, "nz_col_indices", id);
    const universal_array &V = ms.get(GLOBAL_META, "nz_vals", id);
    uint64_t nnz = R.get_len();
    std::vector<uint64_t> nr, nc;
    std::vector<double> nv;
    bool padded = false;
    for (uint64_t i = 0; i < nnz; i++) {
        uint64_t r = R.u[i];
        uint64_t next = (i == nnz - 1) ? row_num : R.u[i + 1];
        nr.push_back(r), nc.push_back(C.u[i]), nv.push_back(V.f[i]);
        // every empty row between r and next gets one explicit zero that reuses this entry's column
        for (uint64_t e = r + 1; e < next; e++) {
            nr.push_back(e), nc.push_back(C.u[i]), nv.push_back(0.0);
            padded = true;
        }
    }
    if (padded) {
        universal_array va = vals_like(V, std::move(nv));
        ms.replace_u(GLOBAL_META, "nz_col_indices", id, std::move(nc));
        ms.replace(GLOBAL_META, "nz_vals", id, std::move(va));
        ms.replace_u(GLOBAL_META, "nz_row_indices", id, std::move(nr));
    }
    if (!padded) return;
    record(rec, "modify_col_indices_by_empty_pad_in_submatrix", id);
    record(rec, "modify_vals_by_empty_pad_in_submatrix", id);
    record(rec, "modify_row_indices_by_empty_pad_in_submatrix", id);
}

void modify_by_col_pad_parent_blk_to_max_row_size(meta_data_set &ms, POS_TYPE parent_pos, int id, bool with_empty, double bound, recs *rec)
{
    uint64_t row_num;
    std::vector<uint64_t> len = row_nnz(ms, id, &row_num);
    const universal_array &C = ms.get(GLOBAL_META, "nz_col_indices", id);
    const universal_array &V = ms.get(GLOBAL_META, "nz_vals", id);
    uint64_t nnz = C.get_len(), after = nnz;
    // parent blocks as row ranges; GLOBAL = the whole sub-matrix
    std::vector<uint64_t> bounds;
    if (parent_pos == GLOBAL_META) bounds = {0, row_num};
    else {
        GS_REQUIRE(parent_pos == TBLOCK_META || parent_pos == WARP_META, "parent_pos");
        bounds = ms.get(parent_pos, "first_row_indices", id).u;
    }
    std::vector<uint64_t> nr, nc;
    std::vector<double> nv;
    uint64_t p = 0;
    bool padded = false;
    for (uint64_t b = 0; b + 1 < bounds.size(); b++) {
        uint64_t mx = 0;
        for (uint64_t r = bounds[b]; r < bounds[b + 1]; r++) mx = std::max(mx, len.at(r));
        for (uint64_t r = bounds[b]; r < bounds[b + 1]; r++) {
            uint64_t l = len[r];
            for (uint64_t k = 0; k < l; k++, p++) nr.push_back(r), nc.push_back(C.u[p]), nv.push_back(V.f[p]);
            bool pad_this = with_empty ? (mx != 0) : (mx != 0 && l != 0);
            if (pad_this && l < mx) {
                after += mx - l;
                check_pad_rate(after, nnz, bound, "modify_col_indices_by_col_pad_parent_blk_to_max_row_size");
                uint64_t last_col = nc.empty() ? C.u[0] : nc.back();
                for (uint64_t k = l; k < mx; k++) nr.push_back(r), nc.push_back(last_col), nv.push_back(0.0);
                padded = true;
            }
        }
    }
    GS_REQUIRE(p == nnz, "parent blocks do not cover every non-zero");
    if (padded) {
        universal_array va = vals_like(V, std::move(nv));
        ms.replace_u(GLOBAL_META, "nz_col_indices", id, std::move(nc));
        ms.replace(GLOBAL_META, "nz_vals", id, std::move(va));
        ms.replace_u(GLOBAL_META, "nz_row_indices", id, std::move(nr));
    }
    if (!padded) return;
    const std::string extra = std::string(",parent_pos:") + pos_name(parent_pos);
    record(rec, "modify_col_indices_by_col_pad_parent_blk_to_max_row_size", id, extra);
    record(rec, "modify_vals_by_col_pad_parent_blk_to_max_row_size", id, extra);
    record(rec, "modify_row_indices_by_col_pad_parent_blk_to_max_row_size", id, extra);
}

// ---------------------------------------------------------------------------- fixed blocking, row direction
static std::string row_dir_name(const char *what, POS_TYPE pos, const char *suffix)
{
    std::string lvl = pos == TBLOCK_META ? "BMTBs" : level_tag(pos);
    return std::string("get_begin_") + what + "_of_" + lvl + "_after_fixed_blocking_in_row_direction" + suffix;
}

void get_begin_rows_after_fixed_blocking_in_row_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t b, recs *rec)
{
    GS_REQUIRE(b >= 1, "fixed_row_block_size must be >= 1");
    uint64_t row_num = ms.real_row_num(id);
    std::vector<uint64_t> v;
    // [0, b, 2b, ..., row_num]; the BMT and BMW/BMTB reference loops give the same sequence
    for (uint64_t r = 0; r < row_num; r += b) v.push_back(r);
    v.push_back(row_num);
    ms.add_u(pos, "first_row_indices", id, std::move(v));
    record(rec, row_dir_name("rows", pos, pos == WARP_META ? "_without_BMTB" : ""), id, ",fixed_row_block_size:" + std::to_string(b));
}

void get_begin_nzs_after_fixed_blocking_in_row_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t b, recs *rec)
{
    GS_REQUIRE(b >= 1, "fixed_row_block_size must be >= 1");
    uint64_t row_num;
    std::vector<uint64_t> len = row_nnz(ms, id, &row_num);
    std::vector<uint64_t> v{0};
    uint64_t acc = 0;
    for (uint64_t r = 0; r < row_num; r++) {
        acc += len[r];
        if ((r + 1) % b == 0 || r == row_num - 1) v.push_back(acc);
    }
    ms.add_u(pos, "first_nz_indices", id, std::move(v));
    record(rec, row_dir_name("nzs", pos, pos == WARP_META ? "_without_BMTB" : ""), id, ",fixed_row_block_size:" + std::to_string(b));
}

void get_begin_rows_after_fixed_blocking_in_row_direction_in_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t b, recs *rec)
{
    const universal_array &P = ms.get(parent, "first_row_indices", id);
    std::vector<uint64_t> v;
    for (uint64_t i = 0; i + 1 < P.get_len(); i++)
        for (uint64_t r = P.u[i]; r < P.u[i + 1]; r += b) v.push_back(r);
    v.push_back(P.u.back());
    ms.add_u(pos, "first_row_indices", id, std::move(v));
    record(rec, row_dir_name("rows", pos, parent == TBLOCK_META ? "_in_BMTB" : "_in_BMW"), id, ",fixed_row_block_size:" + std::to_string(b));
}

void get_begin_nzs_after_fixed_blocking_in_row_direction_in_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t b, recs *rec)
{
    const universal_array &P = ms.get(parent, "first_row_indices", id);
    std::vector<uint64_t> len = row_nnz(ms, id);
    std::vector<uint64_t> v{0};
    for (uint64_t i = 0; i + 1 < P.get_len(); i++)
        for (uint64_t r0 = P.u[i]; r0 < P.u[i + 1]; r0 += b) {
            uint64_t n = 0;
            for (uint64_t r = r0; r < P.u[i + 1] && r < r0 + b; r++) n += len.at(r);
            v.push_back(v.back() + n);
        }
    ms.add_u(pos, "first_nz_indices", id, std::move(v));
    record(rec, row_dir_name("nzs", pos, parent == TBLOCK_META ? "_in_BMTB" : "_in_BMW"), id, ",fixed_row_block_size:" + std::to_string(b));
}

void get_begin_rows_after_fixed_blocking_in_row_direction_relative_to_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t b, recs *rec)
{
    const universal_array &P = ms.get(parent, "first_row_indices", id);
    std::vector<uint64_t> v;
    for (uint64_t i = 0; i + 1 < P.get_len(); i++)
        for (uint64_t r = P.u[i]; r < P.u[i + 1]; r += b) v.push_back(r - P.u[i]);
    const char *sfx = parent == TBLOCK_META ? "_relative_to_BMTB" : "_relative_to_BMW";
    ms.add_u(pos, std::string("first_row_indices") + sfx, id, std::move(v));
    record(rec, row_dir_name("rows", pos, sfx), id, ",fixed_row_block_size:" + std::to_string(b));
}

void get_begin_nzs_after_fixed_blocking_in_row_direction_relative_to_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t b, recs *rec)
{
    const universal_array &P = ms.get(parent, "first_row_indices", id);
    std::vector<uint64_t> len = row_nnz(ms, id);
    std::vector<uint64_t> v;
    for (uint64_t i = 0; i + 1 < P.get_len(); i++) {
        uint64_t n = 0;
        for (uint64_t r = P.u[i]; r < P.u[i + 1]; r++) {
            if ((r - P.u[i]) % b == 0) v.push_back(n);
            n += len.at(r);
        }
    }
    const char *sfx = parent == TBLOCK_META ? "_relative_to_BMTB" : "_relative_to_BMW";
    ms.add_u(pos, std::string("first_nz_indices") + sfx, id, std::move(v));
    record(rec, row_dir_name("nzs", pos, sfx), id, ",fixed_row_block_size:" + std::to_string(b));
}

static const char *first_child_name(POS_TYPE child) { return child == WARP_META ? "first_BMW_indices" : "first_BMT_indices"; }

void get_begin_children_of_parent_after_blocking_in_row_direction(meta_data_set &ms, POS_TYPE child, POS_TYPE parent, int id, recs *rec)
{
    const universal_array &P = ms.get(parent, "first_row_indices", id);
    const universal_array &Cc = ms.get(child, "first_row_indices", id);
    std::vector<uint64_t> v{0};
    uint64_t c = 0;
    for (uint64_t i = 0; i + 1 < P.get_len(); i++) {
        GS_REQUIRE(c < Cc.get_len() && Cc.u[c] == P.u[i], "child blocks do not start at the parent's first row");
        uint64_t n = 0;
        while (c < Cc.get_len() && Cc.u[c] < P.u[i + 1]) n++, c++;
        v.push_back(v.back() + n);
    }
    ms.add_u(parent, first_child_name(child), id, std::move(v));
    record(rec, child == WARP_META ? "get_begin_BMWs_of_BMTB_after_blocking_in_row_direction"
                                   : "get_begin_BMTs_of_specific_parent_after_blocking_in_row_direction", id,
           child == WARP_META ? ",fixed_row_block_size:" : std::string(",parent_pos:") + pos_name(parent));
}

void get_begin_children_of_parent_after_blocking(meta_data_set &ms, POS_TYPE child, POS_TYPE parent, int id, recs *rec)
{
    const universal_array &P = ms.get(parent, "first_nz_indices", id);
    const universal_array &Cc = ms.get(child, "first_nz_indices", id);
    std::vector<uint64_t> v{0};
    uint64_t c = 0;
    for (uint64_t i = 0; i + 1 < P.get_len(); i++) {
        GS_REQUIRE(c < Cc.get_len() && Cc.u[c] == P.u[i], "child blocks do not start at the parent's first non-zero");
        uint64_t n = 0;
        while (c < Cc.get_len() && Cc.u[c] < P.u[i + 1]) n++, c++;
        v.push_back(v.back() + n);
    }
    ms.add_u(parent, first_child_name(child), id, std::move(v));
    record(rec, child == WARP_META ? "get_begin_BMWs_of_BMTB_after_blocking" : "get_begin_BMTs_of_specific_parent_after_blocking", id,
           child == WARP_META ? "" : std::string(",parent_pos:") + pos_name(parent));
}

// ---------------------------------------------------------------------------- fixed blocking, col direction
void get_begin_rows_after_fixed_blocking_in_col_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t c, recs *rec)
{
    GS_REQUIRE(c >= 1, "col_size must be >= 1");
    uint64_t row_num;
    std::vector<uint64_t> len = row_nnz(ms, id, &row_num);
    std::vector<uint64_t> v;
    for (uint64_t r = 0; r < row_num; r++) v.insert(v.end(), (len[r] + c - 1) / c, r);
    ms.add_u(pos, "first_row_indices_without_ending", id, std::move(v));
    record(rec, std::string("get_begin_rows_of_") + level_tag(pos) + "_after_fixed_blocking_in_col_direction", id, ",col_size:" + std::to_string(c));
}

void get_begin_nzs_after_fixed_blocking_in_col_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t c, recs *rec)
{
    GS_REQUIRE(c >= 1, "col_size must be >= 1");
    uint64_t row_num;
    std::vector<uint64_t> len = row_nnz(ms, id, &row_num);
    std::vector<uint64_t> v{0};
    for (uint64_t r = 0; r < row_num; r++) {
        uint64_t remain = len[r];
        while (remain > 0) {
            uint64_t take = remain < c ? remain : c;
            v.push_back(v.back() + take);
            remain -= take;
        }
    }
    ms.add_u(pos, "first_nz_indices", id, std::move(v));
    record(rec, std::string("get_begin_nzs_of_") + level_tag(pos) + "_after_fixed_blocking_in_col_direction", id, ",col_size:" + std::to_string(c));
}

// ---------------------------------------------------------------------------- fixed blocking, nnz direction
// BMT offsets relative to the parent block that holds them (get_begin_rows_of_BMT_after_fixed_blocking_in_col_direction_
// relative_to_{BMTB,BMW}.cc, get_begin_nzs_of_BMT_after_fixed_blocking_in_col_direction_relative_to_parents.cc):
// one entry per BMT, no closing sentinel
void get_begin_rows_after_fixed_blocking_in_col_direction_relative_to_parent(meta_data_set &ms, POS_TYPE parent, int id, uint64_t c, recs *rec)
{
    GS_REQUIRE(parent == TBLOCK_META || parent == WARP_META, "parent_pos");
    std::vector<uint64_t> len = row_nnz(ms, id);
    const universal_array &P = ms.get(parent, "first_row_indices", id);
    std::vector<uint64_t> v;
    for (uint64_t b = 0; b + 1 < P.get_len(); b++)
        for (uint64_t r = P.u[b]; r < P.u[b + 1]; r++)
            for (uint64_t k = 0; k < (len.at(r) + c - 1) / c; k++) v.push_back(r - P.u[b]);
    const char *sfx = parent == TBLOCK_META ? "_relative_to_BMTB" : "_relative_to_BMW";
    ms.add_u(THREAD_META, std::string("first_row_indices") + sfx, id, std::move(v));
    record(rec, std::string("get_begin_rows_of_BMT_after_fixed_blocking_in_col_direction") + sfx, id, ",col_size:" + std::to_string(c));
}

void get_begin_nzs_after_fixed_blocking_in_col_direction_relative_to_parents(meta_data_set &ms, POS_TYPE parent, int id, uint64_t c, recs *rec)
{
    GS_REQUIRE(parent == TBLOCK_META || parent == WARP_META, "parent_pos");
    std::vector<uint64_t> len = row_nnz(ms, id);
    const universal_array &P = ms.get(parent, "first_row_indices", id);
    std::vector<uint64_t> v;
    for (uint64_t b = 0; b + 1 < P.get_len(); b++) {
        uint64_t off = 0;
        for (uint64_t r = P.u[b]; r < P.u[b + 1]; r++)
            for (uint64_t left = len.at(r); left > 0;) {
                v.push_back(off);
                uint64_t take = std::min(left, c);
                off += take, left -= take;
            }
    }
    ms.add_u(THREAD_META, std::string("first_nz_indices_relative_to_") + (parent == TBLOCK_META ? "BMTB" : "BMW"), id, std::move(v));
    record(rec, "get_begin_nzs_of_BMT_after_fixed_blocking_in_col_direction_relative_to_parents", id,
           ",col_size:" + std::to_string(c) + ",parent_pos:" + pos_name(parent));
}

void get_begin_rows_after_fixed_blocking_in_nnz_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t k, recs *rec)
{
    GS_REQUIRE(k >= 1, "nnz_per_block must be >= 1");
    uint64_t row_num = ms.real_row_num(id);
    const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    std::vector<uint64_t> v;
    for (uint64_t i = 0; i < R.get_len(); i += k) v.push_back(R.u[i]);
    v.push_back(row_num);
    ms.add_u(pos, "first_row_indices", id, std::move(v));
    record(rec, std::string("get_begin_rows_of_") + level_tag(pos) + "_after_fixed_blocking_in_nnz_direction", id, ",nnz_per_" + std::string(level_tag(pos)) + ":" + std::to_string(k));
}

void get_begin_nzs_after_fixed_blocking_in_nnz_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t k, recs *rec)
{
    GS_REQUIRE(k >= 1, "nnz_per_block must be >= 1");
    uint64_t nnz = ms.get(GLOBAL_META, "nz_row_indices", id).get_len();
    std::vector<uint64_t> v;
    for (uint64_t i = 0; i < nnz; i += k) v.push_back(i);
    v.push_back(nnz);
    ms.add_u(pos, "first_nz_indices", id, std::move(v));
    record(rec, std::string("get_begin_nzs_of_") + level_tag(pos) + "_after_fixed_blocking_in_nnz_direction", id, ",nnz_per_" + std::string(level_tag(pos)) + ":" + std::to_string(k));
}

void get_begin_rows_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB(meta_data_set &ms, int id, uint64_t k, recs *rec)
{
    GS_REQUIRE(k >= 1, "nnz_per_block must be >= 1");
    const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    const universal_array &PR = ms.get(TBLOCK_META, "first_row_indices", id), &PN = ms.get(TBLOCK_META, "first_nz_indices", id);
    std::vector<uint64_t> v;
    uint64_t p = 0;
    for (uint64_t i = 0; i < R.get_len(); i += k) {
        GS_REQUIRE(p + 1 < PN.get_len(), "BMW beyond the last BMTB");
        if (i >= PN.u[p + 1]) p++;
        GS_REQUIRE(p < PR.get_len() && R.u[i] >= PR.u[p], "BMW starts above its BMTB's first row");
        v.push_back(R.u[i] - PR.u[p]);
    }
    ms.add_u(WARP_META, "first_row_indices_relative_to_BMTB", id, std::move(v));
    record(rec, "get_begin_rows_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB", id, ",nnz_per_BMW:" + std::to_string(k));
}

void get_begin_nzs_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB(meta_data_set &ms, int id, uint64_t k, recs *rec)
{
    GS_REQUIRE(k >= 1, "nnz_per_block must be >= 1");
    const uint64_t nnz = ms.get(GLOBAL_META, "nz_row_indices", id).get_len();
    const universal_array &PN = ms.get(TBLOCK_META, "first_nz_indices", id);
    std::vector<uint64_t> v;
    uint64_t p = 0;
    for (uint64_t i = 0; i < nnz; i += k) {
        GS_REQUIRE(p + 1 < PN.get_len(), "BMW beyond the last BMTB");
        if (i >= PN.u[p + 1]) {
            GS_REQUIRE(i == PN.u[p + 1], "nnz_per_BMTB is not a multiple of nnz_per_BMW"); // the reference asserts the same
            p++;
        }
        v.push_back(i - PN.u[p]);
    }
    ms.add_u(WARP_META, "first_nz_indices_relative_to_BMTB", id, std::move(v));
    record(rec, "get_begin_nzs_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB", id, ",nnz_per_BMW:" + std::to_string(k));
}

// --------------------------------------------------------------------------------------------------- sizes
void get_BMW_size_of_global(meta_data_set &ms, int id, recs *rec)
{
    const universal_array &W = ms.get(WARP_META, "first_nz_indices", id);
    GS_REQUIRE(W.get_len() >= 2, "no BMW");
    const uint64_t size = W.u[1] - W.u[0];
    for (uint64_t w = 0; w + 1 < W.get_len(); w++) {
        GS_REQUIRE(W.u[w + 1] > W.u[w], "empty BMW");
        if (W.u[w + 1] - W.u[w] != size) return; // not uniform: nothing emitted, nothing recorded
    }
    ms.add_u(GLOBAL_META, "BMW_size_of_each_blk", id, std::vector<uint64_t>{size});
    record(rec, "get_BMW_size_of_each_parent", id, ",parent_pos:GLOBAL_META");
}

void get_BMTB_size(meta_data_set &ms, int id, recs *rec)
{
    const universal_array &T = ms.get(TBLOCK_META, "first_nz_indices", id);
    GS_REQUIRE(T.get_len() >= 2, "no BMTB");
    const uint64_t size = T.u[1] - T.u[0];
    for (uint64_t t = 0; t + 1 < T.get_len(); t++) GS_REQUIRE(T.u[t + 1] > T.u[t] && T.u[t + 1] - T.u[t] == size, "the BMTB sizes are not the same");
    ms.add_u(GLOBAL_META, "BMTB_size_of_each_blk", id, std::vector<uint64_t>{size});
    record(rec, "get_BMTB_size", id, ",parent_pos:GLOBAL_META");
}

void get_BMT_size_of_each_parent(meta_data_set &ms, POS_TYPE parent_pos, int id, bool row_direction_blocking, recs *rec)
{
    const universal_array &T = ms.get(THREAD_META, "first_nz_indices", id);
    std::vector<uint64_t> out;
    if (parent_pos == GLOBAL_META) {
        uint64_t size = 0;
        bool seen = false;
        for (uint64_t t = 0; t + 1 < T.get_len(); t++) {
            uint64_t s = T.u[t + 1] - T.u[t];
            if (!seen) size = s, seen = true;
            else if (s != size) return; // not uniform: the reference prints a note and emits nothing
        }
        GS_REQUIRE(seen, "no BMT");
        out.push_back(size);
    } else {
        const universal_array &P = ms.get(parent_pos, "first_nz_indices", id);
        const universal_array *FB = row_direction_blocking ? &ms.get(parent_pos, "first_BMT_indices", id) : nullptr;
        uint64_t t = 0;
        for (uint64_t p = 0; p + 1 < P.get_len(); p++) {
            uint64_t lo = P.u[p], hi = P.u[p + 1];
            GS_REQUIRE(t < T.get_len() && T.u[t] == lo, "BMT does not start at the parent block's first non-zero");
            uint64_t size = 0;
            bool seen = false;
            uint64_t cur = T.u[t];
            while (cur < hi) {
                GS_REQUIRE(t + 1 < T.get_len(), "BMT offsets end inside a parent block");
                uint64_t s = T.u[t + 1] - T.u[t];
                if (!seen) size = s, seen = true;
                else if (s != size) return;
                t++;
                cur = T.u[t];
            }
            if (FB && lo == hi)
                while (t < FB->u[p + 1]) t++;
            // (the reference asserts that at least one BMT was seen unless the parent is empty)
            out.push_back(size);
        }
    }
    ms.add_u(parent_pos, "BMT_size_of_each_blk", id, std::move(out));
    record(rec, "get_BMT_size_of_each_parent", id, std::string(",parent_pos:") + pos_name(parent_pos)); // nothing emitted, nothing recorded
}

// -------------------------------------------------------------------------------------- bit maps / segments
static std::vector<uint8_t> row_start_bits(const universal_array &R)
{
    std::vector<uint8_t> bits(R.get_len());
    bits[0] = 1;
    for (uint64_t j = 1; j < R.get_len(); j++) bits[j] = R.u[j] != R.u[j - 1];
    return bits;
}

void thread_bit_map(meta_data_set &ms, bool parent_flag, uint64_t parent_size, int id, recs *rec)
{
    const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    const universal_array &T = ms.get(THREAD_META, "first_nz_indices", id);
    std::vector<uint8_t> bits = row_start_bits(R);
    if (parent_flag) {
        GS_REQUIRE(ms.is_exist(GLOBAL_META, "BMT_size_of_each_blk", id), "thread_bit_map with a parent needs uniform BMT sizes");
        GS_REQUIRE(parent_size >= 1, "parent size");
        uint64_t thread_size = T.u[1] - T.u[0];
        // the first non-zero of every `parent_size`-th BMT starts a segment (thread_bit_map.cc: parent_flag loop;
        // the reference loop may step one BMT past the end, which writes nothing observable)
        for (uint64_t i = 0; i < bits.size() / thread_size + 1; i += parent_size)
            if (i * thread_size < bits.size()) bits[i * thread_size] = 1;
    }
    std::vector<uint64_t> out;
    out.reserve(T.get_len() - 1);
    for (uint64_t i = 0; i + 1 < T.get_len(); i++) {
        uint64_t lo = T.u[i], hi = T.u[i + 1];
        GS_REQUIRE(hi - lo <= 64, "a BMT wider than 64 non-zeros does not fit the u64 bit map");
        uint64_t map = 0;
        for (uint64_t j = hi; j-- > lo;) map = (map << 1) | bits[j]; // LSB = first non-zero of the BMT
        out.push_back(map);
    }
    ms.add_u(THREAD_META, "thread_bit_map", id, std::move(out));
    record(rec, "thread_bit_map", id, ",fixed_row_block_size:");
}

void parent_bit_map_of_thread(meta_data_set &ms, POS_TYPE pos, int id, uint64_t vector_width, recs *rec)
{
    const universal_array &F = ms.get(THREAD_META, "first_row_indices_without_ending", id);
    std::vector<uint8_t> bits = row_start_bits(F);
    std::vector<uint64_t> out;
    POS_TYPE store = pos;
    if (pos == WARP_META) {
        GS_REQUIRE(vector_width >= 1 && vector_width <= 64, "VECTOR_WIDTH must be in 1..64");
        for (uint64_t i = 0; i < bits.size(); i += vector_width) bits[i] = 1;
        for (uint64_t i = 0; i < bits.size(); i += vector_width) {
            uint64_t k = std::min<uint64_t>(i + vector_width, bits.size());
            uint64_t map = 0;
            for (uint64_t j = k; j-- > i;) map = (map << 1) | bits[j];
            out.push_back(map);
        }
    } else if (pos == TBLOCK_META) {
        const universal_array &FB = ms.get(TBLOCK_META, "first_BMT_indices", id);
        for (uint64_t i = 0; i < FB.get_len(); i++)
            if (FB.u[i] < bits.size()) bits[FB.u[i]] = 1; // the end sentinel falls one past the array in the reference
        out.assign(bits.begin(), bits.end());
        store = THREAD_META; // one 0/1 per BMT, stored at THREAD level (parent_bit_map_of_thread.cc: pos reassigned)
    } else {
        throw Error("parent_bit_map_of_thread: pos must be WARP_META or TBLOCK_META");
    }
    ms.add_u(store, "bit_map_of_thread", id, std::move(out));
    record(rec, "parent_bit_map_of_thread", id, ",fixed_row_block_size:");
}

void segment_empty_flag(meta_data_set &ms, POS_TYPE, uint64_t size, int id, recs *rec)
{
    GS_REQUIRE(size >= 1 && size <= 64, "segment_empty_flag packs at most 64 BMTs per word");
    const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    const universal_array &T = ms.get(THREAD_META, "first_nz_indices", id);
    std::vector<uint8_t> flag;
    for (uint64_t j = 0; j + 1 < T.get_len(); j++) {
        uint8_t f = 0;
        for (uint64_t i = T.u[j] + 1; i < T.u[j + 1]; i++)
            if (R.u[i] - R.u[i - 1] > 1) {
                f = 1;
                break;
            }
        flag.push_back(f);
    }
    std::vector<uint64_t> out;
    for (uint64_t i = 0; i < flag.size(); i += size) {
        uint64_t k = std::min<uint64_t>(i + size, flag.size());
        uint64_t w = 0;
        for (uint64_t j = k; j-- > i;) w = (w << 1) | flag[j];
        out.push_back(w);
    }
    ms.add_u(THREAD_META, "segment_empty_flag", id, std::move(out));
    record(rec, "segment_empty_flag", id, ",fixed_row_block_size:");
}

void segment_empty_row_indices(meta_data_set &ms, POS_TYPE, int id, recs *rec)
{
    const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    const universal_array &T = ms.get(THREAD_META, "first_nz_indices", id);
    std::vector<uint64_t> out;
    for (uint64_t j = 0; j + 1 < T.get_len(); j++) {
        uint64_t first_row = R.u[T.u[j]];
        out.push_back(0);
        for (uint64_t i = T.u[j] + 1; i < T.u[j + 1]; i++)
            if (R.u[i] != R.u[i - 1]) out.push_back(R.u[i] - first_row);
    }
    ms.add_u(THREAD_META, "segment_empty_row_indices", id, std::move(out));
    record(rec, "segment_empty_row_indices", id, ",fixed_row_block_size:");
}

void segment_offset(meta_data_set &ms, bool parent_flag, uint64_t size, int id, recs *rec)
{
    const universal_array *B = nullptr;
    if (ms.is_exist(THREAD_META, "thread_bit_map", id)) B = &ms.get(THREAD_META, "thread_bit_map", id);
    else if (ms.is_exist(THREAD_META, "bit_map_of_thread", id)) B = &ms.get(THREAD_META, "bit_map_of_thread", id);
    GS_REQUIRE(B != nullptr, "segment_offset needs thread_bit_map or bit_map_of_thread");
    GS_REQUIRE(size >= 1, "size");
    std::vector<uint64_t> out(B->get_len(), 0);
    uint64_t count = 0, prev = 0;
    for (uint64_t j = 1; j < B->get_len(); j++) {
        if (B->u[j] == 0 && ((j % size != 0) || !parent_flag)) count++;
        else {
            out[prev] = count;
            count = 0;
            prev = j;
        }
    }
    // NB: the run that reaches the end of the array is never written back (segment_offset.cc) -- kept
    ms.add_u(THREAD_META, "segment_offset", id, std::move(out));
    record(rec, "segment_offset", id, ",fixed_row_block_size:");
}

void segment_ptr(meta_data_set &ms, POS_TYPE, int id, recs *rec)
{
    const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    const universal_array &T = ms.get(THREAD_META, "first_nz_indices", id);
    std::vector<uint64_t> out{0};
    uint64_t count = 0;
    for (uint64_t j = 0; j + 2 < T.get_len(); j++) {
        count++;
        for (uint64_t i = T.u[j] + 1; i < T.u[j + 1]; i++)
            if (R.u[i] != R.u[i - 1]) count++;
        out.push_back(count);
    }
    ms.add_u(THREAD_META, "segment_ptr", id, std::move(out));
    record(rec, "segment_ptr", id, ",fixed_row_block_size:");
}

// -------------------------------------------------------------------------------------- merge-thread grouping
static const universal_array &bmt_first_rows(const meta_data_set &ms, int id)
{
    if (ms.is_exist(THREAD_META, "first_row_indices", id)) return ms.get(THREAD_META, "first_row_indices", id);
    return ms.get(THREAD_META, "first_row_indices_without_ending", id);
}

static std::vector<uint64_t> sample_with_last(const universal_array &A, uint64_t merge_num)
{
    GS_REQUIRE(merge_num >= 1, "merge_num must be >= 1");
    std::vector<uint64_t> v;
    for (uint64_t j = 0; j + 1 < A.get_len(); j += merge_num) v.push_back(A.u[j]);
    v.push_back(A.u.back());
    return v;
}

void get_begin_rows_after_merge_thread(meta_data_set &ms, POS_TYPE pos, uint64_t merge_num, int id, recs *rec)
{
    ms.add_u(pos, "first_row_indices", id, sample_with_last(bmt_first_rows(ms, id), merge_num));
    record(rec, "get_begin_rows_after_merge_thread", id);
}

void get_begin_nzs_after_merge_thread(meta_data_set &ms, POS_TYPE pos, uint64_t merge_num, int id, recs *rec)
{
    ms.add_u(pos, "first_nz_indices", id, sample_with_last(ms.get(THREAD_META, "first_nz_indices", id), merge_num));
    record(rec, "get_begin_nzs_after_merge_thread", id);
}

static std::vector<uint64_t> relative_in_groups(const universal_array &A, uint64_t merge_num)
{
    GS_REQUIRE(merge_num >= 1, "merge_num must be >= 1");
    std::vector<uint64_t> v;
    for (uint64_t j = 0; j + 1 < A.get_len(); j += merge_num)
        for (uint64_t i = j; i < j + merge_num && i + 1 < A.get_len(); i++) v.push_back(A.u[i] - A.u[j]);
    return v;
}

void get_begin_rows_relative_to_parent_after_merge_thread(meta_data_set &ms, POS_TYPE pos, uint64_t merge_num, int id, recs *rec)
{
    GS_REQUIRE(pos == WARP_META || pos == TBLOCK_META, "pos");
    // stored at the PARENT level in the reference (get_begin_rows_relative_to_parent_after_merge_thread.cc)
    ms.add_u(pos, pos == WARP_META ? "first_row_indices_relative_to_BMW" : "first_row_indices_relative_to_BMTB", id,
             relative_in_groups(bmt_first_rows(ms, id), merge_num));
    record(rec, "get_begin_rows_relative_to_parent_after_merge_thread", id);
}

void get_begin_nzs_relative_to_parent_after_merge_thread(meta_data_set &ms, POS_TYPE pos, uint64_t merge_num, int id, recs *rec)
{
    GS_REQUIRE(pos == WARP_META || pos == TBLOCK_META, "pos");
    // stored at THREAD level in the reference (get_begin_nzs_relative_to_parent_after_merge_thread.cc)
    ms.add_u(THREAD_META, pos == WARP_META ? "first_nz_indices_relative_to_BMW" : "first_nz_indices_relative_to_BMTB", id,
             relative_in_groups(ms.get(THREAD_META, "first_nz_indices", id), merge_num));
    record(rec, "get_begin_nzs_relative_to_parent_after_merge_thread", id);
}

void get_begin_BMTs_after_merge_thread(meta_data_set &ms, POS_TYPE pos, uint64_t merge_num, int id, recs *rec)
{
    GS_REQUIRE(merge_num >= 1, "merge_num must be >= 1");
    uint64_t n = ms.get(THREAD_META, "first_nz_indices", id).get_len() - 1;
    std::vector<uint64_t> v;
    for (uint64_t j = 0; j < n; j += merge_num) v.push_back(j);
    v.push_back(n);
    ms.add_u(pos, "first_BMT_indices", id, std::move(v));
    record(rec, "get_begin_BMTs_after_merge_thread", id);
}

// ------------------------------------------------------------------------------------------ balanced blocking
std::vector<uint64_t> balanced_begin_rows(const std::vector<uint64_t> &len, uint64_t per)
{
    // greedy "whole rows until >= nnz_per_interval" (data_transform_common.cc:960-988)
    std::vector<uint64_t> v{0};
    uint64_t cnt = 0, rows = len.size();
    for (uint64_t i = 0; i < rows; i++) {
        cnt += len[i];
        if (cnt >= per) v.push_back(i + 1), cnt = 0;
    }
    if (v.back() < rows) v.push_back(rows);
    return v;
}

std::vector<uint64_t> balanced_begin_nzs(const std::vector<uint64_t> &len, uint64_t per)
{
    // data_transform_common.cc:934-958
    std::vector<uint64_t> v{0};
    uint64_t cnt = 0, total = 0;
    for (uint64_t i = 0; i < len.size(); i++) {
        cnt += len[i];
        total += len[i];
        if (cnt >= per) v.push_back(total), cnt = 0;
    }
    if (cnt != 0) v.push_back(total);
    return v;
}

void get_begin_rows_after_nnz_blocking_in_row_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t per, recs *rec)
{
    GS_REQUIRE(per >= 1, "nnz_per_interval must be >= 1");
    std::vector<uint64_t> len = row_nnz(ms, id);
    std::vector<uint64_t> cuts = balanced_begin_rows(len, per);
    if (cuts.size() >= 3) { // a closing block that the greedy rule did not fill: it must hold something
        uint64_t tail = 0;
        for (uint64_t r = cuts[cuts.size() - 2]; r < cuts.back(); r++) tail += len[r];
        GS_REQUIRE(tail != 0, "the rows after the last interval are all empty (data_transform_common.cc:984 asserts nz_count != 0)");
    }
    ms.add_u(pos, "first_row_indices", id, std::move(cuts));
    record(rec, std::string("get_begin_rows_of_") + level_tag(pos) + "_after_nnz_blocking_in_row_direction", id, ",nnz_per_interval:" + std::to_string(per));
}

void get_begin_nzs_after_nnz_blocking_in_row_direction(meta_data_set &ms, POS_TYPE pos, int id, uint64_t per, recs *rec)
{
    GS_REQUIRE(per >= 1, "nnz_per_interval must be >= 1");
    std::vector<uint64_t> len = row_nnz(ms, id);
    ms.add_u(pos, "first_nz_indices", id, balanced_begin_nzs(len, per));
    record(rec, std::string("get_begin_nzs_of_") + level_tag(pos) + "_after_nnz_blocking_in_row_direction", id, ",nnz_per_interval:" + std::to_string(per));
}

// ---- merge path (get_begin_{rows,nzs}_of_level_after_merge_path.cc) ------------------------------------------------------
// The path walks the non-empty rows: row j (the j-th non-empty one) ends at path position  sum(nnz of rows 0..j) + j.
// Part p starts at path position p * work_size: in the first row whose end lies beyond it, at non-zero  position - j.
// The reference searches that row from the start for every part (get_begin_rows_of_level_after_merge_path.cc:82-93:
// parts x rows steps -- ~10^13 at 10 M rows and work 128); both sequences are monotone, so one two-pointer walk gives
// the same arrays in rows + parts steps.
static void merge_path_walk(const std::vector<uint64_t> &len, uint64_t work, std::vector<uint64_t> *rows_out, std::vector<uint64_t> *nzs_out)
{
    std::vector<uint64_t> path_end, path_row;
    uint64_t count = 0;
    for (uint64_t i = 0; i < len.size(); i++) {
        if (len[i] == 0) continue;
        count += len[i] + (path_end.empty() ? 0 : 1);
        path_end.push_back(count), path_row.push_back(i);
    }
    uint64_t j = 0;
    for (uint64_t pos = 0; pos < count; pos += work) {
        while (j < path_end.size() && path_end[j] <= pos) j++;
        if (j == path_end.size()) break; // (cannot happen: pos < count = path_end.back())
        if (rows_out) rows_out->push_back(path_row[j]);
        if (nzs_out) nzs_out->push_back(pos - j);
    }
}

void get_begin_rows_of_level_after_merge_path(meta_data_set &ms, POS_TYPE pos, int id, uint64_t work, recs *rec)
{
    GS_REQUIRE(work >= 1, "work_size must be >= 1");
    std::vector<uint64_t> v;
    merge_path_walk(row_nnz(ms, id), work, &v, nullptr);
    GS_REQUIRE(!v.empty(), "merge path over a matrix without non-zeros");
    ms.add_u(pos, "first_row_indices_without_ending", id, std::move(v));
    record(rec, "get_begin_rows_of_level_after_merge_path", id, ",work_size:" + std::to_string(work));
}

void get_begin_nzs_of_level_after_merge_path(meta_data_set &ms, POS_TYPE pos, int id, uint64_t work, recs *rec)
{
    GS_REQUIRE(work >= 1, "work_size must be >= 1");
    std::vector<uint64_t> v;
    merge_path_walk(row_nnz(ms, id), work, nullptr, &v);
    v.push_back(ms.get(GLOBAL_META, "nz_row_indices", id).get_len());
    ms.add_u(pos, "first_nz_indices", id, std::move(v));
    record(rec, "get_begin_nzs_of_level_after_merge_path", id, ",work_size:" + std::to_string(work));
}

// ------------------------------------------------------------------- col / nnz direction: a child level relative to its parent
// get_begin_rows_of_BMW_after_fixed_blocking_in_col_direction_relative_to_BMTB.cc: the BMT rule above one level up
void get_begin_rows_of_BMW_after_fixed_blocking_in_col_direction_relative_to_BMTB(meta_data_set &ms, int id, uint64_t c, recs *rec)
{
    std::vector<uint64_t> len = row_nnz(ms, id);
    const universal_array &P = ms.get(TBLOCK_META, "first_row_indices", id);
    std::vector<uint64_t> v;
    for (uint64_t b = 0; b + 1 < P.get_len(); b++)
        for (uint64_t r = P.u[b]; r < P.u[b + 1]; r++)
            for (uint64_t k = 0; k < (len.at(r) + c - 1) / c; k++) v.push_back(r - P.u[b]);
    GS_REQUIRE(!v.empty(), "no BMW");
    ms.add_u(WARP_META, "first_row_indices_relative_to_BMTB", id, std::move(v));
    record(rec, "get_begin_rows_of_BMW_after_fixed_blocking_in_col_direction_relative_to_BMTB", id, ",col_size:" + std::to_string(c));
}

// get_begin_nzs_of_BMW_after_fixed_blocking_in_col_direction_relative_to_BMTB.cc: the offset of every BMW inside its BMTB, no sentinel
void get_begin_nzs_of_BMW_after_fixed_blocking_in_col_direction_relative_to_BMTB(meta_data_set &ms, int id, uint64_t c, recs *rec)
{
    std::vector<uint64_t> len = row_nnz(ms, id);
    const universal_array &P = ms.get(TBLOCK_META, "first_row_indices", id);
    std::vector<uint64_t> v;
    for (uint64_t b = 0; b + 1 < P.get_len(); b++) {
        v.push_back(0);
        for (uint64_t r = P.u[b]; r < P.u[b + 1]; r++)
            for (uint64_t left = len.at(r); left > 0;) {
                uint64_t take = std::min(left, c);
                v.push_back(v.back() + take);
                left -= take;
            }
        v.pop_back(); // the parent's closing offset is dropped again: one entry per BMW
    }
    ms.add_u(WARP_META, "first_nz_indices_relative_to_BMTB", id, std::move(v));
    record(rec, "get_begin_nzs_of_BMW_after_fixed_blocking_in_col_direction_relative_to_BMTB", id, ",col_size:" + std::to_string(c));
}

// get_begin_{rows,nzs}_of_BMT_after_fixed_blocking_in_nnz_direction_relative_to_{BMW,BMTB}.cc: as the BMW-to-BMTB pair above
void get_begin_rows_of_BMT_after_fixed_blocking_in_nnz_direction_relative_to_parent(meta_data_set &ms, POS_TYPE parent, int id, uint64_t k, recs *rec)
{
    GS_REQUIRE(k >= 1 && (parent == WARP_META || parent == TBLOCK_META), "bad argument");
    const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    const universal_array &PR = ms.get(parent, "first_row_indices", id), &PN = ms.get(parent, "first_nz_indices", id);
    std::vector<uint64_t> v;
    uint64_t p = 0;
    for (uint64_t i = 0; i < R.get_len(); i += k) {
        GS_REQUIRE(p + 1 < PN.get_len(), "BMT beyond the last parent block");
        if (i >= PN.u[p + 1]) p++;
        GS_REQUIRE(p < PR.get_len() && R.u[i] >= PR.u[p], "BMT starts above its parent's first row");
        v.push_back(R.u[i] - PR.u[p]);
    }
    const char *tag = level_tag(parent);
    ms.add_u(THREAD_META, std::string("first_row_indices_relative_to_") + tag, id, std::move(v));
    record(rec, std::string("get_begin_rows_of_BMT_after_fixed_blocking_in_nnz_direction_relative_to_") + tag, id, ",nnz_per_BMT:" + std::to_string(k));
}

void get_begin_nzs_of_BMT_after_fixed_blocking_in_nnz_direction_relative_to_parent(meta_data_set &ms, POS_TYPE parent, int id, uint64_t k, recs *rec)
{
    GS_REQUIRE(k >= 1 && (parent == WARP_META || parent == TBLOCK_META), "bad argument");
    const uint64_t nnz = ms.get(GLOBAL_META, "nz_row_indices", id).get_len();
    const universal_array &PN = ms.get(parent, "first_nz_indices", id);
    std::vector<uint64_t> v;
    uint64_t p = 0;
    for (uint64_t i = 0; i < nnz; i += k) {
        GS_REQUIRE(p + 1 < PN.get_len(), "BMT beyond the last parent block");
        if (i >= PN.u[p + 1]) p++;
        GS_REQUIRE(i >= PN.u[p], "BMT starts above its parent's first non-zero");
        v.push_back(i - PN.u[p]);
    }
    const char *tag = level_tag(parent);
    ms.add_u(THREAD_META, std::string("first_nz_indices_relative_to_") + tag, id, std::move(v));
    record(rec, std::string("get_begin_nzs_of_BMT_after_fixed_blocking_in_nnz_direction_relative_to_") + tag, id, ",nnz_per_BMT:" + std::to_string(k));
}

// get_BMW_size_of_each_parent.cc: one size per parent block (GLOBAL: one in all) when the BMWs of that parent hold the same nnz;
// the first parent whose BMWs differ ends the step with no output and no record
void get_BMW_size_of_each_parent(meta_data_set &ms, POS_TYPE parent, int id, recs *rec)
{
    if (parent == GLOBAL_META) return get_BMW_size_of_global(ms, id, rec);
    GS_REQUIRE(parent == TBLOCK_META, "parent_pos must be GLOBAL_META or TBLOCK_META");
    const universal_array &W = ms.get(WARP_META, "first_nz_indices", id), &P = ms.get(TBLOCK_META, "first_nz_indices", id);
    GS_REQUIRE(W.get_len() >= 2 && W.u[0] == P.u[0] && W.u.back() == P.u.back(), "BMW and BMTB offsets do not span the same non-zeros");
    std::vector<uint64_t> out;
    uint64_t w = 0;
    for (uint64_t b = 0; b + 1 < P.get_len(); b++) {
        GS_REQUIRE(w < W.get_len() && W.u[w] == P.u[b], "BMW does not start at its BMTB's first non-zero");
        uint64_t size = 0;
        bool seen = false;
        while (w + 1 < W.get_len() && W.u[w] < P.u[b + 1]) {
            GS_REQUIRE(W.u[w + 1] > W.u[w], "empty BMW");
            const uint64_t sz = W.u[w + 1] - W.u[w];
            if (!seen) size = sz, seen = true;
            else if (sz != size) return;
            w++;
        }
        out.push_back(size);
    }
    GS_REQUIRE(w == W.get_len() - 1, "BMW offsets run past the last BMTB");
    ms.add_u(TBLOCK_META, "BMW_size_of_each_blk", id, std::move(out));
    record(rec, "get_BMW_size_of_each_parent", id, ",parent_pos:TBLOCK_META");
}

// ---------------------------------------------------------------------- balanced (nnz) blocking inside a parent level
// data_transform_common.cc:794-931: the greedy >= rule restarts in every parent block and never cuts behind a parent's last row
static void balanced_in_parent(const universal_array &PR, const universal_array *PN, const std::vector<uint64_t> &len, uint64_t per, bool relative, bool nzs, uint64_t row_num,
                               std::vector<uint64_t> &out)
{
    for (uint64_t j = 0; j + 1 < PR.get_len(); j++) {
        const uint64_t r0 = PR.u[j], r1 = PR.u[j + 1], nz0 = PN ? PN->u[j] : 0;
        out.push_back(nzs ? (relative ? 0 : nz0) : (relative ? 0 : r0));
        uint32_t count = 0; // `unsigned int nz_count` in the reference
        uint64_t in_block = 0;
        for (uint64_t i = r0; i < r1; i++) {
            count += (uint32_t)len.at(i), in_block += len[i];
            if (count >= per && i != r1 - 1) {
                out.push_back(nzs ? (relative ? in_block : nz0 + in_block) : (relative ? i + 1 - r0 : i + 1));
                count = 0;
            }
        }
    }
    if (!relative) out.push_back(nzs ? PN->u.back() : row_num);
}

void get_begin_rows_after_nnz_blocking_in_row_direction_in_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t per, bool relative, recs *rec)
{
    GS_REQUIRE(per >= 1, "nnz_per_interval must be >= 1");
    uint64_t row_num;
    std::vector<uint64_t> len = row_nnz(ms, id, &row_num), v;
    balanced_in_parent(ms.get(parent, "first_row_indices", id), nullptr, len, per, relative, false, row_num, v);
    ms.add_u(pos, relative ? std::string("first_row_indices_relative_to_") + level_tag(parent) : std::string("first_row_indices"), id, std::move(v));
    record(rec, std::string("get_begin_rows_of_") + level_tag(pos) + "_after_nnz_blocking_in_row_direction_" + (relative ? "relative_to_" : "in_") + level_tag(parent), id,
           ",nnz_per_interval:" + std::to_string(per));
}

void get_begin_nzs_after_nnz_blocking_in_row_direction_in_parent(meta_data_set &ms, POS_TYPE pos, POS_TYPE parent, int id, uint64_t per, bool relative, recs *rec)
{
    GS_REQUIRE(per >= 1, "nnz_per_interval must be >= 1");
    uint64_t row_num;
    std::vector<uint64_t> len = row_nnz(ms, id, &row_num), v;
    const universal_array &PR = ms.get(parent, "first_row_indices", id), &PN = ms.get(parent, "first_nz_indices", id);
    GS_REQUIRE(PR.get_len() == PN.get_len(), "parent row and nz offsets differ in length");
    balanced_in_parent(PR, &PN, len, per, relative, true, row_num, v);
    ms.add_u(pos, relative ? std::string("first_nz_indices_relative_to_") + level_tag(parent) : std::string("first_nz_indices"), id, std::move(v));
    record(rec, std::string("get_begin_nzs_of_") + level_tag(pos) + "_after_nnz_blocking_in_row_direction_" + (relative ? "relative_to_" : "in_") + level_tag(parent), id,
           ",nnz_per_interval:" + std::to_string(per));
}

// ----------------------------------------------------------------------------------------------- interleaved storage
// modify_{col_indices,vals,row_indices}_by_interlance_storage.cc: inside every parent block (GLOBAL: the whole stream) the BMTs, all of
// one size, are stored element-interleaved: new[bmt + j * n_bmt + base] = old[j + bmt * size + base]; the originals stay
void modify_by_interlance_storage(meta_data_set &ms, POS_TYPE parent_pos, int id, recs *rec)
{
    GS_REQUIRE(parent_pos == GLOBAL_META || parent_pos == TBLOCK_META || parent_pos == WARP_META, "parent_pos");
    const universal_array &C = ms.get(GLOBAL_META, "nz_col_indices", id), &V = ms.get(GLOBAL_META, "nz_vals", id), &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    const universal_array &S = ms.get(parent_pos, "BMT_size_of_each_blk", id);
    const uint64_t n = C.get_len();
    std::vector<uint64_t> perm(n); // perm[new] = old
    for (uint64_t i = 0; i < n; i++) perm[i] = i; // positions no block covers keep their element (the reference leaves zeros there: none exist for valid input)
    auto block = [&](uint64_t base, uint64_t n_bmt, uint64_t size) {
        GS_REQUIRE(base + n_bmt * size <= n, "interleave: BMTs run past the stream");
        for (uint64_t b = 0; b < n_bmt; b++)
            for (uint64_t j = 0; j < size; j++) perm[b + j * n_bmt + base] = j + b * size + base;
    };
    if (parent_pos == GLOBAL_META) {
        const uint64_t size = S.u.at(0);
        GS_REQUIRE(size > 0 && n % size == 0, "interleave: the stream is not a whole number of BMTs");
        block(0, n / size, size);
    } else {
        const universal_array &FB = ms.get(parent_pos, "first_BMT_indices", id);
        GS_REQUIRE(S.get_len() + 1 == FB.get_len(), "interleave: one BMT size per parent block expected");
        uint64_t base = 0;
        for (uint64_t p = 0; p + 1 < FB.get_len(); p++) {
            const uint64_t n_bmt = FB.u[p + 1] - FB.u[p], size = S.u[p];
            block(base, n_bmt, size);
            base += n_bmt * size;
        }
        GS_REQUIRE(base == n, "interleave: the parent blocks do not cover the stream");
    }
    std::vector<uint64_t> nc(n), nr(n);
    std::vector<double> nv(n);
    for (uint64_t i = 0; i < n; i++) nc[i] = C.u[perm[i]], nr[i] = R.u[perm[i]], nv[i] = V.f[perm[i]];
    const std::string pp = std::string(",parent_pos:") + pos_name(parent_pos);
    ms.add_u(GLOBAL_META, "nz_col_indices_after_interlance_storage", id, std::move(nc));
    record(rec, "modify_col_indices_by_interlance_storage", id, pp);
    ms.add(GLOBAL_META, "nz_vals_after_interlance_storage", id, vals_like(V, std::move(nv)));
    record(rec, "modify_vals_by_interlance_storage", id, pp);
    ms.add_u(GLOBAL_META, "nz_row_indices_after_interlance_storage", id, std::move(nr));
    record(rec, "modify_row_indices_by_interlance_storage", id, pp);
}

// ------------------------------------------------------------------------------------------------ sub-matrix division
// meta_data_set::get_max_sub_matrix_id_of_data_item (metadata_set.cc:363-424)
static int max_sub_matrix_id(const meta_data_set &ms, POS_TYPE pos, const std::string &name)
{
    const std::string prefix = std::string(pos_name(pos)) + "_" + name + "_";
    int mx = -2;
    for (const std::string &k : ms.keys()) {
        if (k.compare(0, prefix.size(), prefix) != 0) continue;
        const std::string rest = k.substr(prefix.size());
        if (rest.empty() || rest.find_first_not_of("-0123456789") != std::string::npos) continue;
        mx = std::max(mx, atoi(rest.c_str()));
    }
    GS_REQUIRE(mx != -2, "no item " + prefix + "*");
    return mx;
}

// the 7 steps both division operators run, in the reference's order: row start, row end, col start, col end boundaries of every new
// sub-matrix, then the column, value and row streams (each removes the parent's array).  `first_row[b]` = first row of part b (parts
// ascending), `rebase`: store row indices relative to the part (fixed division) or as they are (division by row nnz, as the reference does)
static void divide_rows(meta_data_set &ms, int id, const std::vector<uint64_t> &first_row, const std::vector<uint64_t> &last_row, const std::vector<bool> &has_boundary, bool rebase,
                        const char *const names[7], const std::string &param, recs *rec)
{
    const uint64_t begin_row = ms.scalar(GLOBAL_META, "begin_row_index", id), begin_col = ms.scalar(GLOBAL_META, "begin_col_index", id), end_col = ms.scalar(GLOBAL_META, "end_col_index", id);
    const std::vector<uint64_t> R = ms.get(GLOBAL_META, "nz_row_indices", id).u, C = ms.get(GLOBAL_META, "nz_col_indices", id).u;
    const universal_array V = ms.get(GLOBAL_META, "nz_vals", id);
    const size_t parts = first_row.size();
    // part of every non-zero (rows ascending)
    std::vector<std::vector<uint64_t>> pr(parts), pc(parts);
    std::vector<std::vector<double>> pv(parts);
    size_t b = 0;
    for (uint64_t i = 0; i < R.size(); i++) {
        while (b + 1 < parts && R[i] >= first_row[b + 1]) b++;
        GS_REQUIRE(R[i] >= first_row[b], "division: non-zero above its part");
        pr[b].push_back(rebase ? R[i] - first_row[b] : R[i]), pc[b].push_back(C[i]), pv[b].push_back(V.f[i]);
    }
    auto add_scalars = [&](const char *key, const char *step, auto value) {
        for (size_t q = 0; q < parts; q++)
            if (has_boundary[q]) ms.add_scalar(GLOBAL_META, key, max_sub_matrix_id(ms, GLOBAL_META, key) + 1, value(q));
        record(rec, step, id, param);
    };
    add_scalars("begin_row_index", names[0], [&](size_t q) { return begin_row + first_row[q]; });
    add_scalars("end_row_index", names[1], [&](size_t q) { return begin_row + last_row[q]; });
    add_scalars("begin_col_index", names[2], [&](size_t) { return begin_col; });
    add_scalars("end_col_index", names[3], [&](size_t) { return end_col; });
    for (size_t q = 0; q < parts; q++)
        if (!pc[q].empty()) ms.add_u(GLOBAL_META, "nz_col_indices", max_sub_matrix_id(ms, GLOBAL_META, "nz_col_indices") + 1, std::move(pc[q]));
    ms.remove(GLOBAL_META, "nz_col_indices", id);
    if (std::string(names[4]) == "fixed_div_col_indices_by_corr_row_indices") { // (sic) convert_to_string() of that step drops the "fixed_" before the "::"
        if (rec) rec->push_back(std::string("div_col_indices_by_corr_row_indices::{name:\"") + names[4] + "\",target_matrix_id:" + std::to_string(id) + param + "}");
    } else record(rec, names[4], id, param);
    for (size_t q = 0; q < parts; q++)
        if (!pv[q].empty()) ms.add(GLOBAL_META, "nz_vals", max_sub_matrix_id(ms, GLOBAL_META, "nz_vals") + 1, vals_like(V, std::move(pv[q])));
    ms.remove(GLOBAL_META, "nz_vals", id);
    record(rec, names[5], id, param);
    for (size_t q = 0; q < parts; q++)
        if (!pr[q].empty()) ms.add_u(GLOBAL_META, "nz_row_indices", max_sub_matrix_id(ms, GLOBAL_META, "nz_row_indices") + 1, std::move(pr[q]));
    ms.remove(GLOBAL_META, "nz_row_indices", id);
    record(rec, names[6], id, param);
    // the reference asserts that every kind of item ends at the same sub-matrix id (operator run(), check == true)
    const int m = max_sub_matrix_id(ms, GLOBAL_META, "end_row_index");
    GS_REQUIRE(m == max_sub_matrix_id(ms, GLOBAL_META, "begin_row_index") && m == max_sub_matrix_id(ms, GLOBAL_META, "nz_row_indices") && m == max_sub_matrix_id(ms, GLOBAL_META, "nz_col_indices") &&
                   m == max_sub_matrix_id(ms, GLOBAL_META, "nz_vals"),
               "division: a part without non-zeros got boundaries but no arrays");
}

// fixed_interval_row_matrix_div_operator's steps (modify_*_boundary_after_fixed_div_in_row_direction.cc, fixed_div_*.cc): part s = rows
// [s * gap, (s + 1) * gap); only parts that hold a non-zero become sub-matrices; row indices are stored modulo gap
void fixed_div_in_row_direction(meta_data_set &ms, int id, uint64_t gap, recs *rec)
{
    GS_REQUIRE(gap >= 1, "fixed_row_gap_size must be >= 1");
    const uint64_t begin_row = ms.scalar(GLOBAL_META, "begin_row_index", id), end_row = ms.scalar(GLOBAL_META, "end_row_index", id);
    const uint64_t row_num = end_row - begin_row + 1, n_sub = (row_num + gap - 1) / gap;
    std::vector<bool> used(n_sub, false);
    for (uint64_t r : ms.get(GLOBAL_META, "nz_row_indices", id).u) {
        GS_REQUIRE(r / gap < n_sub, "division: row index beyond end_row_index");
        used[r / gap] = true;
    }
    std::vector<uint64_t> first, last;
    for (uint64_t q = 0; q < n_sub; q++)
        if (used[q]) first.push_back(q * gap), last.push_back(std::min((q + 1) * gap - 1, end_row - begin_row));
    static const char *const names[7] = {"modify_row_start_boundary_after_fixed_div_in_row_direction", "modify_row_end_boundary_after_fixed_div_in_row_direction",
                                         "modify_col_start_boundary_after_fixed_div_in_row_direction", "modify_col_end_boundary_after_fixed_div_in_row_direction",
                                         "fixed_div_col_indices_by_corr_row_indices", "fixed_div_vals_by_corr_row_indices", "fixed_div_row_indices"};
    divide_rows(ms, id, first, last, std::vector<bool>(first.size(), true), true, names, ",fixed_row_gap_size:" + std::to_string(gap), rec);
}

// row_nz_matrix_div_operator's steps (modify_*_boundary_after_div_according_to_row_nz.cc, div_*_by_row_nnz.cc): a new part starts at
// every row whose nnz leaves the geometric window [low, low * expansion) the part opened with; windows start at [0, nz_gap) and stop
// growing past max_gap.  Row indices keep their value (the reference does not rebase them here).
std::vector<uint64_t> div_positions_by_row_nnz(const std::vector<uint64_t> &len, uint64_t nz_gap, uint64_t max_gap, uint64_t expansion)
{
    std::vector<uint64_t> pos;
    if (len.empty()) return pos;
    uint64_t lo = 0, hi = nz_gap;
    auto window = [&](uint64_t nz) {
        lo = 0, hi = nz_gap;
        if (nz < max_gap) {
            while (nz >= hi && hi <= max_gap) lo = hi, hi = hi * expansion;
        } else lo = max_gap, hi = lo * expansion;
    };
    pos.push_back(0);
    window(len[0]);
    for (uint64_t r = 1; r < len.size(); r++) {
        const uint64_t nz = len[r];
        if ((nz >= lo && nz < hi) || (nz >= hi && hi > max_gap)) continue;
        pos.push_back(r);
        window(nz);
    }
    return pos;
}

void div_according_to_row_nz(meta_data_set &ms, int id, uint64_t nz_gap, uint64_t max_gap, uint64_t expansion, recs *rec)
{
    GS_REQUIRE(nz_gap >= 1 && expansion >= 1, "bad division parameters");
    const uint64_t begin_row = ms.scalar(GLOBAL_META, "begin_row_index", id), end_row = ms.scalar(GLOBAL_META, "end_row_index", id);
    const universal_array &R = ms.get(GLOBAL_META, "nz_row_indices", id);
    const uint64_t row_num = end_row - begin_row + 1;
    std::vector<uint64_t> len = get_nnz_of_each_row_in_spec_range(R, 0, row_num - 1, 0, R.get_len() - 1);
    std::vector<uint64_t> first = div_positions_by_row_nnz(len, nz_gap, max_gap, expansion), last;
    for (size_t q = 0; q < first.size(); q++) last.push_back((q + 1 < first.size() ? first[q + 1] : len.size()) - 1);
    static const char *const names[7] = {"modify_row_start_boundary_after_div_according_to_row_nz", "modify_row_end_boundary_after_div_according_to_row_nz",
                                         "modify_col_start_boundary_after_div_according_to_row_nz", "modify_col_end_boundary_after_div_according_to_row_nz",
                                         "div_col_indices_by_row_nnz", "div_val_indices_by_row_nnz", "div_row_indices_by_row_nnz"};
    // begin_row_index of a part is div_position itself, not begin_row + div_position (modify_row_start_boundary_after_div_according_to_row_nz.cc)
    const uint64_t saved = begin_row;
    if (saved != 0) ms.replace_u(GLOBAL_META, "begin_row_index", id, std::vector<uint64_t>{0});
    try {
        divide_rows(ms, id, first, last, std::vector<bool>(first.size(), true), false, names, ",nz_gap_size:" + std::to_string(nz_gap), rec);
    } catch (...) {
        if (saved != 0) ms.replace_u(GLOBAL_META, "begin_row_index", id, std::vector<uint64_t>{saved});
        throw;
    }
    if (saved != 0) ms.replace_u(GLOBAL_META, "begin_row_index", id, std::vector<uint64_t>{saved});
}

void remove_item_of_metadata(meta_data_set &ms, int id, const std::string &name, POS_TYPE pos, recs *rec)
{
    ms.remove(pos, name, id);
    record(rec, "remove_item_of_metadata", id, std::string(", pos:") + pos_name(pos));
}

const std::vector<std::string> &covered_reference_step_names()
{
    static const std::vector<std::string> names = {
        "get_row_order_by_length", "reorder_val_by_index", "reorder_col_by_index", "reorder_row_by_index",
        "remove_empty_row_in_end_of_sub_matrix",
        "modify_col_indices_by_col_pad_in_sub_matrix", "modify_vals_by_col_pad_in_sub_matrix", "modify_row_indices_by_col_pad_in_sub_matrix",
        "modify_col_indices_by_nnz_pad", "modify_vals_by_nnz_pad", "modify_row_indices_by_nnz_pad",
        "modify_col_indices_by_row_pad_in_sub_matrix", "modify_vals_by_row_pad_in_sub_matrix", "modify_row_indices_by_row_pad_in_sub_matrix",
        "modify_col_indices_by_empty_pad_in_submatrix", "modify_vals_by_empty_pad_in_submatrix", "modify_row_indices_by_empty_pad_in_submatrix",
        "get_begin_rows_of_BMTBs_after_fixed_blocking_in_row_direction", "get_begin_nzs_of_BMTBs_after_fixed_blocking_in_row_direction",
        "get_begin_rows_of_BMW_after_fixed_blocking_in_row_direction_without_BMTB", "get_begin_nzs_of_BMW_after_fixed_blocking_in_row_direction_without_BMTB",
        "get_begin_rows_of_BMW_after_fixed_blocking_in_row_direction_in_BMTB", "get_begin_nzs_of_BMW_after_fixed_blocking_in_row_direction_in_BMTB",
        "get_begin_rows_of_BMW_after_fixed_blocking_in_row_direction_relative_to_BMTB", "get_begin_nzs_of_BMW_after_fixed_blocking_in_row_direction_relative_to_BMTB",
        "get_begin_rows_of_BMT_after_fixed_blocking_in_row_direction", "get_begin_nzs_of_BMT_after_fixed_blocking_in_row_direction",
        "get_begin_rows_of_BMT_after_fixed_blocking_in_row_direction_in_BMTB", "get_begin_nzs_of_BMT_after_fixed_blocking_in_row_direction_in_BMTB",
        "get_begin_rows_of_BMT_after_fixed_blocking_in_row_direction_in_BMW", "get_begin_nzs_of_BMT_after_fixed_blocking_in_row_direction_in_BMW",
        "get_begin_rows_of_BMT_after_fixed_blocking_in_row_direction_relative_to_BMTB", "get_begin_nzs_of_BMT_after_fixed_blocking_in_row_direction_relative_to_BMTB",
        "get_begin_rows_of_BMT_after_fixed_blocking_in_row_direction_relative_to_BMW", "get_begin_nzs_of_BMT_after_fixed_blocking_in_row_direction_relative_to_BMW",
        "get_begin_BMWs_of_BMTB_after_blocking_in_row_direction", "get_begin_BMTs_of_specific_parent_after_blocking_in_row_direction",
        "get_begin_BMTs_of_specific_parent_after_blocking", "get_begin_BMWs_of_BMTB_after_blocking",
        "get_begin_rows_of_BMT_after_fixed_blocking_in_col_direction", "get_begin_nzs_of_BMT_after_fixed_blocking_in_col_direction",
        "get_begin_rows_of_BMW_after_fixed_blocking_in_col_direction", "get_begin_nzs_of_BMW_after_fixed_blocking_in_col_direction",
        "get_begin_rows_of_BMTB_after_fixed_blocking_in_col_direction", "get_begin_nzs_of_BMTB_after_fixed_blocking_in_col_direction",
        "get_begin_rows_of_BMT_after_fixed_blocking_in_nnz_direction", "get_begin_nzs_of_BMT_after_fixed_blocking_in_nnz_direction",
        "get_begin_rows_of_BMW_after_fixed_blocking_in_nnz_direction", "get_begin_nzs_of_BMW_after_fixed_blocking_in_nnz_direction",
        "get_begin_rows_of_BMTB_after_fixed_blocking_in_nnz_direction", "get_begin_nzs_of_BMTB_after_fixed_blocking_in_nnz_direction",
        "get_begin_rows_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB",
        "get_begin_nzs_of_BMW_after_fixed_blocking_in_nnz_direction_relative_to_BMTB",
        "get_BMT_size_of_each_parent", "get_BMW_size_of_each_parent", "get_BMTB_size",
        "thread_bit_map", "parent_bit_map_of_thread", "segment_empty_flag", "segment_empty_row_indices", "segment_offset", "segment_ptr",
        "get_begin_rows_after_merge_thread", "get_begin_nzs_after_merge_thread", "get_begin_rows_relative_to_parent_after_merge_thread",
        "get_begin_nzs_relative_to_parent_after_merge_thread", "get_begin_BMTs_after_merge_thread",
        "get_begin_rows_of_BMT_after_nnz_blocking_in_row_direction", "get_begin_nzs_of_BMT_after_nnz_blocking_in_row_direction",
        "get_begin_rows_of_BMW_after_nnz_blocking_in_row_direction", "get_begin_nzs_of_BMW_after_nnz_blocking_in_row_direction",
        "get_begin_rows_of_BMTB_after_nnz_blocking_in_row_direction", "get_begin_nzs_of_BMTB_after_nnz_blocking_in_row_direction",
        "get_begin_rows_of_BMT_after_fixed_blocking_in_col_direction_relative_to_BMTB", "get_begin_rows_of_BMT_after_fixed_blocking_in_col_direction_relative_to_BMW",
        "get_begin_nzs_of_BMT_after_fixed_blocking_in_col_direction_relative_to_parents",
        "modify_col_indices_by_col_pad_parent_blk_to_max_row_size", "modify_vals_by_col_pad_parent_blk_to_max_row_size",
        "modify_row_indices_by_col_pad_parent_blk_to_max_row_size",
        "get_begin_rows_of_level_after_merge_path", "get_begin_nzs_of_level_after_merge_path",
        "remove_item_of_metadata"};
    return names;
}

} // namespace steps
} // namespace gs
