#include "metadata.hpp"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <sys/stat.h>

namespace gs {

const char *pos_name(POS_TYPE p)
{
    switch (p) {
    case GLOBAL_META: return "GLOBAL_META";
    case TBLOCK_META: return "TBLOCK_META";
    case WARP_META: return "WARP_META";
    case THREAD_META: return "THREAD_META";
    default: return "NONE_META";
    }
}

const char *data_type_name(data_type t)
{
    switch (t) {
    case UNSIGNED_CHAR: return "UNSIGNED_CHAR";
    case UNSIGNED_SHORT: return "UNSIGNED_SHORT";
    case UNSIGNED_INT: return "UNSIGNED_INT";
    case UNSIGNED_LONG: return "UNSIGNED_LONG";
    case FLOAT: return "FLOAT";
    default: return "DOUBLE";
    }
}

size_t data_type_size(data_type t)
{
    switch (t) {
    case UNSIGNED_CHAR: return 1;
    case UNSIGNED_SHORT: return 2;
    case UNSIGNED_INT: return 4;
    case UNSIGNED_LONG: return 8;
    case FLOAT: return 4;
    default: return 8;
    }
}

data_type find_most_suitable_data_type(uint64_t m)
{
    if (m <= 255ull) return UNSIGNED_CHAR;
    if (m <= 65535ull) return UNSIGNED_SHORT;
    if (m <= 4294967295ull) return UNSIGNED_INT;
    return UNSIGNED_LONG;
}

universal_array::universal_array(std::vector<double> v, data_type t) : type(t), f(std::move(v))
{
    GS_REQUIRE(t == FLOAT || t == DOUBLE, "value arrays are FLOAT or DOUBLE");
    if (t == FLOAT) // compress_float_precise: values live at f32 precision
        for (double &x : f) x = (double)(float)x;
}

uint64_t universal_array::max_value() const
{
    uint64_t m = 0;
    for (uint64_t x : u) m = x > m ? x : m;
    return m;
}

data_type universal_array::get_compress_data_type(bool data_type_compress) const
{
    if (is_float()) return type;
    if (!data_type_compress) return type;
    return find_most_suitable_data_type(max_value());
}

std::string meta_data_set::item_name(POS_TYPE pos, const std::string &name, int id)
{
    return std::string(pos_name(pos)) + "_" + name + "_" + std::to_string(id);
}

bool meta_data_set::is_exist(POS_TYPE pos, const std::string &name, int id) const
{
    return items_.count(item_name(pos, name, id)) != 0;
}

universal_array &meta_data_set::get(POS_TYPE pos, const std::string &name, int id)
{
    auto it = items_.find(item_name(pos, name, id));
    GS_REQUIRE(it != items_.end(), "no metadata item " + item_name(pos, name, id));
    return it->second.arr;
}

const universal_array &meta_data_set::get(POS_TYPE pos, const std::string &name, int id) const
{
    auto it = items_.find(item_name(pos, name, id));
    GS_REQUIRE(it != items_.end(), "no metadata item " + item_name(pos, name, id));
    return it->second.arr;
}

const universal_array *meta_data_set::find_key(const std::string &key) const
{
    auto it = items_.find(key);
    return it == items_.end() ? nullptr : &it->second.arr;
}

void meta_data_set::add(POS_TYPE pos, const std::string &name, int id, universal_array arr)
{
    std::string key = item_name(pos, name, id);
    GS_REQUIRE(items_.count(key) == 0, "metadata item already exists: " + key);
    GS_REQUIRE(arr.get_len() > 0, "empty metadata array: " + key);
    items_.emplace(key, entry{pos, name, id, std::move(arr)});
    order_.push_back(key);
}

void meta_data_set::remove(POS_TYPE pos, const std::string &name, int id)
{
    std::string key = item_name(pos, name, id);
    auto it = items_.find(key);
    GS_REQUIRE(it != items_.end(), "cannot remove missing item " + key);
    items_.erase(it);
    order_.erase(std::find(order_.begin(), order_.end(), key));
}

void meta_data_set::replace_u(POS_TYPE pos, const std::string &name, int id, std::vector<uint64_t> v)
{
    remove(pos, name, id);
    add_u(pos, name, id, std::move(v));
}

void meta_data_set::replace(POS_TYPE pos, const std::string &name, int id, universal_array arr)
{
    remove(pos, name, id);
    add(pos, name, id, std::move(arr));
}

std::vector<std::string> meta_data_set::all_item_of_metadata_of_diff_pos(POS_TYPE pos, int id) const
{
    std::vector<std::string> out;
    for (const auto &k : order_) {
        const entry &e = items_.at(k);
        if (e.pos == pos && e.id == id) out.push_back(e.name);
    }
    return out;
}

std::vector<std::string> meta_data_set::keys() const { return order_; }

bool meta_data_set::check() const
{
    for (const auto &kv : items_) {
        const entry &e = kv.second;
        if (kv.first != item_name(e.pos, e.name, e.id)) return false;
        if (e.arr.get_len() == 0) return false;
        if (e.arr.is_float() && !e.arr.u.empty()) return false;
        if (!e.arr.is_float() && !e.arr.f.empty()) return false;
    }
    return true;
}

void meta_data_set::output_item_to_file(const std::string &key, const std::string &path) const
{
    const universal_array *a = find_key(key);
    GS_REQUIRE(a != nullptr, "no metadata item " + key);
    std::ofstream out(path, std::ios::out | std::ios::trunc);
    GS_REQUIRE((bool)out, "cannot open " + path);
    if (a->is_float()) {
        for (double v : a->f) {
            // the reference clamps everything <= 1e-10 to 0 when it writes (struct.cc:2018-2022)
            if (v <= 0.0000000001) v = 0.0;
            out << v << '\n';
        }
    } else {
        for (uint64_t v : a->u) out << v << '\n';
    }
}

void meta_data_set::output_2_dir(const std::string &dir) const
{
    mkdir(dir.c_str(), 0755);
    for (const auto &k : order_) output_item_to_file(k, dir + "/" + k);
}

uint64_t meta_data_set::real_row_num(int id) const
{
    uint64_t begin = scalar(GLOBAL_META, "begin_row_index", id);
    uint64_t end = scalar(GLOBAL_META, "end_row_index", id);
    GS_REQUIRE(end >= begin, "end_row_index < begin_row_index");
    const universal_array &r = get(GLOBAL_META, "nz_row_indices", id);
    uint64_t real_end = begin + r.u.back();
    if (end < real_end) end = real_end;
    return end - begin + 1;
}

std::vector<uint64_t> get_nnz_of_each_row_in_spec_range(const universal_array &row_index, uint64_t begin_row_bound,
                                                        uint64_t end_row_bound, uint64_t begin_nz_bound, uint64_t end_nz_bound)
{
    GS_REQUIRE(end_row_bound >= begin_row_bound, "row bounds");
    GS_REQUIRE(end_nz_bound < row_index.u.size() && begin_nz_bound <= end_nz_bound, "nz bounds");
    std::vector<uint64_t> n(end_row_bound - begin_row_bound + 1, 0);
    uint64_t last = 0;
    for (uint64_t i = begin_nz_bound; i <= end_nz_bound; i++) {
        uint64_t r = row_index.u[i];
        GS_REQUIRE(r >= last, "row indices must be non-decreasing");
        GS_REQUIRE(r >= begin_row_bound && r <= end_row_bound, "row index outside the stated range");
        n[r - begin_row_bound]++;
        last = r;
    }
    return n;
}

static void init_set(meta_data_set &ms, uint64_t max_row_index, uint64_t max_col_index, std::vector<uint64_t> rows,
                     std::vector<uint64_t> cols, std::vector<double> vals, data_type vt)
{
    GS_REQUIRE(!rows.empty(), "the matrix has no non-zeros (the reference path cannot represent it)");
    uint64_t nnz = rows.size();
    ms.add_scalar(GLOBAL_META, "origin_row_num", -1, max_row_index + 1);
    ms.add_scalar(GLOBAL_META, "origin_col_num", -1, max_col_index + 1);
    ms.add_scalar(GLOBAL_META, "origin_nnz_num", -1, nnz);
    ms.add_scalar(GLOBAL_META, "begin_row_index", 0, 0);
    ms.add_scalar(GLOBAL_META, "begin_col_index", 0, 0);
    ms.add_scalar(GLOBAL_META, "end_row_index", 0, max_row_index);
    ms.add_scalar(GLOBAL_META, "end_col_index", 0, max_col_index);
    ms.add_u(GLOBAL_META, "nz_row_indices", 0, std::move(rows));
    ms.add_u(GLOBAL_META, "nz_col_indices", 0, std::move(cols));
    ms.add(GLOBAL_META, "nz_vals", 0, universal_array(std::move(vals), vt));
}

static data_type parse_precise(const std::string &p)
{
    if (p == "float") return FLOAT;
    if (p == "double") return DOUBLE;
    throw Error("create_init_metadata_set: invalid precise, float_precise: " + p);
}

std::shared_ptr<meta_data_set> create_init_metadata_set_from_file(const std::string &file_name, const std::string &matrix_name,
                                                                 const std::string &float_precise, bool reference_ones)
{
    data_type vt = parse_precise(float_precise);
    std::ifstream in(file_name);
    GS_REQUIRE(in.is_open(), "cannot open file " + file_name);
    std::vector<uint64_t> rows, cols;
    std::vector<double> vals;
    uint64_t max_row = 0, max_col = 0;
    bool first = true;
    std::string line;
    while (std::getline(in, line)) {
        // the reference skips empty lines, lines that start with white space, and '%' comments (struct.cc:97-100)
        if (line.empty() || isspace((unsigned char)line[0]) || line[0] == '%') continue;
        const char *p = line.c_str();
        char *e;
        long a = strtol(p, &e, 10);
        long b = strtol(e, &e, 10);
        if (first) {
            max_row = (uint64_t)(a - 1);
            max_col = (uint64_t)(b - 1);
            first = false;
            continue;
        }
        double v = strtod(e, &e);
        uint64_t r = (uint64_t)(a - 1), c = (uint64_t)(b - 1);
        GS_REQUIRE(rows.empty() || r >= rows.back(), "input must be row-major sorted COO (struct.cc:125)");
        rows.push_back(r);
        cols.push_back(c);
        vals.push_back(reference_ones ? 1.0 : v);
        if (r > max_row) max_row = r;
        if (c > max_col) max_col = c;
    }
    auto ms = std::make_shared<meta_data_set>();
    ms->matrix_name = matrix_name;
    init_set(*ms, max_row, max_col, std::move(rows), std::move(cols), std::move(vals), vt);
    return ms;
}

std::shared_ptr<meta_data_set> create_init_metadata_set_from_coo(uint64_t nrows, uint64_t ncols, uint64_t nnz, const int64_t *row,
                                                                const int64_t *col, const float *vals, const std::string &float_precise)
{
    data_type vt = parse_precise(float_precise);
    GS_REQUIRE(nrows > 0 && ncols > 0, "empty shape");
    std::vector<uint64_t> rows(nnz), cols(nnz);
    std::vector<double> v(nnz);
    uint64_t max_row = nrows - 1, max_col = ncols - 1;
    for (uint64_t i = 0; i < nnz; i++) {
        GS_REQUIRE(row[i] >= 0 && col[i] >= 0, "negative index");
        GS_REQUIRE(i == 0 || row[i] >= row[i - 1], "input must be row-major sorted COO (struct.cc:125)");
        rows[i] = (uint64_t)row[i];
        cols[i] = (uint64_t)col[i];
        v[i] = vals ? (double)vals[i] : 1.0;
        if (rows[i] > max_row) max_row = rows[i];
        if (cols[i] > max_col) max_col = cols[i];
    }
    auto ms = std::make_shared<meta_data_set>();
    init_set(*ms, max_row, max_col, std::move(rows), std::move(cols), std::move(v), vt);
    return ms;
}

} // namespace gs
