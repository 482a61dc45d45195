// Metadata store of the SpMM path: the B200-native counterpart of the reference's
// meta_data_set / meta_data_item / universal_array (metadata_set.hpp:103-153,
// code_source_data.hpp:11-76).  Same key names ("<POS>_<name>_<sub_matrix_id>",
// metadata_set.cc:147-151), same in-memory convention (every index array is u64, values are
// f32 or f64), same dtype-compression rule (code_source_data.cc:383-416 +
// op_manager.cc:985-1016) and the same one-value-per-line text dump (struct.cc:1991).
// Unlike the reference, arrays are flat std::vector and are read without a per-element
// dtype switch; failures throw gs::Error instead of assert(false).
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace gs {

struct Error : std::runtime_error {
    using std::runtime_error::runtime_error;
};
#define GS_REQUIRE(cond, msg)                                                      \
    do {                                                                           \
        if (!(cond)) throw ::gs::Error(std::string(__func__) + ": " + (msg));      \
    } while (0)

// POS_TYPE of the reference (metadata_set.hpp:12-22); only the four levels the path uses.
enum POS_TYPE { GLOBAL_META = 0, TBLOCK_META = 1, WARP_META = 2, THREAD_META = 3, NONE_META = 4 };
const char *pos_name(POS_TYPE p);

// data_type subset the path produces (struct.hpp data_type enum): what a dump/upload uses.
enum data_type { UNSIGNED_CHAR = 0, UNSIGNED_SHORT = 1, UNSIGNED_INT = 2, UNSIGNED_LONG = 3, FLOAT = 4, DOUBLE = 5 };
const char *data_type_name(data_type t);
size_t data_type_size(data_type t);
data_type find_most_suitable_data_type(uint64_t max_index_number); // op_manager.cc:985

// universal_array: an owning typed buffer.  Integers live as u64, values as f64 (rounded to
// f32 on store when the logical type is FLOAT, universal_array::compress_float_precise).
struct universal_array {
    data_type type = UNSIGNED_LONG; // logical in-memory type: UNSIGNED_LONG, FLOAT or DOUBLE
    std::vector<uint64_t> u;
    std::vector<double> f;

    universal_array() = default;
    explicit universal_array(std::vector<uint64_t> v) : type(UNSIGNED_LONG), u(std::move(v)) {}
    universal_array(std::vector<double> v, data_type t);
    bool is_float() const { return type == FLOAT || type == DOUBLE; }
    uint64_t get_len() const { return is_float() ? f.size() : u.size(); }
    uint64_t read_integer_from_arr(uint64_t i) const { return u[i]; }
    double read_float_from_arr(uint64_t i) const { return f[i]; }
    uint64_t max_value() const;
    data_type get_compress_data_type(bool data_type_compress = true) const;
};

class meta_data_set {
  public:
    std::string matrix_name;

    static std::string item_name(POS_TYPE pos, const std::string &name, int sub_matrix_id);
    bool is_exist(POS_TYPE pos, const std::string &name, int id) const;
    universal_array &get(POS_TYPE pos, const std::string &name, int id);
    const universal_array &get(POS_TYPE pos, const std::string &name, int id) const;
    const universal_array *find_key(const std::string &key) const;
    uint64_t scalar(POS_TYPE pos, const std::string &name, int id) const { return get(pos, name, id).u.at(0); }
    // add_element: the key must not exist (metadata_set.cc add_element asserts the same)
    void add(POS_TYPE pos, const std::string &name, int id, universal_array arr);
    void add_u(POS_TYPE pos, const std::string &name, int id, std::vector<uint64_t> v) { add(pos, name, id, universal_array(std::move(v))); }
    void add_scalar(POS_TYPE pos, const std::string &name, int id, uint64_t v) { add_u(pos, name, id, std::vector<uint64_t>{v}); }
    void remove(POS_TYPE pos, const std::string &name, int id);
    void replace_u(POS_TYPE pos, const std::string &name, int id, std::vector<uint64_t> v);
    void replace(POS_TYPE pos, const std::string &name, int id, universal_array arr);
    // names (without pos/id decoration) of every item at `pos` of sub-matrix `id`
    std::vector<std::string> all_item_of_metadata_of_diff_pos(POS_TYPE pos, int id) const;
    size_t count_of_metadata_of_diff_pos(POS_TYPE pos, int id) const { return all_item_of_metadata_of_diff_pos(pos, id).size(); }
    std::vector<std::string> keys() const;
    // invariants the reference asserts after every DISTRIBUTING operator (metadata_set.cc:334-361, :806+)
    bool check() const;
    // dump: one decimal value per line, ofstream default precision -- byte-identical to
    // meta_data_item::output_2_file / print_arr_to_file_with_data_type (struct.cc:1991)
    void output_item_to_file(const std::string &key, const std::string &path) const;
    void output_2_dir(const std::string &dir) const;

    // widened row count of a sub-matrix: end-begin+1, or begin+last(nz_row_indices) when
    // padding created rows past end_row_index (e.g. get_begin_rows_of_BMT_..._row_direction.cc:52-58)
    uint64_t real_row_num(int id) const;

  private:
    struct entry {
        POS_TYPE pos;
        std::string name;
        int id;
        universal_array arr;
    };
    std::map<std::string, entry> items_;
    std::vector<std::string> order_; // insertion order, for stable dumps
};

// create_init_metadata_set_from_file (metadata_set.cc:612-707) + get_matrix_index_and_val_from_file
// (struct.cc:49-261): 1-based row-major-sorted COO text; '%' comments skipped; first data line =
// "rows cols [nnz]".  reference_ones=true reproduces the reference (every value := 1, struct.cc:188-195);
// false keeps the file's values (what the pruned-weight configs need).
std::shared_ptr<meta_data_set> create_init_metadata_set_from_file(const std::string &file_name, const std::string &matrix_name,
                                                                 const std::string &float_precise = "float",
                                                                 bool reference_ones = true);
// same store built from in-memory COO (row-major sorted, 0-based); vals may be null (ones)
std::shared_ptr<meta_data_set> create_init_metadata_set_from_coo(uint64_t rows, uint64_t cols, uint64_t nnz, const int64_t *row,
                                                                const int64_t *col, const float *vals,
                                                                const std::string &float_precise = "float");

// nnz of each row of a row-sorted index array restricted to [begin_row, end_row] x [begin_nz, end_nz]
// (data_transform_common.cc:7-49)
std::vector<uint64_t> get_nnz_of_each_row_in_spec_range(const universal_array &row_index, uint64_t begin_row_bound,
                                                        uint64_t end_row_bound, uint64_t begin_nz_bound, uint64_t end_nz_bound);

} // namespace gs
